"""The BASELINE.json configs as RenderOverrides values (sizes come only from overrides in the
reference: src/config.rs:50-62 applied by each book's apply_overrides).

    key -> (book, the_next_week scene number or None, overrides, default multi-GPU shard mode)

C4's depth is scene 9's shipped 40 (the_next_week/mod.rs:78); BASELINE.json names none.
F_PATH[key] = algorithmic FP32-equivalent flops per camera path, counted by the
oracle's op counters on the reference-shaped scene graph with scripts/count_flops.py and frozen
here so the roofline figure is a property of the reference algorithm, not of our kernel.
"""
SCENE_SEED = 0x5EED1234  # the constant the reference's GPU path seeds its scene with, src/gpu/mod.rs:199

CONFIGS = {
    "C1": ("in_one_weekend", None, dict(image_width=400, samples_per_pixel=100, max_depth=50), "sample"),
    "C2": ("in_one_weekend", None, dict(image_width=1200, samples_per_pixel=500, max_depth=50), "sample"),
    "C3a": ("the_next_week", 1, dict(image_width=1920, samples_per_pixel=1000, max_depth=50), "sample"),
    "C3b": ("the_next_week", 3, dict(image_width=1920, samples_per_pixel=1000, max_depth=50), "sample"),
    "C3c": ("the_next_week", 4, dict(image_width=1920, samples_per_pixel=1000, max_depth=50), "sample"),
    "C4": ("the_next_week", 9, dict(image_width=800, samples_per_pixel=10000, max_depth=40), "sample"),
    "C5": ("the_rest_of_your_life", None, dict(image_width=3840, samples_per_pixel=4096, max_depth=50), "tile"),
}

DESCRIPTION = {
    "C1": "book-1 final random spheres 400x225, 100 spp, depth 50",
    "C2": "book-1 final random spheres 1200x675, 500 spp, depth 50, defocus 0.6",
    "C3a": "book-2 bouncing spheres (motion blur, checker) 1920x1080, 1000 spp, depth 50",
    "C3b": "book-2 earth (image texture) 1920x1080, 1000 spp, depth 50",
    "C3c": "book-2 perlin spheres 1920x1080, 1000 spp, depth 50",
    "C4": "book-2 final scene 800x800, 10000 spp, depth 40",
    "C5": "book-3 cornell box, mixture-pdf lights, 3840x3840, 4096 spp, depth 50",
}

# flops per path, filled from scripts/count_flops.py output (tests/golden/flops_per_path.json)
F_PATH = {}


def _load_f_path():
    import json
    import os
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "flops_per_path.json")
    if os.path.exists(p):
        with open(p) as f:
            F_PATH.update({k: v["flops_per_path"] for k, v in json.load(f).items()})


_load_f_path()


def build(key, **override_changes):
    """BookScene for a named config (overrides can be changed, e.g. samples_per_pixel=16)."""
    from .api import BookScene
    book, scene_no, ov, _ = CONFIGS[key]
    ov = dict(ov)
    ov.update(override_changes)
    return BookScene(book, scene_no, seed=SCENE_SEED, overrides=ov)
