"""Small renders of every feature class for compute-sanitizer (memcheck):
    compute-sanitizer --tool memcheck python scripts/sanitize_small.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rustraytrace_b200 as R

ctx = R.Context(0)
for book, no in [("in_one_weekend", None), ("the_next_week", 1), ("the_next_week", 3), ("the_next_week", 4),
                 ("the_next_week", 8), ("the_next_week", 9), ("the_rest_of_your_life", None)]:
    s = R.BookScene(book, no, overrides=dict(image_width=48, samples_per_pixel=4, max_depth=12))
    gs = ctx.scene(s.desc)
    acc = ctx.render(gs, s.camera, s.integrator, seed=1)
    for r in range(2):
        ctx.render(gs, s.camera, s.integrator, seed=1, shard_mode=R._ffi.SHARD_TILE, rank=r, world=2)
    ctx.probe(gs, s.camera, 64)
    ctx.probe(gs, s.camera, 32)
    ctx.resolve(acc, 4)
    gs.close()
    print(book, no, "ok", flush=True)
ctx.close()
