"""CPU tier: the C-ABI libraries load, export every symbol include/*.h declares, and their
host-only entry points (flattener check, shard plan, Philox KAT) behave.  No compute calls."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import rustraytrace_b200 as R

F = R._ffi
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

ALL_SCENES = [("in_one_weekend", None), ("in_one_weekend_gpu", None), ("the_rest_of_your_life", None)] + \
             [("the_next_week", i) for i in range(1, 10)] + [("the_next_week", None)]


def declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rrth?_[a-z0-9_]+)\s*\(", src)))


@pytest.mark.parametrize("header,loader,table", [("rrt.h", F.rrt, F.RRT_SYMBOLS), ("rrt_host.h", F.rrth, F.RRTH_SYMBOLS)])
def test_library_exports_every_declared_symbol(header, loader, table):
    lib = loader()
    names = declared(header)
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), f"{n} declared in {header} but not exported"
    assert sorted(t[0] for t in table) == names, "ctypes table and header disagree"


def test_struct_sizes_match_header_layout():
    assert C.sizeof(F.Hittable) == 16 + 72
    assert C.sizeof(F.Material) == 8 + 24 + 16
    assert C.sizeof(F.Texture) == 24 + 24 + 8
    assert C.sizeof(F.Camera) == 8 + 16 + 24 + 8 + 72 + 16
    assert C.sizeof(F.RenderParams) == 24
    assert C.sizeof(F.Perlin) == 256 * 24 + 3 * 256 * 4


def test_no_gpu_is_a_loud_error_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(R.RrtError) as e:
        R.Context(0)
    assert e.value.status == F.ERR_NO_DEVICE and "no CPU fallback" in str(e.value)


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "rustraytrace_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle_ffi" not in text and "liboracle" not in text and "oracle/" not in text, f


def test_philox_known_answers():
    """Philox4x32-10 KATs from Random123's kat_vectors."""
    def ph(c, k):
        out = (C.c_uint32 * 4)()
        F.rrt().rrt_philox_kat(*c, *k, out)
        return [hex(x) for x in out]
    assert ph((0, 0, 0, 0), (0, 0)) == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert ph((0xffffffff,) * 4, (0xffffffff,) * 2) == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    assert ph((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0)) == \
        ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


@pytest.mark.parametrize("book,scene_no", ALL_SCENES)
def test_flattener_self_check_on_every_book_scene(book, scene_no):
    s = R.BookScene(book, scene_no, overrides=dict(image_width=64))
    c = R.flatten_check(s.desc)
    assert c["n_prims"] >= 1
    if book == "the_next_week" and scene_no in (9, None):
        assert c["n_prims"] == 2400 + 1000 + 6 + 1 and c["n_media"] == 2
    if book == "the_next_week" and scene_no == 8:
        assert c["n_media"] == 2 and c["n_prims"] == 6
    if book == "the_rest_of_your_life":
        assert c["n_prims"] == 13 and c["n_lights"] == 2


def test_scene_feature_bits_select_the_kernel_instance():
    """rrt_scene_features: which render_kernel<INTEGRATOR, FEAT> instance a scene needs (csrc/kernels.h)."""
    def feat(book, no=None):
        s = R.BookScene(book, no, overrides=dict(image_width=64))
        return R.scene_features(s.desc)
    assert feat("in_one_weekend") == F.SF_GRID              # a field of spheres: walked through a grid, no tree
    assert feat("the_next_week", 1) == F.SF_GRID | F.SF_GRID_MOVING | F.SF_MOTION | F.SF_TEXTURES   # bouncing spheres on a checker ground
    assert feat("the_next_week", 3) == F.SF_TEXTURES        # earth (image texture)
    assert feat("the_next_week", 4) == F.SF_TEXTURES | F.SF_NOISE   # perlin
    assert feat("the_next_week", 6) == F.SF_QUADS | F.SF_TEXTURES | F.SF_NOISE   # simple light over perlin spheres
    assert feat("the_next_week", 5) == F.SF_QUADS
    assert feat("the_next_week", 7) == F.SF_QUADS           # cornell box
    assert feat("the_next_week", 8) == F.SF_QUADS | F.SF_MEDIA   # cornell smoke
    assert feat("the_next_week", 9) == F.SF_ALL | F.SF_GRID   # final scene: every feature, two grids
    assert feat("the_rest_of_your_life") == F.SF_QUADS      # cornell box + glass sphere


def test_axis_aligned_boxes_become_box_leaves(monkeypatch):
    """The 400 ground boxes of the book-2 final scene (quad.rs box_(), 6 quads each) are one BVH leaf
    each; the rotated boxes of the Cornell scenes are not axis-aligned and stay plain quads."""
    s = R.BookScene("the_next_week", 9, overrides=dict(image_width=64))
    with_boxes = R.flatten_check(s.desc)
    assert R.scene_features(s.desc) & F.SF_BOXES
    monkeypatch.setenv("RRT_NO_BOX_LEAVES", "1")
    without = R.flatten_check(s.desc)
    assert not (R.scene_features(s.desc) & F.SF_BOXES)
    monkeypatch.delenv("RRT_NO_BOX_LEAVES")
    assert with_boxes["n_prims"] == without["n_prims"] == 3407
    # 400 boxes: 2400 single-quad leaves and their 2000 inner nodes collapse into 400 leaves
    assert without["n_nodes"] - with_boxes["n_nodes"] >= 1700
    for no in (7, 8):
        c = R.BookScene("the_next_week", no, overrides=dict(image_width=64))
        assert not (R.scene_features(c.desc) & F.SF_BOXES)


def test_outsized_primitives_leave_the_tree(monkeypatch):
    """Front run (flatten.cpp): the r = 1000 ground sphere and the three r = 1 spheres of the book-1
    scene (in_one_weekend/mod.rs:59-63, :93-105) are tested ahead of the traversal instead of sitting
    in the BVH; the checker still finds every primitive exactly once.  The book-2 final scene has no
    such outliers (its largest item is 4x the median); its front run is what its two grids leave over
    (seven primitives, tests/test_grid.py), and without grids everything stays in the tree."""
    s = R.BookScene("in_one_weekend", overrides=dict(image_width=64))
    front = R.flatten_check(s.desc)
    monkeypatch.setenv("RRT_NO_FRONT", "1")
    plain = R.flatten_check(s.desc)
    monkeypatch.delenv("RRT_NO_FRONT")
    assert front["n_prims"] == plain["n_prims"]
    assert plain["n_nodes"] == plain["n_prims"] - 1          # single-primitive leaves: n - 1 inner nodes
    assert front["n_nodes"] == plain["n_nodes"] - 4          # four primitives left the tree
    s9 = R.BookScene("the_next_week", 9, overrides=dict(image_width=64))
    a = R.flatten_check(s9.desc)
    monkeypatch.setenv("RRT_NO_FRONT", "1")
    b = R.flatten_check(s9.desc)
    assert a["n_nodes"] == b["n_nodes"] - 7 and a["n_prims"] == b["n_prims"]
    monkeypatch.delenv("RRT_NO_FRONT")
    monkeypatch.setenv("RRT_NO_GRID", "1")
    assert R.flatten_check(s9.desc) == b


def test_flattener_rejects_malformed_descriptions():
    b = R.DescBuilder()
    m = b.lambertian((0.5, 0.5, 0.5))
    sp = b.sphere((0, 0, -1), 0.5, m)
    with pytest.raises(R.RrtError) as e:
        R.flatten_check(b.desc(b.list([sp, 99])))
    assert e.value.status == F.ERR_INVALID
    b2 = R.DescBuilder()
    bad = b2.sphere((0, 0, 0), 1.0, 7)  # material index out of range
    with pytest.raises(R.RrtError):
        R.flatten_check(b2.desc(b2.list([bad])))
    b3 = R.DescBuilder()  # empty world is legal (every ray misses)
    assert R.flatten_check(b3.desc(b3.list([])))["n_prims"] == 0


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_shard_plan_partitions_the_frame(world):
    cam = R.make_camera(image_width=200, aspect_ratio=16 / 9, samples_per_pixel=37)
    H = F.rrt().rrt_image_height(C.byref(cam))
    cover = np.zeros((H, 200), dtype=np.int32)
    got = []
    for r in range(world):
        t = R.shard_plan(cam, F.INT_BOOK2, F.SHARD_TILE, r, world, mask=True)
        assert t["mask"].sum() == t["n_pixels"] and (t["sample_begin"], t["sample_end"]) == (0, 37)
        cover += t["mask"]
        s = R.shard_plan(cam, F.INT_BOOK2, F.SHARD_SAMPLE, r, world)
        assert s["n_pixels"] == 200 * H
        got.append((s["sample_begin"], s["sample_end"]))
    assert np.all(cover == 1), "every pixel belongs to exactly one rank"
    assert got[0][0] == 0 and got[-1][1] == 37 and all(got[i][1] == got[i + 1][0] for i in range(world - 1))
    # book 3 shards its floor(sqrt(spp))^2 strata (camera.rs:114-116)
    s3 = [R.shard_plan(cam, F.INT_BOOK3, F.SHARD_SAMPLE, r, world) for r in range(world)]
    assert s3[-1]["sample_end"] == 36


def test_image_height_truncates_like_camera_initialize():
    # SURVEY.md §4 KAT: 400->225, 1200->675, 1920->1080 at 16/9; 800->800, 3840->3840 at 1.0
    for w, ar, h in [(400, 16 / 9, 225), (1200, 16 / 9, 675), (1920, 16 / 9, 1080), (800, 1.0, 800), (3840, 1.0, 3840), (1, 16 / 9, 1)]:
        cam = R.make_camera(image_width=w, aspect_ratio=ar)
        assert F.rrt().rrt_image_height(C.byref(cam)) == h


C_TO_RUST = {"int32_t": "i32", "uint64_t": "u64", "uint32_t": "u32", "double": "f64", "float": "f32", "int": "i32"}


def c_struct_fields(header_text, name):
    """[(field, rust type)] of `typedef struct name { ... } name;` in a C header (arrays and pointers included)."""
    body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), header_text, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    out = []
    for decl in body.split(";"):
        decl = " ".join(decl.split())
        if not decl:
            continue
        m = re.match(r"(const )?(\w+)(\*?) (.+)$", decl)
        const, ctype, star, names = m.groups()
        for nm in names.split(","):
            nm = nm.strip()
            ptr = star == "*" or nm.startswith("*")
            nm = nm.lstrip("*")
            dims = [int(d) for d in re.findall(r"\[(\d+)\]", nm)]
            nm = nm.split("[")[0]
            base = C_TO_RUST.get(ctype) or "Rrt" + "".join(w.capitalize() for w in ctype.split("_")[1:])
            if ctype == "uint8_t":
                base = "u8"
            ty = base
            for d in reversed(dims):
                ty = f"[{ty}; {d}]"
            if ptr:
                ty = f"*const {ty}"
            out.append((nm, ty))
    return out


def test_rust_binding_declares_every_entry_point():
    """bindings/rrt_ffi.rs (unbuilt here: no rustc) must name every function of include/rrt.h, and its
    #[repr(C)] structs must list the header's fields in the header's order with the matching Rust types."""
    rs = open(os.path.join(ROOT, "bindings", "rrt_ffi.rs")).read()
    hdr = open(os.path.join(ROOT, "include", "rrt.h")).read()
    for n in declared("rrt.h"):
        assert re.search(r"pub fn %s\(" % n, rs), f"{n} missing from bindings/rrt_ffi.rs"
    for cname, rust in [("rrt_hittable", "RrtHittable"), ("rrt_material", "RrtMaterial"), ("rrt_texture", "RrtTexture"),
                        ("rrt_image", "RrtImage"), ("rrt_perlin", "RrtPerlin"), ("rrt_scene_desc", "RrtSceneDesc"),
                        ("rrt_camera", "RrtCamera"), ("rrt_render_params", "RrtRenderParams"), ("rrt_stats", "RrtStats")]:
        body = re.search(r"pub struct %s \{(.*?)\}\n" % rust, rs, re.S).group(1)
        got = [(a, " ".join(b.split())) for a, b in re.findall(r"pub (\w+): ([^,}]+?)\s*(?:,|$)", body.replace("\n", " "))]
        want = c_struct_fields(hdr, cname)
        assert got == want, (rust, got, want)
    # the flattener and the call site are complete files: every helper they call exists
    fl = open(os.path.join(ROOT, "bindings", "flatten.rs")).read()
    for fn in ("hittable", "container", "material", "texture", "image", "perlin", "finish"):
        assert re.search(r"fn %s\(" % fn, fl), fn
    assert "pub struct Keep" in fl
    cs = open(os.path.join(ROOT, "bindings", "call_site.rs")).read()
    assert "pub fn render_with_rrt(" in cs and "-> Result<(), String>" in cs and "rrt_init_multi" in cs
