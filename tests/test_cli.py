"""GPU tier: the `rrt` command line (host/main.cpp, mirror of src/main.rs:14-99) end to end."""
import os
import subprocess

import numpy as np
import pytest

import rustraytrace_b200 as R

pytestmark = pytest.mark.gpu
RRT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "rustraytrace_b200", "bin", "rrt")


def read_p3(text):
    tok = text.split()
    assert tok[0] == "P3" and tok[3] == "255"
    w, h = int(tok[1]), int(tok[2])
    return np.array(tok[4:], dtype=np.int64).reshape(h, w, 3)


def test_cli_renders_the_same_image_as_the_library(tmp_path):
    p = subprocess.run([RRT, "--cuda", "the_next_week", "3", "--width", "96", "--spp", "64", "--depth", "10", "--sample-seed", "5"],
                       capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    img = read_p3(p.stdout)
    assert img.shape == (54, 96, 3)
    s = R.BookScene("the_next_week", 3, overrides=dict(image_width=96, samples_per_pixel=64, max_depth=10))
    ctx = R.Context(0)
    gs = ctx.scene(s.desc)
    want = ctx.resolve(ctx.render(gs, s.camera, s.integrator, seed=5), 64)
    # sample chunks of a tile are summed with float atomics, so two runs can differ in the last bit
    assert np.abs(img - want.astype(np.int64)).max() <= 1 and (img != want).mean() < 0.01
    gs.close(); ctx.close()


def test_cli_book_aliases_probe_and_errors(tmp_path):
    out = str(tmp_path / "p.txt")
    p = subprocess.run([RRT, "rest", "--width", "32", "--probe", "-o", out], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    lines = open(out).read().splitlines()
    assert lines[0] == "# probe 32 32" and len(lines) == 1 + 32 * 32
    s = R.BookScene("the_rest_of_your_life", overrides=dict(image_width=32))
    ctx = R.Context(0)
    gs = ctx.scene(s.desc)
    ids, t = ctx.probe(gs, s.camera, 64)
    got = np.array([ln.split()[0] for ln in lines[1:]], dtype=np.int32).reshape(32, 32)
    assert np.array_equal(got, ids)
    gs.close(); ctx.close()
    bad = subprocess.run([RRT, "no_such_book"], capture_output=True, text=True, timeout=60)
    assert bad.returncode == 1 and "books: in_one_weekend, the_next_week, the_rest_of_your_life" in bad.stderr
    bad = subprocess.run([RRT, "--backend", "vulkan", "weekend"], capture_output=True, text=True, timeout=60)
    assert bad.returncode == 2


@pytest.mark.parametrize("devices,shard", [("0", "sample"), ("0", "tile"), ("0,1", "sample"), ("0,1", "tile")])
def test_cli_devices_flag_renders_on_a_group(devices, shard):
    """rrt --devices: the multi-GPU group of rrt_init_multi from the command line, no Python in the path."""
    import torch
    if len(devices.split(",")) > torch.cuda.device_count():
        pytest.skip("needs two GPUs")
    args = ["the_next_week", "2", "--width", "100", "--spp", "32", "--depth", "8", "--sample-seed", "3"]
    one = subprocess.run([RRT] + args, capture_output=True, text=True, timeout=120)
    many = subprocess.run([RRT] + args + ["--devices", devices, "--shard", shard], capture_output=True, text=True, timeout=180)
    assert one.returncode == 0 and many.returncode == 0, many.stderr
    assert f"{len(devices.split(','))} GPU(s)" in many.stderr
    a, b = read_p3(one.stdout), read_p3(many.stdout)
    assert a.shape == b.shape and np.abs(a - b).max() <= 1 and (a != b).mean() < 0.01
    bad = subprocess.run([RRT, "weekend", "--devices", "0,0"], capture_output=True, text=True, timeout=60)
    assert bad.returncode == 1 and "listed twice" in bad.stderr


def test_cli_binary_ppm_holds_the_same_pixels_as_the_text_ppm(tmp_path):
    """--p6: the resolve runs on the GPU (rrt_resolve_rgb8) and the file is a binary PPM; same bytes as the P3 numbers
    (write_color, color.rs:19-36), up to the last-bit freedom of the atomically summed sample chunks."""
    args = ["the_next_week", "2", "--width", "80", "--spp", "32", "--depth", "8", "--sample-seed", "3"]
    text = subprocess.run([RRT] + args, capture_output=True, text=True, timeout=120)
    assert text.returncode == 0, text.stderr
    out = str(tmp_path / "img.ppm")
    binary = subprocess.run([RRT] + args + ["--p6", "-o", out], capture_output=True, timeout=120)
    assert binary.returncode == 0, binary.stderr
    raw = open(out, "rb").read()
    head = raw.split(b"\n", 3)
    assert head[0] == b"P6" and head[2] == b"255"
    w, h = (int(x) for x in head[1].split())
    img = np.frombuffer(head[3], dtype=np.uint8).reshape(h, w, 3).astype(np.int64)
    want = read_p3(text.stdout)
    assert img.shape == want.shape == (45, 80, 3)
    assert np.abs(img - want).max() <= 1 and (img != want).mean() < 0.01
