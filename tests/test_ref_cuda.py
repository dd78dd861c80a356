"""GPU tier: our kernel against the REFERENCE'S OWN CUDA kernel (src/cuda/mod.rs:15-335, built by
oracle/ref_cuda/build.sh into oracle/_ref/libref_cuda.so) on the reference's seeded GPU-path scene
(build_in_one_weekend_scene, src/gpu/mod.rs:124-301 — SmallRng::seed_from_u64(0x5EED_1234)).
This is the one place a piece of the real reference runs on the box: its f32 brute-force renderer
and ours must converge to the same image (north star: PSNR >= 40 dB, mean radiance within 0.5 %)."""
import numpy as np
import pytest

import ref_cuda_ffi as RC
import rustraytrace_b200 as R

F = R._ffi
pytestmark = pytest.mark.gpu


def psnr8(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10.0 * np.log10(255.0 ** 2 / mse)


@pytest.fixture(scope="module")
def ctx():
    if not RC.available():
        pytest.skip("oracle/_ref/libref_cuda.so not built (needs /root/reference at build time)")
    c = R.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("overrides", [dict(), dict(background=(0.7, 0.8, 1.0))], ids=["sky", "solid-background"])
def test_converged_render_matches_reference_cuda_kernel(ctx, overrides):
    ov = dict(image_width=96, samples_per_pixel=8192, max_depth=50)
    ov.update(overrides)
    s = R.BookScene("in_one_weekend_gpu", overrides=ov)
    ref, ms, passes = RC.render(s, seed=0xC0FFEE)
    ref2, _, _ = RC.render(s, seed=0xBEEF)          # the reference against itself: the noise floor
    assert passes == 32 and np.allclose(ref[..., 3], 8192)
    gs = ctx.scene(s.desc)
    ours = ctx.render(gs, s.camera, s.integrator, seed=9)
    img = lambda a: ctx.resolve(a, 8192)
    floor = psnr8(img(ref), img(ref2))
    got = psnr8(img(ours), img(ref))
    assert got >= min(40.0, floor - 1.0), f"PSNR {got:.2f} dB (reference-vs-reference floor {floor:.2f} dB)"
    m_ref = 0.5 * (ref[..., :3].mean(axis=(0, 1)) + ref2[..., :3].mean(axis=(0, 1))) / 8192
    m_ours = ours[..., :3].mean(axis=(0, 1)) / 8192
    noise = np.abs(ref[..., :3].mean(axis=(0, 1)) - ref2[..., :3].mean(axis=(0, 1))) / 8192 / m_ref
    rel = np.abs(m_ours - m_ref) / m_ref
    assert np.all(rel <= np.maximum(0.005, 3 * noise)), (rel, noise)
    gs.close()


def test_reference_pods_have_the_reference_layout():
    s = R.BookScene("in_one_weekend_gpu", overrides=dict(image_width=1200))
    cu, S, M, W, H = RC.pods_from_scene(s)
    assert (W, H) == (1200, 675) and cu.nbytes == 144 and S.itemsize == 32 and M.itemsize == 32
    assert S[0]["cr"].tolist() == [0.0, -1000.0, 0.0, 1000.0] and M[S[0]["mat"]]["kind"] == 0   # ground, src/gpu/mod.rs:205
    assert S[-3]["cr"].tolist() == [0.0, 1.0, 0.0, 1.0] and M[S[-3]["mat"]]["kind"] == 2         # glass
    assert S[-1]["cr"].tolist() == [4.0, 1.0, 0.0, 1.0] and M[S[-1]["mat"]]["kind"] == 1         # metal, fuzz 0
