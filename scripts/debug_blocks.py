"""Where a GPU render and a converged oracle fixture disagree, per 4x4-pixel block (diagnostics):
    python scripts/debug_blocks.py the_next_week_1"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import rustraytrace_b200 as R
import test_converged_golden as T

key = sys.argv[1]
m = T.MANIFEST[key]
gold, var = T.load(key)
ctx = R.Context(0)
s = T.build(key, m["n"])
gs = ctx.scene(s.desc)
imgs = []
for seed in (77, 78):
    acc = ctx.render(gs, s.camera, s.integrator, seed=seed).astype(np.float64)
    imgs.append(acc[..., :3] / acc[..., 3:4])
n = m["n"]
def blocks(a, b, na, nb):
    se2 = var * (1 / na + 1 / nb) + (4e-6 * np.abs(gold)) ** 2
    H, W = a.shape[0] // 4 * 4, a.shape[1] // 4 * 4
    sh = (H // 4, 4, W // 4, 4, 3)
    d = (a - b)[:H, :W].reshape(sh).sum(axis=(1, 3)); s_ = se2[:H, :W].reshape(sh).sum(axis=(1, 3))
    return d / np.sqrt(s_)
for name, z in (("gpu77 vs gold", blocks(imgs[0], gold, n, n)), ("gpu78 vs gold", blocks(imgs[1], gold, n, n)), ("gpu77 vs gpu78", blocks(imgs[0], imgs[1], n, n))):
    print(name, "rms", float(np.sqrt((z * z).mean())), "mean", float(z.mean()))
    zz = np.abs(z).max(axis=2)
    idx = np.argsort(zz.ravel())[::-1][:8]
    for i in idx:
        y, x = divmod(int(i), zz.shape[1])
        print("   block", y, x, "z", np.round(z[y, x], 2), "gold", np.round(gold[4 * y:4 * y + 4, 4 * x:4 * x + 4].mean(axis=(0, 1)), 4),
              "gpu", np.round(imgs[0][4 * y:4 * y + 4, 4 * x:4 * x + 4].mean(axis=(0, 1)), 4))
np.set_printoptions(linewidth=250, precision=1, suppress=True)
print(np.abs(blocks(imgs[0], gold, n, n)).max(axis=2))
