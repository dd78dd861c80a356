import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import rustraytrace_b200 as R
for book, no in (("in_one_weekend", None), ("the_next_week", 1), ("the_rest_of_your_life", None)):
    for depth in (1, 2, 10):
        s = R.BookScene(book, no, overrides=dict(image_width=100, samples_per_pixel=49, max_depth=depth))
        c = R.Context(0); gs = c.scene(s.desc)
        a = c.render(gs, s.camera, s.integrator, seed=9).astype(np.float64); gs.close(); c.close()
        os.environ["RRT_POOL"] = "1"
        p = R.Context(0); del os.environ["RRT_POOL"]
        gp = p.scene(s.desc)
        b = p.render(gp, s.camera, s.integrator, seed=9).astype(np.float64); gp.close(); p.close()
        d = np.abs(a[..., :3] - b[..., :3]).max(axis=2)
        bad = d > 1e-3
        ys, xs = np.nonzero(bad)
        print(book, no, "depth", depth, "bad", int(bad.sum()), "of", bad.size, "max", float(d.max()), "mean a", a[..., :3].mean(), "mean b", b[..., :3].mean(),
              "first", list(zip(ys[:6].tolist(), xs[:6].tolist())), [ (a[y, x, :3].round(3).tolist(), b[y, x, :3].round(3).tolist()) for y, x in zip(ys[:3], xs[:3])])
