"""Probe mismatches against the oracle, with the primitives involved:
    python scripts/probe_debug.py the_next_week 9 200      (scene number "-" for books 1 and 3)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oracle_ffi as O  # noqa: E402
import rustraytrace_b200 as R  # noqa: E402

book, sc, w = sys.argv[1], (None if sys.argv[2] == "-" else int(sys.argv[2])), int(sys.argv[3])
s = R.BookScene(book, sc, overrides=dict(image_width=w))
ctx = R.Context(0)
gs = ctx.scene(s.desc)
ids, t = ctx.probe(gs, s.camera, 64)
o = O.OracleScene(s.desc)
oids, ot = o.probe(s.camera)
bad = np.argwhere(ids != oids)
print("mismatches", len(bad))
for y, x in bad[:10]:
    print((y, x), "gpu", ids[y, x], repr(t[y, x]), "oracle", oids[y, x], repr(ot[y, x]))
    for i in (ids[y, x], oids[y, x]):
        if i >= 0:
            h = s.desc.hittables[i]
            print("   id", i, "kind", h.kind, list(h.a), list(h.b), list(h.c))
