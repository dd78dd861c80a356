"""Committed fixtures (tests/golden/, made by scripts/make_golden.py): the oracle must reproduce
them on CPU; the CUDA path must match them on the GPU without the oracle being present."""
import os

import numpy as np
import pytest

import rustraytrace_b200 as R

F = R._ffi
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCENES = [("in_one_weekend", None), ("in_one_weekend_gpu", None), ("the_rest_of_your_life", None)] + \
         [("the_next_week", i) for i in range(1, 10)]


def name(book, no):
    return book if no is None else f"{book}_{no}"


def load(book, no):
    z = np.load(os.path.join(GOLD, f"probe_{name(book, no)}.npz"))
    return z["ids"].astype(np.int32), z["t"]


@pytest.mark.parametrize("book,no", SCENES)
def test_oracle_reproduces_golden_probe(book, no):
    import oracle_ffi as O
    s = R.BookScene(book, no, overrides=dict(image_width=96))
    ids, t = O.OracleScene(s.desc).probe(s.camera)
    gids, gt = load(book, no)
    assert np.array_equal(ids, gids)
    assert np.allclose(t, gt, rtol=1e-13, atol=0)   # libm ulp differences only


@pytest.mark.gpu
@pytest.mark.parametrize("book,no", SCENES)
def test_gpu_probe_matches_golden(book, no):
    s = R.BookScene(book, no, overrides=dict(image_width=96))
    ctx = R.Context(0)
    gs = ctx.scene(s.desc)
    ids, t = ctx.probe(gs, s.camera, 64)
    gids, gt = load(book, no)
    assert np.array_equal(ids, gids)
    hit = gids >= 0
    assert np.max(np.abs(t[hit] - gt[hit]) / gt[hit]) <= 1e-5
    gs.close(); ctx.close()
