"""The Hogbom oracle (oracle/clean.py) against outputs of the reference's own hogbom()
(tests/golden/make_golden_clean.py).  CPU only, bit-exact."""
import os

import numpy as np
import pytest

from oracle import clean as oc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "clean_case.npz")


@pytest.mark.parametrize("name", ["a", "b", "c", "d"])
def test_hogbom_bit_exact(name):
    g = np.load(GOLD)
    gain, thresh, niter, frac = g[f"{name}_par"]
    window = g[f"{name}_window"] if f"{name}_window" in g.files else None
    comps, res, it = oc.hogbom(g[f"{name}_dirty"], g[f"{name}_psf"], window, gain, thresh, int(niter), frac)
    assert np.array_equal(comps, g[f"{name}_comps"])
    assert np.array_equal(res, g[f"{name}_res"])
    assert 1 <= it <= int(niter)
