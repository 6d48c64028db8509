"""Hogbom CLEAN on the GPU (csrc/clean_kernels.cu through the C ABI): bit-exact against golden
vectors from the reference's own hogbom() and against the numpy oracle on a larger seeded image
(ties, windows, PSF smaller / larger than the image, early stop, niter exhaustion)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cl(built):
    import torch
    assert torch.cuda.is_available()
    from ska_sdp_func_python_b200 import cleaners
    return cleaners


@pytest.mark.parametrize("name", ["a", "b", "c", "d"])
def test_golden_bit_exact(cl, name):
    g = np.load(os.path.join(GOLDEN, "clean_case.npz"))
    gain, thresh, niter, frac = g[f"{name}_par"]
    window = g[f"{name}_window"] if f"{name}_window" in g.files else None
    dirty = g[f"{name}_dirty"].copy()
    comps, res = cl.hogbom(dirty, g[f"{name}_psf"], window, gain, thresh, int(niter), frac)
    assert isinstance(comps, np.ndarray) and comps.dtype == np.float64
    assert np.array_equal(comps, g[f"{name}_comps"])
    assert np.array_equal(res, g[f"{name}_res"])
    assert np.array_equal(dirty, g[f"{name}_dirty"])          # input not modified


def test_larger_image_against_oracle(cl):
    import torch
    from oracle import clean as oc
    rng = np.random.default_rng(8)
    n = 512
    yy, xx = np.meshgrid(np.arange(n) - n // 2, np.arange(n) - n // 2)
    r = np.hypot(xx, yy)
    psf = np.exp(-(r / 2.0) ** 2) + 0.05 * np.cos(r / 3.1) * np.exp(-r / 40.0) * (r > 0)
    dirty = 0.02 * rng.normal(size=(n, n))
    for _ in range(40):
        x, y = rng.integers(20, n - 20, 2)
        dirty[x - 10:x + 10, y - 10:y + 10] += rng.uniform(0.5, 4.0) * psf[n // 2 - 10:n // 2 + 10, n // 2 - 10:n // 2 + 10]
    dirty[100, 100] = dirty[300, 50] = 9.0                      # exact tie: the first one wins
    comps, res, it = oc.hogbom(dirty, psf, None, 0.2, 0.0, 400, 0.0)
    gc, gr = cl.hogbom(torch.from_numpy(dirty).cuda(), torch.from_numpy(psf).cuda(), None, 0.2, 0.0, 400, 0.0)
    assert gc.is_cuda and np.array_equal(gc.cpu().numpy(), comps) and np.array_equal(gr.cpu().numpy(), res)
    assert comps[100, 100] != 0 and it == 400
    # early stop through the fractional threshold, boolean window
    window = np.zeros((n, n), dtype=bool)
    window[64:448, 64:448] = True
    dirty[100, 100] = dirty[300, 50] = 0.0                      # back to PSF-shaped sources only
    comps, res, it = oc.hogbom(dirty, psf, window, 0.3, 0.0, 5000, 0.3)
    gc, gr = cl.hogbom(dirty, psf, window, 0.3, 0.0, 5000, 0.3)
    assert it < 5000
    assert np.array_equal(gc, comps) and np.array_equal(gr, res)
    dirty[100, 100] = dirty[300, 50] = 9.0
    v, (x, y) = cl.image_peak(dirty)
    assert v == 9.0 and (x, y) == (100, 100)


def test_argument_checks(cl):
    d = np.zeros((8, 8))
    p = np.zeros((8, 8))
    p[4, 4] = 1.0
    with pytest.raises(AssertionError):
        cl.hogbom(d, p, None, 2.5, 0.0, 10, 0.0)
    with pytest.raises(AssertionError):
        cl.hogbom(d, p, None, 0.1, 0.0, 0, 0.0)
    with pytest.raises(AssertionError):
        cl.hogbom(d, 0.5 * p, None, 0.1, 0.0, 10, 0.0)          # PSF peak not 1
    comps, res = cl.hogbom(d, p, None, 0.1, 0.0, 10, 0.0)       # all-zero image: niter no-op iterations
    assert not comps.any() and not res.any()
