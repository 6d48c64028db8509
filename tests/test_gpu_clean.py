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


def test_hogbom_kernel_list(cl):
    """Image-level loop over channels and polarisations (image/deconvolution.py:543-643)."""
    from oracle import clean as oc
    from ska_sdp_func_python_b200 import minimal_models as mm
    g = np.load(os.path.join(GOLDEN, "clean_case.npz"))
    pc = mm.SkyCoord.from_deg(0.0, -30.0)
    pf = mm.PolarisationFrame("stokesIQ")
    dirty = mm.create_image(64, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    dirty["pixels"].data[0, 0] = g["a_dirty"]
    dirty["pixels"].data[0, 1] = 0.5 * g["a_dirty"][::-1]
    psf = mm.create_image(64, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    psf["pixels"].data[0, 0] = g["a_psf"]                      # pol 1 PSF all zero: skipped
    comps, resid = cl.hogbom_kernel_list([dirty], "test", [psf], None, gain=0.1, niter=300,
                                         fractional_threshold=0.01)
    assert len(comps) == len(resid) == 1 and isinstance(comps[0], mm.Image)
    assert np.array_equal(comps[0]["pixels"].data[0, 0], g["a_comps"])
    assert np.array_equal(resid[0]["pixels"].data[0, 0], g["a_res"])
    assert not comps[0]["pixels"].data[0, 1].any() and not resid[0]["pixels"].data[0, 1].any()
    psf["pixels"].data[0, 1] = g["a_psf"]
    win = mm.create_image(64, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    win["pixels"].data[0, :, 16:48, 16:48] = 1.0
    comps, resid = cl.hogbom_kernel_list([dirty], "test", [psf], [win], gain=0.2, niter=50)
    oc_c, oc_r, _ = oc.hogbom(dirty["pixels"].data[0, 1], g["a_psf"], win["pixels"].data[0, 1], 0.2,
                              0.0, 50, 0.01)
    assert np.array_equal(comps[0]["pixels"].data[0, 1], oc_c)
    assert np.array_equal(resid[0]["pixels"].data[0, 1], oc_r)
    with pytest.raises(ValueError):
        cl.hogbom_kernel_list([dirty], "test", [psf], None, gain=2.0)
