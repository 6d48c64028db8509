"""Image-level deconvolution entry points on the GPU: deconvolve_cube (hogbom / msclean with
windows and PSF support) against the kernels' golden cases, restore_cube against a direct
convolution (oracle/clean.py; parity unpinned for the restoring beam, see its header)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dc(built):
    import torch
    assert torch.cuda.is_available()
    from ska_sdp_func_python_b200 import deconvolution
    return deconvolution


def image(npix, nchan=1, npol=1, cell=1e-3):
    from ska_sdp_func_python_b200 import minimal_models as mm
    pf = mm.PolarisationFrame("stokesI" if npol == 1 else "stokesIQ")
    return mm.create_image(npix, cell, mm.SkyCoord.from_deg(10.0, -40.0),
                           frequency=1e8 + 1e6 * np.arange(nchan), polarisation_frame=pf)


def test_deconvolve_cube_hogbom_matches_kernel_golden(dc):
    g = np.load(os.path.join(GOLDEN, "clean_case.npz"))
    dirty, psf = image(64, nchan=2), image(64, nchan=2)
    dirty["pixels"].data[0, 0] = g["a_dirty"]
    dirty["pixels"].data[1, 0] = -g["a_dirty"][::-1, ::-1]
    psf["pixels"].data[:, 0] = g["a_psf"]
    comp, resid = dc.deconvolve_cube(dirty, psf, algorithm="hogbom", gain=0.1, niter=300,
                                     fractional_threshold=0.01, threshold=0.0)
    assert comp["pixels"].data.shape == (2, 1, 64, 64)
    assert np.array_equal(comp["pixels"].data[0, 0], g["a_comps"])
    assert np.array_equal(resid["pixels"].data[0, 0], g["a_res"])
    assert comp["pixels"].data[1, 0].any()


def test_deconvolve_cube_windows_and_support(dc):
    from oracle import clean as oc
    g = np.load(os.path.join(GOLDEN, "clean_case.npz"))
    dirty, psf = image(64), image(64)
    dirty["pixels"].data[0, 0] = g["a_dirty"]
    psf["pixels"].data[0, 0] = g["a_psf"]
    comp, resid = dc.deconvolve_cube(dirty, psf, algorithm="hogbom", gain=0.2, niter=80,
                                     window_shape="quarter", psf_support=20)
    window = np.zeros((64, 64))
    window[17:48, 17:48] = 1.0
    o_c, o_r, _ = oc.hogbom(g["a_dirty"], g["a_psf"][12:52, 12:52], window, 0.2, 0.0, 80, 0.01)
    assert np.array_equal(comp["pixels"].data[0, 0], o_c) and np.array_equal(resid["pixels"].data[0, 0], o_r)
    assert not comp["pixels"].data[0, 0][window == 0].any()
    with pytest.raises(ValueError):
        dc.deconvolve_cube(dirty, psf, algorithm="mmclean")          # one channel < 2 (nmoment - 1) + 1
    with pytest.raises(ValueError):
        dc.deconvolve_cube(dirty, psf, algorithm="hogbom-complex")   # needs stokesIQUV
    with pytest.raises(ValueError):
        dc.deconvolve_cube(dirty, psf, algorithm="nonsense")
    psf["pixels"].data *= 0.5
    with pytest.raises(AssertionError):
        dc.deconvolve_cube(dirty, psf, algorithm="hogbom")


def test_deconvolve_cube_msclean_default_algorithm(dc):
    g = np.load(os.path.join(GOLDEN, "msclean_case.npz"))
    dirty, psf = image(64), image(64)
    dirty["pixels"].data[0, 0] = g["a_dirty"]
    psf["pixels"].data[0, 0] = g["a_psf"]
    comp, resid = dc.deconvolve_cube(dirty, psf, gain=0.7, niter=60, fractional_threshold=0.01,
                                     scales=[0, 3, 10])
    assert np.abs(comp["pixels"].data[0, 0] - g["a_comps"]).max() < 1e-11
    assert np.abs(resid["pixels"].data[0, 0] - g["a_res"]).max() < 1e-11


@pytest.mark.parametrize("beam_pix", [(1.5, 1.5, 0.0), (1.2, 2.8, 0.5)])
def test_restore_cube(dc, beam_pix):
    from oracle import clean as oc
    rng = np.random.default_rng(5)
    model, resid = image(48, nchan=2, npol=2), image(48, nchan=2, npol=2)
    model["pixels"].data[:, :, rng.integers(0, 48, 9), rng.integers(0, 48, 9)] = rng.uniform(0.5, 2.0, 9)
    model["pixels"].data[0, 0, 0, 47] = 3.0                     # wraps round the edge
    resid["pixels"].data[:] = 0.01 * rng.normal(size=(2, 2, 48, 48))
    beam = dc.convert_clean_beam_to_degrees(model, beam_pix)
    out = dc.restore_cube(model, None, resid, clean_beam=beam)
    bp = dc.convert_clean_beam_to_pixels(model, beam)
    for chan in range(2):
        for pol in range(2):
            want = oc.restore_plane(model["pixels"].data[chan, pol], resid["pixels"].data[chan, pol], *bp)
            assert np.abs(out["pixels"].data[chan, pol] - want).max() < 1e-13
    assert out.attrs["clean_beam"] == beam and out["pixels"].data.dtype == np.float64
    # without a residual; beam fitted from a PSF
    psf = image(48, nchan=2, npol=2)
    y, x = np.mgrid[:48, :48]
    psf["pixels"].data[:, :] = dc._gaussian2d((1.0, 24, 24, 2.0, 2.0, 0.0), x.astype(float), y.astype(float))
    out2 = dc.restore_cube(model, psf, None)
    want = oc.restore_plane(model["pixels"].data[1, 1], None, 2.0, 2.0, 0.0)
    assert np.abs(out2["pixels"].data[1, 1] - want).max() < 1e-6
    with pytest.raises(ValueError):
        dc.restore_list([model], None, None, None)


@pytest.mark.parametrize("nmoment,use_window", [(2, False), (3, True), (1, False)])
def test_deconvolve_cube_mmclean(dc, nmoment, use_window):
    """The MSMFS wrapper (image/deconvolution.py:646-824): moments of the channel images, the device
    kernel, channel images rebuilt from the moment model -- against the same steps written out with
    the numpy oracle."""
    from oracle import clean as oc
    rng = np.random.default_rng(31)
    nchan, n = 7, 48
    g = np.load(os.path.join(GOLDEN, "clean_case.npz"))
    psf0 = g["b_psf"]                                             # 33 x 41 -> embed in 48 x 48
    base = np.zeros((n, n))
    base[8:41, 4:45] = psf0
    base = np.roll(base, (24 - (8 + 16), 24 - (4 + 20)), axis=(0, 1))
    assert base[24, 24] == 1.0
    dirty, psf = image(n, nchan=nchan), image(n, nchan=nchan)
    freq = 1e8 + 1e6 * np.arange(nchan)
    wc = (freq - freq[nchan // 2]) / freq[nchan // 2]
    src = rng.integers(8, 40, (5, 2))
    for c in range(nchan):
        psf["pixels"].data[c, 0] = base
        sky = np.zeros((n, n))
        for k, (x, y) in enumerate(src):
            sky[x, y] = (1.0 + k) * (1.0 + 20.0 * (k - 2) * wc[c])
        shifted = np.roll(base, (-24, -24), axis=(0, 1))
        dirty["pixels"].data[c, 0] = np.real(np.fft.ifft2(np.fft.fft2(sky) * np.fft.fft2(shifted))) \
            + 0.005 * rng.normal(size=(n, n))
    kw = dict(algorithm="mmclean", nmoment=nmoment, gain=0.5, niter=30, scales=[0, 3],
              fractional_threshold=0.01, threshold=0.0)
    if use_window:
        kw["window_shape"] = "no_edge"
    comp, resid = dc.deconvolve_cube(dirty, psf, **kw)
    assert comp["pixels"].data.shape == (nchan, 1, n, n)
    # the same, step by step, with the oracle
    dm = np.array([sum(wc[c] ** t * dirty["pixels"].data[c, 0] for c in range(nchan)) for t in range(nmoment)])
    npm = 2 * nmoment if nmoment > 1 else 1
    pm = np.array([sum(wc[c] ** t * psf["pixels"].data[c, 0] for c in range(nchan)) for t in range(npm)])
    peak = pm.max()
    window = None
    if use_window:
        window = np.zeros((n, n))
        window[17:32, 17:32] = 1.0
        window = window * nchan / nchan
    o_model, o_res, done = oc.msmfsclean(dm / peak, pm / peak, window, None, 0.5, 0.0, 30, [0, 3], 0.01)
    assert done > 5
    for c in range(nchan):
        want_c = sum(wc[c] ** t * o_model[t] for t in range(nmoment))
        want_r = sum(wc[c] ** t * o_res[t] for t in range(nmoment))
        assert np.abs(comp["pixels"].data[c, 0] - want_c).max() < 1e-9
        assert np.abs(resid["pixels"].data[c, 0] - want_r).max() < 1e-9
