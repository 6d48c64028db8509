"""Host-side pieces of the deconvolution entry points (no GPU): clean-beam conversions, the PSF fit
(image/deconvolution.py:1113-1159; analytic cases -- astropy is not installed, parity unpinned),
windows, PSF support, channel scatter / gather."""
import numpy as np
import pytest

from ska_sdp_func_python_b200 import deconvolution as dc
from ska_sdp_func_python_b200 import minimal_models as mm


def image(npix=64, nchan=1, npol=1, cell=1e-3):
    pf = mm.PolarisationFrame("stokesI" if npol == 1 else "stokesIQ")
    return mm.create_image(npix, cell, mm.SkyCoord.from_deg(10.0, -40.0), frequency=1e8 + 1e6 * np.arange(nchan),
                           polarisation_frame=pf)


def gaussian_psf(npix, sx, sy, theta):
    im = image(npix)
    y, x = np.mgrid[:npix, :npix]
    im["pixels"].data[0, 0] = dc._gaussian2d((1.0, npix // 2, npix // 2, sx, sy, theta), x.astype(float), y.astype(float))
    return im


@pytest.mark.parametrize("sx,sy,theta", [(2.0, 2.0, 0.0), (1.5, 3.0, 0.4), (3.5, 1.2, -0.7)])
def test_fit_psf_recovers_gaussian(sx, sy, theta):
    psf = gaussian_psf(64, sx, sy, theta)
    beam = dc.fit_psf(psf)
    cell_deg = np.degrees(1e-3)
    fwhm = np.sqrt(8.0 * np.log(2.0))
    assert beam["bmaj"] == pytest.approx(max(sx, sy) * fwhm * cell_deg, rel=1e-6)
    assert beam["bmin"] == pytest.approx(min(sx, sy) * fwhm * cell_deg, rel=1e-6)
    if sx != sy:
        # the fitted ellipse reproduces the PSF patch whatever branch of theta the fit took
        bp = dc.convert_clean_beam_to_pixels(psf, beam)
        y, x = np.mgrid[25:40, 25:40]
        model = dc._gaussian2d((1.0, 32, 32, bp[0], bp[1], bp[2]), x.astype(float), y.astype(float))
        assert np.abs(model - psf["pixels"].data[0, 0, 25:40, 25:40]).max() < 1e-6


def test_fit_psf_degenerate_falls_back_to_one_pixel():
    psf = image(64)
    psf["pixels"].data[0, 0, 32, 32] = 1.0            # a delta: stddev collapses
    beam = dc.fit_psf(psf)
    assert np.isfinite(beam["bmaj"]) and beam["bmaj"] > 0 and beam["bmin"] > 0


def test_beam_conversions_round_trip():
    im = image(64)
    beam = dc.convert_clean_beam_to_degrees(im, (1.5, 3.0, 0.3))
    assert beam["bmaj"] > beam["bmin"] and beam["bpa"] == pytest.approx(np.degrees(0.3))
    assert dc.convert_clean_beam_to_pixels(im, beam) == pytest.approx((1.5, 3.0, 0.3))
    beam = dc.convert_clean_beam_to_degrees(im, (3.0, 1.5, 0.3))
    assert beam["bpa"] == pytest.approx(np.degrees(0.3) + 90.0)
    assert dc.convert_clean_beam_to_pixels(im, beam)[:2] == pytest.approx((1.5, 3.0))


def test_beam_kernel_shape_and_peak():
    k = dc._beam_kernel((1.0, 2.0, 0.0))
    assert k.shape == (17, 17) and k.max() == 1.0 and k[8, 8] == 1.0     # 8 * 2 = 16 -> odd 17
    assert k[8, 9] == pytest.approx(np.exp(-0.5))                        # x (last axis) has stddev 1
    assert k[9, 8] == pytest.approx(np.exp(-0.125))
    assert dc._beam_kernel((1.1, 1.1, 0.0)).shape == (9, 9)              # ceil(8.8) = 9


def test_windows():
    d = [image(64, npol=2)]
    assert dc.find_window_list(d, "t") is None
    w = dc.find_window_list(d, "t", window_shape="quarter")[0]["pixels"].data
    assert w.shape == (1, 2, 64, 64) and w.sum() == 2 * 31 * 31 and w[0, 0, 17, 17] == 1 and w[0, 0, 16, 17] == 0
    w = dc.find_window_list(d, "t", window_shape="no_edge", window_edge=4)[0]["pixels"].data
    assert w[0, 0, 5, 5] == 1 and w[0, 0, 4, 5] == 0 and w[0, 0, 59, 59] == 1 and w[0, 0, 60, 59] == 0
    mask = image(64, npol=2)
    mask["pixels"].data[..., 10:20, 10:20] = 1.0
    w = dc.find_window_list(d, "t", window_shape="quarter", mask=mask)[0]["pixels"].data
    assert w.sum() == 200
    with pytest.raises(ValueError):
        dc.find_window_list(d, "t", window_shape="circle")


def test_psf_support_and_peak_check():
    d = [image(32)]
    psf = gaussian_psf(64, 2.0, 2.0, 0.0)
    dc.check_psf_peak([psf])
    cut = dc.bound_psf_list(d, "t", [psf])[0]["pixels"].data
    assert cut.shape == (1, 1, 32, 32) and cut[0, 0, 16, 16] == 1.0
    cut = dc.bound_psf_list(d, "t", [psf], psf_support=5)[0]["pixels"].data
    assert cut.shape == (1, 1, 10, 10) and cut[0, 0, 5, 5] == 1.0
    assert dc.bound_psf_list(d, "t", [psf], psf_support=40)[0] is psf      # too small: whole PSF
    psf["pixels"].data *= 0.99
    with pytest.raises(AssertionError):
        dc.check_psf_peak([psf])


def test_scatter_gather_channels():
    im = image(16, nchan=3, npol=2)
    im["pixels"].data[:] = np.random.default_rng(0).normal(size=(3, 2, 16, 16))
    parts = dc.image_scatter_channels(im)
    assert len(parts) == 3 and parts[1]["pixels"].data.shape == (1, 2, 16, 16)
    assert parts[2].image_acc.wcs.wcs.crval[3] == pytest.approx(1e8 + 2e6)
    back = dc.image_gather_channels(parts)
    assert np.array_equal(back["pixels"].data, im["pixels"].data)
    assert back.image_acc.wcs.wcs.crval[3] == pytest.approx(1e8)


def test_frequency_moments_and_back():
    """image/taylor_terms.py:160-288: moment k = sum_c w_c^k image_c with w_c = (nu_c - nu_ref) / nu_ref,
    nu_ref = the middle channel; rebuilding evaluates the polynomial in w_c."""
    from ska_sdp_func_python_b200 import taylor_terms as tt
    rng = np.random.default_rng(2)
    nchan = 5
    im = image(16, nchan=nchan, npol=2)
    im["pixels"].data[:] = rng.normal(size=(nchan, 2, 16, 16))
    parts = dc.image_scatter_channels(im)
    freq = 1e8 + 1e6 * np.arange(nchan)
    w = (freq - freq[2]) / freq[2]
    mom = tt.calculate_image_list_frequency_moments(parts, nmoment=3)
    assert mom["pixels"].data.shape == (3, 2, 16, 16) and mom.image_acc.wcs.wcs.ctype[3] == "MOMENT"
    for k in range(3):
        want = sum(w[c] ** k * im["pixels"].data[c] for c in range(nchan))
        assert np.allclose(mom["pixels"].data[k], want, rtol=0, atol=1e-14)
    back = tt.calculate_image_list_from_frequency_taylor_terms(parts, mom)
    assert len(back) == nchan
    for c in range(nchan):
        want = sum(w[c] ** k * mom["pixels"].data[k] for k in range(3))
        assert np.allclose(back[c]["pixels"].data[0], want, rtol=0, atol=1e-14)
    with pytest.raises(ValueError):
        tt.calculate_image_list_frequency_moments(parts, nmoment=6)
    with pytest.raises(ValueError):
        tt.calculate_image_list_frequency_moments(parts, nmoment=0)
    assert tt.calculate_image_list_frequency_moments([None], nmoment=1) is None
