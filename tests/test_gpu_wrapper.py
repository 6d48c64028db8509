"""invert_ng / predict_ng drop-in against golden vectors produced by THE REFERENCE'S OWN
ng.py / base.py / visibility/base.py (tests/golden/make_golden.py; ducc0 replaced by the exact
DFT there, so the goldens are exact and the gridder must be within its epsilon of them)."""
import numpy as np
import pytest

from conftest import load_ng_case, rel_l2, rel_max

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def imaging(built):
    import torch
    assert torch.cuda.is_available()
    from ska_sdp_func_python_b200 import imaging
    return imaging


def test_case_a_invert_variants(imaging):
    g, vis, im = load_ng_case("ng_case_a.npz")
    vis0 = vis["vis"].data.copy()
    d, sw = imaging.invert_ng(vis, im, dopsf=False, normalise=True, epsilon=1e-10)
    assert np.array_equal(sw, g["sumwt"])                       # bit-exact
    assert sw.dtype == np.float64 and sw.shape == (1, 1)
    assert d["pixels"].data.dtype == np.float64
    assert rel_max(d["pixels"].data, g["dirty_norm"]) < 1e-9
    assert np.array_equal(vis["vis"].data, vis0) and np.all(im["pixels"].data == 0)  # no mutation
    d, _ = imaging.invert_ng(vis, im, normalise=False, epsilon=1e-10)
    assert rel_max(d["pixels"].data, g["dirty_raw"]) < 1e-9
    d, sw = imaging.invert_ng(vis, im, dopsf=True, epsilon=1e-10)
    assert rel_max(d["pixels"].data, g["psf"]) < 1e-9
    assert abs(d["pixels"].data[0, 0, 32, 32] - 1.0) < 1e-9    # PSF peak after normalisation
    d, _ = imaging.invert_visibility(vis, im, context="2d", epsilon=1e-10)
    assert rel_max(d["pixels"].data, g["dirty_2d"]) < 1e-9


def test_case_a_default_epsilon_fp64_bound(imaging):
    """Defaults of the reference (epsilon=1e-12, fp64): within 1e-10 of the peak."""
    g, vis, im = load_ng_case("ng_case_a.npz")
    d, _ = imaging.invert_ng(vis, im)
    assert rel_max(d["pixels"].data, g["dirty_norm"]) < 1e-10
    model = im.copy(deep=True)
    model["pixels"].data[...] = g["model"]
    p = imaging.predict_ng(vis, model)
    assert rel_max(p["vis"].data, g["predicted"]) < 1e-10


def test_case_a_predict(imaging):
    g, vis, im = load_ng_case("ng_case_a.npz")
    model = im.copy(deep=True)
    model["pixels"].data[...] = g["model"]
    p = imaging.predict_ng(vis, model, epsilon=1e-10)
    assert p is not vis and p["vis"].data.dtype == np.complex128
    assert p["vis"].data.shape == vis["vis"].data.shape
    # predict ignores flags and weights (ng.py:103): flagged samples are predicted too
    assert rel_max(p["vis"].data, g["predicted"]) < 1e-9
    p2 = imaging.predict_visibility(vis, model, context="2d", epsilon=1e-10)
    assert rel_max(p2["vis"].data, g["predicted_2d"]) < 1e-9


def test_case_a_single_precision_bound(imaging):
    """fp32 opt-in: max-abs error <= 10*epsilon of the peak (north_star)."""
    g, vis, im = load_ng_case("ng_case_a.npz")
    eps = 1e-5
    d, sw = imaging.invert_ng(vis, im, epsilon=eps, precision="single")
    assert np.array_equal(sw, g["sumwt"])
    assert rel_max(d["pixels"].data, g["dirty_norm"]) < 10 * eps
    model = im.copy(deep=True)
    model["pixels"].data[...] = g["model"]
    p = imaging.predict_ng(vis, model, epsilon=eps, precision="single")
    assert rel_max(p["vis"].data, g["predicted"]) < 10 * eps


def test_case_b_cube_polarisation_phase_shift(imaging):
    """linear -> stokesIQUV, 4 visibility channels onto 2 image channels, image centre 1 degree
    from the visibility phase centre (shift_vis_to_image both directions)."""
    g, vis, im = load_ng_case("ng_case_b.npz")
    d, sw = imaging.invert_ng(vis, im, epsilon=1e-10)
    assert np.array_equal(sw, g["sumwt"])
    for c in range(2):
        for p in range(4):
            assert rel_max(d["pixels"].data[c, p], g["dirty_norm"][c, p]) < 1e-8, (c, p)
    d, sw = imaging.invert_ng(vis, im, dopsf=True, normalise=False, epsilon=1e-10)
    assert np.array_equal(sw, g["sumwt_psf"])
    assert rel_max(d["pixels"].data[:, 0], g["psf_raw"][:, 0]) < 1e-9
    assert np.all(d["pixels"].data[:, 1:] == 0)                 # ng.py:231-233, :267
    model = im.copy(deep=True)
    model["pixels"].data[...] = g["model"]
    p = imaging.predict_ng(vis, model, epsilon=1e-10)
    assert rel_max(p["vis"].data, g["predicted"]) < 1e-9
    pc = p.attrs["phasecentre"]                                  # base.py:90
    assert np.allclose([pc.ra.rad, pc.dec.rad], g["predicted_pc"], atol=1e-15)


def test_sumwt_on_device_option(imaging):
    g, vis, im = load_ng_case("ng_case_b.npz")
    _, sw = imaging.invert_ng(vis, im, epsilon=1e-8, sumwt_on_device=True)
    assert np.allclose(sw, g["sumwt"], rtol=1e-13)


def test_invert_of_zero_vis_keeps_sumwt(imaging):
    """Gridding is skipped when all visibilities vanish but sumwt is still accumulated
    (ng.py:238, :258)."""
    g, vis, im = load_ng_case("ng_case_a.npz")
    z = vis.copy(deep=True, zero=True)
    d, sw = imaging.invert_ng(z, im, epsilon=1e-8)
    assert np.all(d["pixels"].data == 0) and np.array_equal(sw, g["sumwt"])


def test_predict_then_invert_roundtrip_like_reference_test(imaging):
    """tests/imaging/test_imaging_ng.py:60-77 of the reference: predict of a zero model gives
    zero visibilities; inverting (predicted - original) stays bounded."""
    g, vis, im = load_ng_case("ng_case_a.npz")
    p = imaging.predict_ng(vis, im, epsilon=1e-8)
    assert np.all(p["vis"].data == 0)
    p["vis"].data = p["vis"].data - vis["vis"].data
    d, sw = imaging.invert_ng(p, im, epsilon=1e-8)
    assert np.max(np.abs(d["pixels"].data)) > 0.0 and np.isfinite(d["pixels"].data).all()
