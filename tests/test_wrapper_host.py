"""Host-side logic of the invert_ng / predict_ng drop-in that needs no GPU: geometry helpers
against the stand-in SkyCoord / WCS (which the reference's own shift_vis_to_image ran on when the
goldens were made), polarisation matrices, error behaviour (ng.py:52-56, :78, :172-173)."""
import numpy as np
import pytest

from conftest import load_ng_case

from ska_sdp_func_python_b200 import imaging, minimal_models as mm, polarisation


def test_image_centre_and_lmn_match_standins():
    pc = mm.SkyCoord.from_deg(180.0, -35.0)
    ipc = mm.SkyCoord.from_deg(181.0, -34.5)
    im = mm.create_image(48, 1.5e-3, ipc, [1e8])
    ra, dec = imaging.image_centre_radec(im)
    sc = mm.pixel_to_skycoord(48 // 2 + 1, 48 // 2 + 1, im.image_acc.wcs, origin=1)
    assert abs(ra - sc.ra.rad) < 1e-15 and abs(dec - sc.dec.rad) < 1e-15
    assert abs(imaging.separation(pc.ra.rad, pc.dec.rad, ra, dec) - pc.separation(sc).rad) < 1e-16
    l, m, nm1 = imaging.skycoord_to_lmn(ra, dec, pc.ra.rad, pc.dec.rad)
    dc = sc.transform_to(pc.skyoffset_frame()).represent_as(None)
    assert np.allclose((l, m, nm1), (dc.y.value, dc.z.value, dc.x.value - 1), rtol=0, atol=1e-15)


def test_image_centre_off_reference_pixel():
    """crpix away from the FFT centre: SIN deprojection equals the stand-in WCS."""
    pc = mm.SkyCoord.from_deg(10.0, -60.0)
    im = mm.create_image(64, 1e-3, pc, [1e8])
    im.attrs["wcs"].wcs.crpix[:2] = [20.0, 40.0]
    ra, dec = imaging.image_centre_radec(im)
    sc = mm.pixel_to_skycoord(33, 33, im.image_acc.wcs, origin=1)
    assert abs(ra - sc.ra.rad) < 1e-15 and abs(dec - sc.dec.rad) < 1e-15
    assert pc.separation(sc).rad > 1e-3


def test_shift_vector_noop_when_centres_coincide():
    g, vis, im = load_ng_case("ng_case_a.npz")
    assert imaging._shift_vector(vis, im) == (None, False)
    g, vis, im = load_ng_case("ng_case_b.npz")
    shift, update = imaging._shift_vector(vis, im)
    assert update and shift is not None and abs(shift[0]) > 1e-3


@pytest.mark.parametrize("a,b", [("linear", "stokesIQUV"), ("circular", "stokesIQUV"),
                                 ("linearnp", "stokesIQ"), ("circularnp", "stokesIV")])
def test_pol_matrices_invert_each_other(a, b):
    fa, fb = mm.PolarisationFrame(a), mm.PolarisationFrame(b)
    M, Mi = polarisation.pol_matrix(fa, fb), polarisation.pol_matrix(fb, fa)
    assert np.allclose(M @ Mi, np.eye(fa.npol), atol=1e-15)


def test_pol_matrix_linear_convention():
    """XX = I + Q, XY = U + iV, YX = U - iV, YY = I - Q."""
    M = polarisation.pol_matrix(mm.PolarisationFrame("stokesIQUV"), mm.PolarisationFrame("linear"))
    iquv = np.array([1.0, 0.2, 0.3, 0.4])
    assert np.allclose(M @ iquv, [1.2, 0.3 + 0.4j, 0.3 - 0.4j, 0.8])
    assert np.array_equal(polarisation.pol_matrix(mm.PolarisationFrame("stokesI"),
                                                  mm.PolarisationFrame("stokesI")), np.eye(1))
    with pytest.raises(ValueError):
        polarisation.pol_matrix(mm.PolarisationFrame("linear"), mm.PolarisationFrame("stokesIV"))


def test_vis_to_im_mapping():
    g, vis, im = load_ng_case("ng_case_b.npz")
    assert list(imaging._vis_to_im(im, vis.frequency.data)) == [0, 0, 1, 1]


def test_error_behaviour_matches_reference():
    g, vis, im = load_ng_case("ng_case_a.npz")
    assert imaging.predict_ng(vis, None) is vis                      # ng.py:52-53
    with pytest.raises(AssertionError):
        imaging.invert_ng(vis, "not an image")                       # ng.py:172
    with pytest.raises(AssertionError):
        imaging.predict_ng(vis, object())                            # ng.py:55
    bad = im.copy(deep=True)
    bad.attrs["wcs"].wcs.ctype[0] = "RA---TAN"
    with pytest.raises(AssertionError):
        imaging.invert_ng(vis, bad)                                  # ng.py:173 is_canonical
    with pytest.raises(ValueError):
        imaging.invert_visibility(vis, im, context="nonsense")       # imaging.py:105
    with pytest.raises(ValueError):
        imaging.predict_visibility(vis, im, context="nonsense")      # imaging.py:55


def test_npol_mismatch_asserts():
    g, vis, im = load_ng_case("ng_case_a.npz")
    im4 = mm.create_image(64, 1e-3, vis.phasecentre, [1e8],
                          polarisation_frame=mm.PolarisationFrame("stokesIQUV"))
    with pytest.raises((AssertionError, Exception)) as exc:
        imaging.predict_ng(vis, im4)                                 # ng.py:78
    assert exc.type is AssertionError or "CUDA" in str(exc.value)


def test_no_cpu_fallback_in_wrapper():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from ska_sdp_func_python_b200.wgridder import WgridderError
    g, vis, im = load_ng_case("ng_case_a.npz")
    with pytest.raises(WgridderError):
        imaging.invert_ng(vis, im)
    with pytest.raises(WgridderError):
        imaging.predict_ng(vis, im)


def test_inputs_are_not_mutated_by_standin_copy():
    g, vis, im = load_ng_case("ng_case_a.npz")
    c = vis.copy(deep=True, zero=True)
    assert np.all(c["vis"].data == 0) and np.any(vis["vis"].data != 0)
    assert np.array_equal(c["uvw"].data, vis["uvw"].data)
    c["uvw"].data[...] = 0
    assert np.any(vis["uvw"].data != 0)


# ---- pins that do not go through the repo's own tables / formulas ------------------------------
def test_pol_matrices_on_a_typed_in_stokes_vector():
    """Thompson, Moran & Swenson eq. 4.58 / 4.59 as ska_sdp_datamodels applies them, on the Stokes
    vector (I, Q, U, V) = (1.0, 0.3, -0.2, 0.1) with the expected correlations typed in by hand:
    XX = I + Q, XY = U + iV, YX = U - iV, YY = I - Q;  RR = I + V, RL = -iQ + U, LR = -iQ - U,
    LL = I - V.  Independent of polarisation.py's tables."""
    from ska_sdp_func_python_b200.polarisation import pol_matrix_table
    stokes = np.array([1.0, 0.3, -0.2, 0.1], dtype=complex)
    linear = np.array([1.3, -0.2 + 0.1j, -0.2 - 0.1j, 0.7])
    circular = np.array([1.1, -0.2 - 0.3j, 0.2 - 0.3j, 0.9])
    assert np.allclose(pol_matrix_table("stokesIQUV", "linear") @ stokes, linear, atol=1e-15)
    assert np.allclose(pol_matrix_table("stokesIQUV", "circular") @ stokes, circular, atol=1e-15)
    assert np.allclose(pol_matrix_table("linear", "stokesIQUV") @ linear, stokes, atol=1e-15)
    assert np.allclose(pol_matrix_table("circular", "stokesIQUV") @ circular, stokes, atol=1e-15)
    # two-product frames: (XX, YY) = (I + Q, I - Q), (RR, LL) = (I + V, I - V)
    assert np.allclose(pol_matrix_table("stokesIQ", "linearnp") @ np.array([1.0, 0.3]), [1.3, 0.7])
    assert np.allclose(pol_matrix_table("stokesIV", "circularnp") @ np.array([1.0, 0.1]), [1.1, 0.9])
    assert np.allclose(pol_matrix_table("linearnp", "stokesIQ") @ np.array([1.3, 0.7]), [1.0, 0.3])
    assert np.allclose(pol_matrix_table("circularnp", "stokesIV") @ np.array([1.1, 0.9]), [1.0, 0.1])
    assert np.array_equal(pol_matrix_table("stokesI", "stokesI"), np.eye(1))


def test_geometry_against_typed_in_closed_form_values():
    """SIN deprojection of an off-centre FFT-centre pixel and (l, m, n-1) of a sky position, against
    values computed independently with 30-digit arithmetic from the rotation of the tangent-plane
    unit vector into equatorial coordinates (x = n cos d0 - m sin d0, y = l, z = n sin d0 + m cos d0;
    ra = ra0 + atan2(y, x), dec = asin z) and typed in here."""
    from ska_sdp_func_python_b200 import imaging
    from ska_sdp_func_python_b200 import minimal_models as mm
    # image 64 x 64, reference pixel moved so that the FFT centre (33, 33 one-based) sits at
    # (dx, dy) = (+3, -2) pixels from crpix; cdelt = (-0.01, +0.01) deg; crval = (15, -35) deg
    wcs = mm.WCS(crval=[15.0, -35.0, 1.0, 1e8], crpix=[30.0, 35.0, 1.0, 1.0], cdelt=[-0.01, 0.01, 1.0, 1e6])
    im = mm.Image(np.zeros((1, 1, 64, 64)), wcs, mm.PolarisationFrame("stokesI"))
    ra, dec = imaging.image_centre_radec(im)
    assert abs(ra - 0.261160035410117501731) < 1e-15
    assert abs(dec - (-0.611214208049086070517)) < 1e-15
    l, m, nm1 = imaging.skycoord_to_lmn(np.radians(15.5), np.radians(-34.2), np.radians(15.0), np.radians(-35.0))
    assert abs(l - 0.00721754799142246370904) < 1e-16
    assert abs(m - 0.0139441168719772890360) < 1e-16
    assert abs(nm1 - (-0.000123273295326594406871)) < 2e-16


def test_uvw_reprojection_matrix_equals_the_two_step_rotation():
    """phaserotate_visibility(tangent=False): base._reproject_matrix against the reference's two
    steps written out (uvw -> earth xyz at hour angle -ra_old, xyz -> uvw at -ra_new;
    util/coordinate_support.py:335-393), plus a typed-in case."""
    from ska_sdp_func_python_b200 import base

    def uvw_to_xyz(uvw, ha, dec):
        u, v, w = uvw
        v0 = v * np.sin(dec) - w * np.cos(dec)
        z = v * np.cos(dec) + w * np.sin(dec)
        return np.array([u * np.cos(ha) + v0 * np.sin(ha), -u * np.sin(ha) + v0 * np.cos(ha), z])

    def xyz_to_uvw(xyz, ha, dec):
        x, y, z = xyz
        v0 = x * np.sin(ha) + y * np.cos(ha)
        return np.array([x * np.cos(ha) - y * np.sin(ha), z * np.cos(dec) + v0 * np.sin(dec),
                         z * np.sin(dec) - v0 * np.cos(dec)])

    rng = np.random.default_rng(4)
    for _ in range(10):
        ra0, dec0, ra1, dec1 = rng.uniform(-1.4, 1.4, 4)
        b = rng.normal(size=3) * 1e3
        want = xyz_to_uvw(uvw_to_xyz(b, -ra0, dec0), -ra1, dec1)
        R = base._reproject_matrix(ra0, dec0, ra1, dec1)
        assert np.abs(R @ b - want).max() < 1e-12
        assert np.abs(R @ R.T - np.eye(3)).max() < 1e-15
    # phase centre (ra 0, dec 0) -> (ra 90 deg, dec 0): old w (towards ra 0) becomes -u, old u
    # (east at ra 0, i.e. towards ra 90 deg) becomes w, v keeps pointing north
    R = base._reproject_matrix(0.0, 0.0, np.pi / 2, 0.0)
    assert np.allclose(R, [[0.0, 0.0, -1.0], [0.0, 1.0, 0.0], [1.0, 0.0, 0.0]], atol=1e-15)
