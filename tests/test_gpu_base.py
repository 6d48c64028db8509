"""Stand-alone shift_vis_to_image / phaserotate_visibility / normalise_sumwt (SURVEY.md section 8a rows
a5, a6) against the reference's formulas written out in numpy (visibility/base.py:27-45, :60-96;
imaging/base.py:48-156)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
C = 299792458.0


@pytest.fixture(scope="module")
def mods(built):
    import torch
    assert torch.cuda.is_available()
    from ska_sdp_func_python_b200 import base, minimal_models as mm
    return base, mm


def make_vis(mm, npol=4, seed=3):
    rng = np.random.default_rng(seed)
    nt, nbl, nchan = 3, 21, 5
    pf = mm.PolarisationFrame("linear" if npol == 4 else "stokesI")
    vis = rng.normal(size=(nt, nbl, nchan, npol)) + 1j * rng.normal(size=(nt, nbl, nchan, npol))
    uvw = rng.normal(size=(nt, nbl, 3)) * [800.0, 800.0, 60.0]
    return mm.Visibility(vis, uvw, 1e8 + 2e6 * np.arange(nchan), mm.SkyCoord.from_deg(15.0, -35.0), pf)


def numpy_phasor(vis, l, m):
    s = np.array([l, m, np.sqrt(1 - l * l - m * m) - 1.0])
    uvw_lambda = vis["uvw"].data[:, :, None, :] * (vis.frequency.data / C)[None, None, :, None]
    return np.exp(-2j * np.pi * np.einsum("tbfs,s->tbf", uvw_lambda, s))[..., np.newaxis]


@pytest.mark.parametrize("npol", [1, 4])
def test_shift_vis_to_image_both_directions(mods, npol):
    base, mm = mods
    from ska_sdp_func_python_b200.imaging import image_centre_radec, skycoord_to_lmn
    vis = make_vis(mm, npol)
    pf = mm.PolarisationFrame("stokesIQUV" if npol == 4 else "stokesI")
    im = mm.create_image(64, 1e-3, mm.SkyCoord.from_deg(15.6, -34.7), frequency=[1e8], polarisation_frame=pf)
    ra_i, dec_i = image_centre_radec(im)
    l, m, _ = skycoord_to_lmn(ra_i, dec_i, vis.phasecentre.ra.rad, vis.phasecentre.dec.rad)
    ph = numpy_phasor(vis, l, m)
    v0 = vis["vis"].data.copy()
    fwd = base.shift_vis_to_image(vis, im, tangent=True, inverse=False)
    assert fwd is not vis and np.array_equal(vis["vis"].data, v0)          # input untouched
    assert np.abs(fwd["vis"].data - v0 * np.conj(ph)).max() < 1e-12
    assert fwd.attrs["phasecentre"].separation(im.image_acc.phasecentre).rad < 1e-15
    back = base.phaserotate_visibility(fwd_like(mm, vis, fwd), mm.SkyCoord(ra_i, dec_i), inverse=True)
    assert np.abs(back["vis"].data - v0).max() < 1e-12


def test_phaserotate_off_the_tangent_plane(mods):
    """tangent=False (visibility/base.py:98-125): same phasor, uvw re-projected onto the new phase
    centre, phase centre moved.  Checked against the geometry itself: the new w of a baseline is
    its projection on the new direction, u stays perpendicular to the new direction and to the
    celestial pole, lengths are preserved, and rotating back returns the original uvw."""
    base, mm = mods
    from ska_sdp_func_python_b200.imaging import skycoord_to_lmn
    vis = make_vis(mm, 1)
    new = mm.SkyCoord.from_deg(17.5, -33.0)
    ra0, dec0 = vis.phasecentre.ra.rad, vis.phasecentre.dec.rad
    l, m, nm1 = skycoord_to_lmn(new.ra.rad, new.dec.rad, ra0, dec0)
    out = base.phaserotate_visibility(vis, new, tangent=False)
    same = base.phaserotate_visibility(vis, new, tangent=True)
    assert np.array_equal(out["vis"].data, same["vis"].data)                 # identical phasor
    assert out.attrs["phasecentre"].separation(new).rad < 1e-15
    assert same.attrs["phasecentre"].separation(vis.phasecentre).rad < 1e-15
    uvw0, uvw1 = vis["uvw"].data.reshape(-1, 3), out["uvw"].data.reshape(-1, 3)
    assert np.allclose(np.linalg.norm(uvw1, axis=1), np.linalg.norm(uvw0, axis=1), rtol=1e-13)
    # w' = b . s_new, with s_new = (l, m, n) in the old frame
    assert np.allclose(uvw1[:, 2], uvw0 @ np.array([l, m, nm1 + 1.0]), rtol=0, atol=1e-9)
    # typed-in check of one rotation: phase centre (0, 0) -> (90 deg, 0): the old w axis (towards
    # ra 0) becomes -u, the old u axis (east at ra 0 = towards ra 90 deg) becomes w, v stays north
    R = base._reproject_matrix(0.0, 0.0, np.pi / 2, 0.0)
    assert np.allclose(R, [[0.0, 0.0, -1.0], [0.0, 1.0, 0.0], [1.0, 0.0, 0.0]], atol=1e-15)
    back = base.phaserotate_visibility(out, vis.phasecentre, tangent=False)
    assert np.allclose(back["uvw"].data, vis["uvw"].data, rtol=0, atol=1e-9)
    assert np.abs(out.visibility_acc.uvw_lambda.data[1, 2, 3]
                  - out["uvw"].data[1, 2] * vis.frequency.data[3] / C).max() < 1e-9


def fwd_like(mm, vis, fwd):
    """fwd's data with vis's original phase centre, so that the inverse rotation sees the same shift."""
    out = vis.copy(deep=True)
    out["vis"].data = fwd["vis"].data.copy()
    return out


def test_shift_is_a_no_op_on_the_phase_centre(mods):
    base, mm = mods
    vis = make_vis(mm, 1)
    im = mm.create_image(64, 1e-3, vis.phasecentre, frequency=[1e8])
    assert base.shift_vis_to_image(vis, im) is vis


def test_normalise_sumwt(mods):
    base, mm = mods
    rng = np.random.default_rng(0)
    im = mm.create_image(16, 1e-3, mm.SkyCoord.from_deg(0, -30), frequency=[1e8, 1.1e8],
                         polarisation_frame=mm.PolarisationFrame("stokesIQ"))
    im["pixels"].data[:] = rng.normal(size=(2, 2, 16, 16))
    d0 = im["pixels"].data.copy()
    sumwt = np.array([[4.0, 0.0], [2.5, -1.0]])
    out = base.normalise_sumwt(im, sumwt)
    assert out is im
    assert np.array_equal(im["pixels"].data[0, 0], d0[0, 0] / 4.0) and np.array_equal(im["pixels"].data[1, 0], d0[1, 0] / 2.5)
    assert not im["pixels"].data[0, 1].any() and not im["pixels"].data[1, 1].any()
    wt = im.copy(deep=True)
    wt["pixels"].data[:] = rng.uniform(0.5, 2.0, size=(2, 2, 16, 16))
    im["pixels"].data[:] = d0
    mx = wt["pixels"].data.max()
    base.normalise_sumwt(im, wt)
    assert np.allclose(im["pixels"].data, d0 / mx, rtol=1e-15)
    with pytest.raises(ValueError):
        base.normalise_sumwt(im, [1.0])
