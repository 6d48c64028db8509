"""apply_beam_to_skycomponent (sky_component/operations.py:366-450) on the SIN projection's definition:
host code, no GPU."""
import numpy as np
import pytest

from ska_sdp_func_python_b200 import minimal_models as mm
from ska_sdp_func_python_b200.sky_component import apply_beam_to_skycomponent


def beam_image(npix=32, nchan=2, npol=1):
    im = mm.create_image(npix, 2e-3, mm.SkyCoord.from_deg(40.0, -55.0), frequency=[1e8, 1.1e8],
                         polarisation_frame=mm.PolarisationFrame("stokesI"))
    rng = np.random.default_rng(5)
    im["pixels"].data[:] = rng.uniform(0.2, 1.0, im["pixels"].data.shape)
    return im


def comp_at(beam, ix, iy, flux=2.0):
    return mm.SkyComponent(beam.image_acc.wcs.pixel_to_skycoord(ix, iy, origin=0), [1e8, 1.1e8],
                           np.full((2, 1), flux), mm.PolarisationFrame("stokesI"))


def test_pixel_lookup_and_quirk():
    beam = beam_image()
    pix = beam["pixels"].data
    # the reference rounds the 1-BASED pixel position and indexes with it: pixel (ix, iy) -> [iy + 1, ix + 1]
    for ix, iy in ((16, 16), (3, 20), (29, 7)):
        out = apply_beam_to_skycomponent([comp_at(beam, ix, iy)], beam)
        assert np.allclose(out[0].flux, 2.0 * pix[:, :, iy + 1, ix + 1], rtol=1e-15)
    single = apply_beam_to_skycomponent(comp_at(beam, 10, 11), beam)
    assert not isinstance(single, list) and np.allclose(single.flux, 2.0 * pix[:, :, 12, 11])


def test_inverse_outside_and_input_untouched():
    beam = beam_image()
    pix = beam["pixels"].data
    c = comp_at(beam, 8, 9)
    f0 = c.flux.copy()
    inv = apply_beam_to_skycomponent([c], beam, inverse=True)[0]
    assert np.allclose(inv.flux, 2.0 / pix[:, :, 10, 9]) and np.array_equal(c.flux, f0)
    pix[:, :, 10, 9] = 0.0                                   # zero beam: inverse falls back to the product
    assert not apply_beam_to_skycomponent([c], beam, inverse=True)[0].flux.any()
    far = mm.SkyComponent(mm.SkyCoord.from_deg(48.0, -55.0), [1e8, 1.1e8], np.ones((2, 1)), mm.PolarisationFrame("stokesI"))
    out = apply_beam_to_skycomponent([c, far], beam)
    assert len(out) == 2 and not out[1].flux.any()           # outside the raster: zero flux, kept
    behind = mm.SkyComponent(mm.SkyCoord.from_deg(220.0, 55.0), [1e8, 1.1e8], np.ones((2, 1)), mm.PolarisationFrame("stokesI"))
    assert len(apply_beam_to_skycomponent([c, behind], beam)) == 1    # behind the tangent plane: dropped
    gauss = comp_at(beam, 5, 5)
    gauss.shape = "Gaussian"
    with pytest.raises(AssertionError):
        apply_beam_to_skycomponent([gauss], beam)


def test_non_sin_raster_uses_the_phase_centre():
    beam = beam_image()
    c = comp_at(beam, 12, 14)
    want = apply_beam_to_skycomponent([c], beam)[0].flux
    pc = beam.image_acc.phasecentre
    beam.image_acc.wcs.wcs.ctype[0] = "AZELGEO long"
    got = apply_beam_to_skycomponent([c], beam, phasecentre=pc)[0].flux
    assert np.array_equal(got, want)


def test_subtract_visibility():
    from ska_sdp_func_python_b200.skymodel_imaging import subtract_visibility
    rng = np.random.default_rng(2)
    shape = (2, 3, 4, 1)
    mk = lambda: mm.Visibility(rng.normal(size=shape) + 1j * rng.normal(size=shape), rng.normal(size=(2, 3, 3)),  # noqa: E731
                               1e8 + 1e6 * np.arange(4), mm.SkyCoord.from_deg(0, -30), mm.PolarisationFrame("stokesI"))
    a, b = mk(), mk()
    a0 = a["vis"].data.copy()
    r = subtract_visibility(a, b)
    assert r is not a and np.array_equal(r["vis"].data, a0 - b["vis"].data) and np.array_equal(a["vis"].data, a0)
    r2 = subtract_visibility(a, b, inplace=True)
    assert r2 is a and np.array_equal(a["vis"].data, a0 - b["vis"].data)
    c = mm.Visibility(np.zeros((2, 3, 5, 1), dtype=complex), rng.normal(size=(2, 3, 3)), 1e8 + 1e6 * np.arange(5),
                      mm.SkyCoord.from_deg(0, -30), mm.PolarisationFrame("stokesI"))
    with pytest.raises(AssertionError):
        subtract_visibility(a, c)
