"""``apply_beam_to_skycomponent`` (sky_component/operations.py:366-450 of the reference) without its
astropy dependency: the sky-to-pixel step is the FITS SIN projection written out on the image's
``wcs.wcs.crval / crpix / cdelt`` (canonical images: no rotation matrix), so it works on the
datamodels' astropy WCS and on ``minimal_models.WCS`` alike.  Component lists are a handful of
entries: host code, used by the major-cycle glue (``skymodel_imaging``) when a sky model combines
components with a mask or a primary beam.

Parity note: the reference takes the pixel position from ``astropy.wcs.utils.skycoord_to_pixel``,
which is not installed in the build container, so this function is checked against the SIN
projection's definition only (tests/test_sky_component.py): PARITY UNPINNED.  Kept from the
reference: the 1-based pixel position (``origin=1``) is rounded and used as a 0-based index.
"""
import collections.abc
import copy
import logging

import numpy

log = logging.getLogger("func-python-logger")

__all__ = ["apply_beam_to_skycomponent"]


def _sin_pixel(ra, dec, wcs, centre=None):
    """1-based pixel coordinates of (ra, dec) [rad] in a SIN-projected image; NaN behind the
    tangent plane."""
    w = wcs.wcs
    ra0, dec0 = (numpy.radians(w.crval[0]), numpy.radians(w.crval[1])) if centre is None else centre
    dra = ra - ra0
    if numpy.sin(dec) * numpy.sin(dec0) + numpy.cos(dec) * numpy.cos(dec0) * numpy.cos(dra) < 0.0:
        return float("nan"), float("nan")
    x = numpy.degrees(numpy.cos(dec) * numpy.sin(dra))
    y = numpy.degrees(numpy.sin(dec) * numpy.cos(dec0) - numpy.cos(dec) * numpy.sin(dec0) * numpy.cos(dra))
    return w.crpix[0] + x / w.cdelt[0], w.crpix[1] + y / w.cdelt[1]


def _radec(direction):
    ra, dec = direction.ra, direction.dec
    return (float(getattr(ra, "rad", getattr(ra, "radian", ra))),
            float(getattr(dec, "rad", getattr(dec, "radian", dec))))


def apply_beam_to_skycomponent(sc, beam, phasecentre=None, inverse=False):
    """Multiply (or, with ``inverse``, divide) component fluxes by the beam value at their pixel.
    Components outside the image get zero flux; components behind the tangent plane are dropped."""
    single = not isinstance(sc, collections.abc.Iterable)
    if single:
        sc = [sc]
    pix = beam["pixels"].data
    ny, nx = pix.shape[2], pix.shape[3]
    centre = None
    if beam.image_acc.wcs.wcs.ctype[0] != "RA---SIN":
        centre = _radec(phasecentre)                     # treat the raster as SIN about phasecentre
    newsc = []
    for comp in sc:
        assert comp.shape == "Point", f"Cannot handle shape {comp.shape}"
        ra, dec = _radec(comp.direction)
        px, py = _sin_pixel(ra, dec, beam.image_acc.wcs, centre)
        if numpy.isnan(px) or numpy.isnan(py):
            continue
        x, y = int(round(float(px))), int(round(float(py)))
        if 0 <= x < nx and 0 <= y < ny:
            if inverse and (pix[:, :, y, x] != 0.0).all():
                flux = comp.flux / pix[:, :, y, x]
            else:
                flux = comp.flux * pix[:, :, y, x]
        else:
            flux = 0.0 * comp.flux
        new = copy.copy(comp)
        new.flux = flux
        newsc.append(new)
    log.debug("apply_beam_to_skycomponent: %d components", len(newsc))
    return newsc[0] if single else newsc
