"""Stand-alone forms of the two small steps of the path that ``invert_ng`` / ``predict_ng`` otherwise
fuse into their device kernels, for callers that use them
on their own:

* ``shift_vis_to_image`` / ``phaserotate_visibility`` (imaging/base.py:48-92, visibility/base.py:27-125):
  vis *= conj(phasor) (or phasor for ``inverse``), phasor =
  exp(-2 pi i uvw_lambda . [l, m, sqrt(1 - l^2 - m^2) - 1]), on the device through ``b2g_vis_finish``
  (exact range reduction of the phase, ``sincospi``);
* ``normalise_sumwt`` (imaging/base.py:95-156).
"""
import ctypes
import logging

import numpy

from . import _lib
from .imaging import (_Device, _dptr, _is_image, _ptr, _radec, _rows_per_chunk, image_centre_radec,
                      separation, skycoord_to_lmn)

log = logging.getLogger("func-python-logger")

__all__ = ["shift_vis_to_image", "phaserotate_visibility", "normalise_sumwt"]


def _rotate(vis, lmn, inverse):
    """vis["vis"] *= exp(-+ 2 pi i (u l + v m + w (n - 1)) f / c) on the device, chunk by chunk."""
    data = vis["vis"].data
    nrows, nbl, nchan, npol = data.shape
    nrow = nrows * nbl
    D = _Device(numpy.asarray(vis["uvw"].data, dtype=numpy.float64).reshape(nrow, 3),
                numpy.asarray(vis.frequency.data))
    torch = D.torch
    # b2g_vis_finish multiplies by exp(-2 pi i uvw . lmn): that is the phasor itself (inverse);
    # the forward rotation multiplies by its conjugate, i.e. the phasor of -lmn
    sgn = 1.0 if inverse else -1.0
    lmn = numpy.array([sgn * lmn[0], sgn * lmn[1], sgn * lmn[2]], dtype=numpy.float64)
    ident = numpy.ascontiguousarray(numpy.eye(npol, dtype=numpy.complex128))
    flat = numpy.ascontiguousarray(data.reshape(nrow, nchan, npol), dtype=numpy.complex128)
    out = numpy.empty_like(flat)
    step = max(1, _rows_per_chunk(nchan, npol, 16) // 2)
    L = _lib.lib()
    stage = D.stager()
    for r0 in range(0, nrow, step):
        r1 = min(nrow, r0 + step)
        rows = r1 - r0
        # the kernel reads polarisation planes [npol][rows][nchan] and writes [rows][nchan][npol]
        planes = stage.put(flat[r0:r1]).permute(2, 0, 1).contiguous()
        chunk = torch.empty((rows, nchan, npol), device=D.dev, dtype=torch.complex128)
        _lib.check(L.b2g_vis_finish(
            rows, nchan, npol, npol, _ptr(planes), _lib.F64, _dptr(ident.view(numpy.float64)),
            _dptr(lmn), _ptr(D.uvw[r0:r1]), _ptr(D.fscale), 0, rows, _ptr(chunk), _lib.F64,
            ctypes.c_void_p(D.stream)))
        stage.get(chunk, out[r0:r1])
    stage.finish()
    return out.reshape(data.shape)


def _reproject_matrix(ra_old, dec_old, ra_new, dec_new):
    """The 3 x 3 rotation uvw(old phase centre) -> uvw(new phase centre): out of the old frame
    into earth coordinates (hour angle -ra_old, declination dec_old) and into the new one, as the
    reference's uvw_to_xyz followed by xyz_to_uvw (util/coordinate_support.py:335-393) -- two
    rotations each, about the z axis by the hour angle and about the u axis by 90 deg - dec."""
    def to_xyz(ha, dec):      # rows: x, y, z as combinations of (u, v, w)
        sh, ch, sd, cd = numpy.sin(ha), numpy.cos(ha), numpy.sin(dec), numpy.cos(dec)
        return numpy.array([[ch, sh * sd, -sh * cd],
                            [-sh, ch * sd, -ch * cd],
                            [0.0, cd, sd]])
    # xyz_to_uvw is the inverse rotation, i.e. the transpose
    return to_xyz(-ra_new, dec_new).T @ to_xyz(-ra_old, dec_old)


def _phaserotate(vis, ra_n, dec_n, inverse, tangent=True):
    ra_v, dec_v = _radec(vis.phasecentre)
    l, m, nm1 = skycoord_to_lmn(ra_n, dec_n, ra_v, dec_v)
    if abs(nm1) < 1e-15:                                   # "no significant change", visibility/base.py:78
        return vis
    newvis = vis.copy(deep=True)
    s = (l, m, float(numpy.sqrt(1.0 - l * l - m * m) - 1.0))
    newvis["vis"].data = _rotate(newvis, s, inverse)
    if not tangent:
        # leave the tangent plane: re-project the baselines onto the new phase centre and move the
        # phase centre there (visibility/base.py:98-125; the reference rotates the same way for
        # both values of ``inverse``); uvw_lambda follows from uvw and frequency on demand
        uvw = numpy.asarray(newvis["uvw"].data, dtype=numpy.float64)
        R = _reproject_matrix(ra_v, dec_v, ra_n, dec_n)
        newvis["uvw"].data[...] = (uvw.reshape(-1, 3) @ R.T).reshape(uvw.shape)
    return newvis


def phaserotate_visibility(vis, newphasecentre, tangent=True, inverse=False):
    """Phase rotate to a new phase centre (visibility/base.py:60-125).  Returns ``vis`` itself when
    the change is insignificant, else a rotated deep copy; with ``tangent=False`` the uvw are
    re-projected onto the new phase centre and ``attrs["phasecentre"]`` moves there."""
    ra_n, dec_n = _radec(newphasecentre)
    out = _phaserotate(vis, ra_n, dec_n, inverse, tangent)
    if not tangent and out is not vis:
        out.attrs["phasecentre"] = newphasecentre
    return out


def shift_vis_to_image(vis, im, tangent=True, inverse=False):
    """Shift the visibility phases to the image's FFT centre pixel (imaging/base.py:48-92); sets
    ``attrs["phasecentre"]`` as the reference does."""
    ra_i, dec_i = image_centre_radec(im)
    ra_v, dec_v = _radec(vis.phasecentre)
    if separation(ra_v, dec_v, ra_i, dec_i) > 1e-15:
        vis = _phaserotate(vis, ra_i, dec_i, inverse, tangent)
        vis.attrs["phasecentre"] = im.image_acc.phasecentre
    return vis


def _normalise_by_weight_image(im, sumwt, min_weight, flat_sky):
    """The Image-valued ``sumwt`` case (imaging/base.py:129-149) as whole-cube expressions; both
    arguments are updated in place like the reference does.  Its per-plane loop takes the square
    root of the WHOLE weight cube once per (chan, pol) plane, so plane j of K (loop order) leaves as
    (W_j^(2^-j) / max W)^(2^-(K-j)); with ``flat_sky`` the weights are left alone."""
    w = sumwt["pixels"].data
    pix = im["pixels"].data
    maxwt = numpy.max(w)
    nchan, npol, ny, nx = w.shape
    if flat_sky:
        norm = numpy.sqrt(w[:, :, ny // 2:ny // 2 + 1, nx // 2:nx // 2 + 1] * w)
        good = norm > min_weight * maxwt
        pix[...] = numpy.where(good, pix / numpy.where(good, norm, 1.0), pix / maxwt)
        return im
    k = nchan * npol
    j = numpy.arange(k, dtype=float).reshape(nchan, npol, 1, 1)
    pix /= maxwt
    sumwt["pixels"].data = (w ** (0.5 ** j) / maxwt) ** (0.5 ** (k - j))
    return im


def normalise_sumwt(im, sumwt, min_weight=0.1, flat_sky=False):
    """Divide out the sum of weights (imaging/base.py:95-156): per (channel, polarisation) for a
    2-D ``sumwt`` (planes with sumwt <= 0 are zeroed), pixel by pixel / by its maximum for an Image."""
    nchan, npol, _, _ = im["pixels"].data.shape
    assert sumwt is not None
    if isinstance(sumwt, numpy.ndarray):
        assert nchan == sumwt.shape[0]
        assert npol == sumwt.shape[1]
        pix = im["pixels"].data
        for chan in range(nchan):
            for pol in range(npol):
                if sumwt[chan, pol] > 0.0:
                    pix[chan, pol, :, :] = pix[chan, pol, :, :] / sumwt[chan, pol]
                else:
                    pix[chan, pol, :, :] = 0.0
        return im
    if _is_image(sumwt) and im["pixels"].data.shape == sumwt["pixels"].data.shape:
        return _normalise_by_weight_image(im, sumwt, min_weight, flat_sky)
    raise ValueError("sumwt is not a 2D or 4D array - cannot perform normalisation")
