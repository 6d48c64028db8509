"""Drop-in density weighting / tapering of the imaging weights on a B200 (SURVEY.md section 8f,
row N1: the step immediately before ``invert_ng``).

Mirrors, with the same names, argument meaning and assertion behaviour,
  * ``weight_visibility``                     /root/reference/src/ska_sdp_func_python/imaging/weighting.py:35-65
  * ``taper_visibility_gaussian`` / ``_tukey`` imaging/weighting.py:68-140
  * ``grid_visibility_weight_to_griddata``    grid_data/gridding.py:258-334
  * ``griddata_visibility_reweight``          grid_data/gridding.py:362-499
The reference runs interpreted per-row loops (gridding.py:301-324); here the weight cube is
uploaded once and three streaming kernels (csrc/weight_kernels.cu) do the nearest-cell histogram,
the uniform / robust re-weighting and the tapers.  There is no CPU fallback.

Like the reference these functions modify ``vis`` in place (``imaging_weight``) and return it.
The array-level entry points ``grid_weights`` / ``reweight`` / ``taper`` work on numpy arrays or
CUDA torch tensors ([nrow, nchan, npol] cubes) and return torch tensors on the device.
"""
import ctypes

import numpy

from . import _lib
from .imaging import (C_M_S, _flags_for_device, _is_image, _ptr, _torch, _vis_to_im)

__all__ = ["weight_visibility", "taper_visibility_gaussian", "taper_visibility_tukey",
           "grid_visibility_weight_to_griddata", "griddata_visibility_reweight",
           "grid_weights", "reweight", "taper"]


class WeightGrid:
    """Device context of one weighting job: uvw, f/c, channel map and the UU / VV axes."""

    def __init__(self, uvw, freq, vis_to_im, nchan_img, nx, ny, cdelt_u, cdelt_v, crpix_u, crpix_v):
        torch = _torch()
        self.torch = torch
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        self.uvw = self._up(uvw, torch.float64).reshape(-1, 3)
        freq = numpy.ascontiguousarray(freq, dtype=numpy.float64)
        self.kscale = self._up(freq / C_M_S, torch.float64)   # as visibility/base.py:54
        self.wave = self._up(C_M_S / freq, torch.float64)     # as weighting.py:89, :122
        self.vis_to_im = self._up(numpy.ascontiguousarray(vis_to_im, dtype=numpy.int32), torch.int32)
        self.nchan_img, self.nx, self.ny = int(nchan_img), int(nx), int(ny)
        self.axes = (float(cdelt_u), float(cdelt_v), float(crpix_u), float(crpix_v))
        self.nrow, self.nchan = self.uvw.shape[0], len(freq)

    def _up(self, a, dtype=None):
        torch = self.torch
        if isinstance(a, torch.Tensor):
            t = a.to(self.dev)
        else:
            a = numpy.asarray(a)
            if a.nbytes >= (1 << 24):
                return self._up_staged(a, dtype)
            t = torch.from_numpy(numpy.ascontiguousarray(a)).to(self.dev)
        if dtype is not None and t.dtype != dtype:
            t = t.to(dtype)
        return t.contiguous()

    def _up_staged(self, a, dtype, chunk_bytes=1 << 28):
        """Large host arrays go up in chunks through pinned buffers (_xfer.Stager), converted to
        ``dtype`` during the staging copy."""
        from ._xfer import Stager
        torch = self.torch
        a = numpy.ascontiguousarray(a)
        flat = a.reshape(-1)
        dt = dtype or torch.from_numpy(flat[:1]).dtype
        out = torch.empty(a.shape, device=self.dev, dtype=dt)
        oflat = out.reshape(-1)
        st = Stager(torch, self.dev, slots=3)
        step = max(1, chunk_bytes // a.dtype.itemsize)
        for i0 in range(0, flat.shape[0], step):
            oflat[i0:i0 + step].copy_(st.put(flat[i0:i0 + step], dt), non_blocking=True)
        st.finish()
        return out

    def cube(self, a):
        """[nrow, nchan, npol] float64 on the device."""
        t = self._up(a, self.torch.float64)
        npol = t.shape[-1] if t.dim() >= 3 else 1
        return t.reshape(self.nrow, self.nchan, npol)

    def flags(self, flags):
        """(device tensor or None, B2G_FLAG_* code)"""
        if flags is None:
            return None, _lib.FLAG_NONE
        if isinstance(flags, self.torch.Tensor):
            codes = {self.torch.uint8: _lib.FLAG_U8, self.torch.bool: _lib.FLAG_U8,
                     self.torch.int32: _lib.FLAG_I32, self.torch.int64: _lib.FLAG_I64,
                     self.torch.float64: _lib.FLAG_F64}
            if flags.dtype not in codes:
                flags = flags.to(self.torch.float64)
            return flags.to(self.dev).contiguous(), codes[flags.dtype]
        f, code = _flags_for_device(numpy.asarray(flags))
        if f.dtype.kind in "iu" and f.dtype.itemsize > 1 and f.size:
            lo, hi = self.torch.from_numpy(f.reshape(-1)).aminmax()
            if int(lo) >= 0 and int(hi) <= 255:     # 0 / 1 flags travel as bytes
                return self._up(f, self.torch.uint8), _lib.FLAG_U8
        return self._up(f), code


def grid_weights(G, weight, flags=None):
    """Histogram of weight*(1-flags) on the grid of ``G``.
    Returns (grid [nchan_img, npol, ny, nx] float64 CUDA tensor, sumwt [nchan_img, npol] numpy,
    sum(grid**2) float)."""
    torch = G.torch
    w = G.cube(weight)
    npol = w.shape[2]
    f, fcode = G.flags(flags)
    grid = torch.empty((G.nchan_img, npol, G.ny, G.nx), device=G.dev, dtype=torch.float64)
    sumwt = numpy.zeros((G.nchan_img, npol))
    sumsq = ctypes.c_double(0.0)
    _lib.check(_lib.lib().b2g_weight_grid(
        G.nrow, G.nchan, npol, _ptr(G.uvw), _ptr(G.kscale), _ptr(w), _ptr(f), fcode,
        _ptr(G.vis_to_im), G.nchan_img, G.nx, G.ny, *G.axes, _ptr(grid),
        sumwt.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), ctypes.byref(sumsq), G.stream))
    return grid, sumwt, float(sumsq.value)


def reweight(G, weight, flags, grid, weighting="uniform", f2=0.0):
    """imaging_weight [nrow, nchan, npol] (CUDA tensor) from the gridded weights."""
    torch = G.torch
    w = G.cube(weight)
    npol = w.shape[2]
    f, fcode = G.flags(flags)
    out = torch.empty_like(w)
    mode = {"uniform": 0, "robust": 1}[weighting]
    _lib.check(_lib.lib().b2g_weight_apply(
        G.nrow, G.nchan, npol, _ptr(G.uvw), _ptr(G.kscale), _ptr(w), _ptr(f), fcode,
        _ptr(G.vis_to_im), G.nchan_img, G.nx, G.ny, *G.axes, _ptr(grid), mode, float(f2),
        _ptr(out), G.stream))
    return out


def taper(G, imaging_weight, flags, kind, param):
    """imaging_weight * (1-flags) * taper; ``kind`` "gaussian" (param = beam FWHM, rad) or
    "tukey" (param = transition fraction).  Returns a new CUDA tensor."""
    imw = G.cube(imaging_weight).clone()
    f, fcode = G.flags(flags)
    if kind == "gaussian":
        mode, p = 0, numpy.pi ** 2 * param ** 2 / (4.0 * numpy.log(2.0))
    elif kind == "tukey":
        mode, p = 1, param
    else:
        raise ValueError(f"unknown taper {kind}")
    _lib.check(_lib.lib().b2g_weight_taper(G.nrow, G.nchan, imw.shape[2], _ptr(G.uvw), _ptr(G.wave),
                                           _ptr(f), fcode, mode, float(p), _ptr(imw), G.stream))
    return imw


# ---- data-model level (drop-ins) ---------------------------------------------------------------
def _grid_from_wcs(vis, gwcs, shape):
    """WeightGrid for a Visibility and a GridData-style WCS (UU, VV, STOKES, FREQ)."""
    w = gwcs.wcs
    nchan_img, _, ny, nx = shape
    freq = numpy.asarray(vis.frequency.data)
    vis_to_im = numpy.round(gwcs.sub([4]).wcs_world2pix(freq, 0)[0]).astype("int")
    return WeightGrid(numpy.asarray(vis.uvw.data), freq, vis_to_im, nchan_img, nx, ny,
                      w.cdelt[0], w.cdelt[1], w.crpix[0], w.crpix[1])


def _grid_from_image(vis, model):
    """What create_griddata_from_image (external datamodels) derives from the image WCS:
    cdelt = 1 / (n * cdelt_image[rad]), crpix = n // 2 + 1."""
    nchan_img, _, ny, nx = model["pixels"].data.shape
    w = model.image_acc.wcs.wcs
    d2r = numpy.pi / 180.0
    freq = numpy.asarray(vis.frequency.data)
    return WeightGrid(numpy.asarray(vis.uvw.data), freq, _vis_to_im(model, freq), nchan_img, nx, ny,
                      1.0 / (nx * w.cdelt[0] * d2r), 1.0 / (ny * w.cdelt[1] * d2r),
                      nx // 2 + 1, ny // 2 + 1)


def _store(vis, imw_dev, chunk_bytes=1 << 28):
    """imaging_weight of ``vis`` <- device tensor, in place, through pinned buffers."""
    dst = vis["imaging_weight"].data
    if not (isinstance(dst, numpy.ndarray) and dst.flags.c_contiguous and dst.dtype == numpy.float64
            and dst.flags.writeable):
        dst[...] = imw_dev.cpu().numpy().reshape(dst.shape)
        return vis
    from ._xfer import Stager
    torch = _torch()
    st = Stager(torch, imw_dev.device, slots=3)
    flat, src = dst.reshape(-1), imw_dev.reshape(-1)
    step = max(1, chunk_bytes // 8)
    for i0 in range(0, flat.shape[0], step):
        st.get(src[i0:i0 + step], flat[i0:i0 + step])
    st.finish()
    return vis


def _robust_f2(robustness, sumwt_total, sumsq):
    return float((5.0 * numpy.power(10.0, -robustness)) ** 2 * sumwt_total / sumsq)


def grid_visibility_weight_to_griddata(vis, griddata):
    """Grid Visibility weight onto a GridData (gridding.py:258-334).  Returns (griddata, sumwt)."""
    assert vis.visibility_acc.polarisation_frame == griddata.griddata_acc.polarisation_frame
    G = _grid_from_wcs(vis, griddata.griddata_acc.griddata_wcs, griddata["pixels"].data.shape)
    grid, sumwt, _ = grid_weights(G, vis["weight"].data, vis["flags"].data)
    griddata["pixels"].data = grid.cpu().numpy().astype("complex")
    return griddata, sumwt


def griddata_visibility_reweight(vis, griddata, weighting="uniform", robustness=0.0, sumwt=None):
    """Reweight visibility weight using the weights in griddata (gridding.py:362-499)."""
    if griddata is not None:
        assert vis.visibility_acc.polarisation_frame == griddata.griddata_acc.polarisation_frame
    assert weighting in ["natural", "uniform", "robust"], f"Weighting {weighting} not supported"
    if weighting == "natural":
        vis["imaging_weight"].data[...] = vis["weight"].data[...]
        return vis
    G = _grid_from_wcs(vis, griddata.griddata_acc.griddata_wcs, griddata["pixels"].data.shape)
    real_gd = numpy.real(griddata["pixels"].data)
    grid = G._up(real_gd, G.torch.float64)
    f2 = 0.0
    if weighting == "robust":
        sumlocwt = numpy.sum(real_gd ** 2)  # the grid is a host array here: the reference's own sum
        if sumwt is None:
            total = numpy.sum(vis.visibility_acc.flagged_weight) * 2
        else:
            total = numpy.sum(sumwt)
        f2 = _robust_f2(robustness, total, sumlocwt)
    return _store(vis, reweight(G, vis["weight"].data, vis["flags"].data, grid, weighting, f2))


def weight_visibility(vis, model, weighting="uniform", robustness=0.0):
    """Weight the visibility data (weighting.py:35-65): natural, uniform or robust.  The weight
    grid never leaves the device."""
    assert _is_image(model), model
    assert model.image_acc.is_canonical()
    assert weighting in ["natural", "uniform", "robust"], f"Weighting {weighting} not supported"
    if weighting == "natural":
        vis["imaging_weight"].data[...] = vis["weight"].data[...]
        return vis
    G = _grid_from_image(vis, model)
    w = G.cube(vis["weight"].data)
    f, _ = G.flags(vis["flags"].data)
    grid, sumwt, sumsq = grid_weights(G, w, f)
    f2 = _robust_f2(robustness, numpy.sum(sumwt), sumsq) if weighting == "robust" else 0.0
    return _store(vis, reweight(G, w, f, grid, weighting, f2))


def _taper_vis(vis, kind, param):
    nchan = len(vis.frequency.data)
    G = WeightGrid(numpy.asarray(vis.uvw.data), numpy.asarray(vis.frequency.data),
                   numpy.zeros(nchan, dtype=numpy.int32), 1, 1, 1, 1.0, 1.0, 1.0, 1.0)
    return _store(vis, taper(G, vis["imaging_weight"].data, vis["flags"].data, kind, param))


def taper_visibility_gaussian(vis, beam=None):
    """Taper the imaging weights with a Gaussian of FWHM ``beam`` radians (weighting.py:68-101)."""
    if beam is None:
        raise ValueError("Beam size not specified for Gaussian taper")
    return _taper_vis(vis, "gaussian", beam)


def taper_visibility_tukey(vis, tukey=0.1):
    """Tukey taper of the imaging weights over the uv radius (weighting.py:104-140)."""
    return _taper_vis(vis, "tukey", tukey)
