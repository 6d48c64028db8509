"""Drop-in ``invert_ng`` / ``predict_ng`` (and the ``invert_visibility`` / ``predict_visibility``
dispatch for the ``"ng"`` and ``"2d"`` contexts) on a B200.

Mirrors /root/reference/src/ska_sdp_func_python/imaging/ng.py:38-143 (predict_ng), :146-294
(invert_ng) and imaging/imaging.py:28-105 (dispatch): same names, argument meaning, kwargs read
with ``.get`` (``threads``, ``epsilon``, ``do_wstacking``, ``verbosity``; unknown keys ignored),
same ``dopsf`` / ``normalise`` / ``sumwt`` semantics, same assertion behaviour, inputs never
mutated, outputs fresh objects.  What differs is where the work happens: the Visibility cube is
uploaded once, ONE streaming kernel applies phase rotation + flags + polarisation conversion
(csrc/wrapper_kernels.cu), ONE plan (uv/w sort) serves every polarisation, and all gridding /
degridding runs in the CUDA library behind include/b200_wgridder.h.  There is no CPU fallback.

New optional kwargs (absent in the reference): ``precision`` ("double" default as the reference
forces fp64, ng.py:198/:203; "single" is the fp32 opt-in for epsilon >= 2e-6), ``nthreads``
(alias of ``threads``; both ignored on the GPU), ``shard`` / ``group`` / ``src`` (multi-GPU: "channel" or
"row" = ``bvis`` is this rank's block of the observation, "wplane" = every rank holds the observation
and works on its own |w| interval; one all-reduce of the un-normalised image + sumwt in invert, one
broadcast of the model image in predict; see ``_shard_mode``), ``sumwt_on_device`` (default None = automatic, see ``_invert``; False: sumwt is
computed with the reference's own numpy expression so that it is bit-exact).

The ``Visibility`` / ``Image`` arguments are used through the attributes the reference touches
(SURVEY.md section 8 a9); they may be the real ``ska_sdp_datamodels`` classes or any object
with the same attributes.  astropy is not required: the image centre and (l, m, n) come from
the SIN-projection formulas below, which equal ``pixel_to_skycoord`` / ``skycoord_to_lmn``
(imaging/base.py:68-70, util/coordinate_support.py:436-459) for the canonical images the
reference asserts (``is_canonical()``: RA---SIN, DEC--SIN, STOKES, FREQ).
"""
import ctypes
import logging

import numpy

from . import _lib, wgridder
from .polarisation import pol_matrix

log = logging.getLogger("func-python-logger")

__all__ = ["invert_ng", "predict_ng", "residual_ng", "invert_visibility", "predict_visibility"]

C_M_S = 299792458.0
_CHUNK_BYTES = 1 << 29  # raw Visibility bytes staged per prepare / finish launch


# ---- small host-side geometry (replaces astropy on this path) ------------------------------
def _is_image(model):
    try:
        from ska_sdp_datamodels.image.image_model import Image
        return isinstance(model, Image)
    except ImportError:
        return hasattr(model, "image_acc") and "pixels" in model


def _radec(pc):
    return float(pc.ra.rad), float(pc.dec.rad)


def image_centre_radec(im):
    """Sky position of the FFT centre pixel (nx//2, ny//2 zero-based): what
    pixel_to_skycoord(nx//2+1, ny//2+1, wcs, origin=1) returns (imaging/base.py:68-70)."""
    ny, nx = im["pixels"].data.shape[2], im["pixels"].data.shape[3]
    w = im.image_acc.wcs.wcs
    ra0, dec0 = numpy.radians(w.crval[0]), numpy.radians(w.crval[1])
    dx = (nx // 2 + 1) - w.crpix[0]
    dy = (ny // 2 + 1) - w.crpix[1]
    if dx == 0 and dy == 0:
        return float(ra0), float(dec0)
    # SIN (orthographic) deprojection of the intermediate world coordinates
    l = numpy.radians(w.cdelt[0] * dx)
    m = numpy.radians(w.cdelt[1] * dy)
    n = numpy.sqrt(max(0.0, 1.0 - l * l - m * m))
    dec = numpy.arcsin(m * numpy.cos(dec0) + n * numpy.sin(dec0))
    ra = ra0 + numpy.arctan2(l, n * numpy.cos(dec0) - m * numpy.sin(dec0))
    return float(ra), float(dec)


def separation(ra1, dec1, ra2, dec2):
    """Angular separation (Vincenty formula, as astropy's SkyCoord.separation)."""
    dl = ra2 - ra1
    num = numpy.hypot(numpy.cos(dec2) * numpy.sin(dl),
                      numpy.cos(dec1) * numpy.sin(dec2) - numpy.sin(dec1) * numpy.cos(dec2) * numpy.cos(dl))
    den = numpy.sin(dec1) * numpy.sin(dec2) + numpy.cos(dec1) * numpy.cos(dec2) * numpy.cos(dl)
    return float(numpy.arctan2(num, den))


def skycoord_to_lmn(ra, dec, ra0, dec0):
    """(l, m, n-1) of (ra, dec) relative to the phase centre (ra0, dec0); restates
    util/coordinate_support.py:436-459."""
    dra = ra - ra0
    l = numpy.cos(dec) * numpy.sin(dra)
    m = numpy.sin(dec) * numpy.cos(dec0) - numpy.cos(dec) * numpy.sin(dec0) * numpy.cos(dra)
    n = numpy.sin(dec) * numpy.sin(dec0) + numpy.cos(dec) * numpy.cos(dec0) * numpy.cos(dra)
    return float(l), float(m), float(n - 1.0)


def _shift_vector(vis, im):
    """None if shift_vis_to_image would be a no-op, else the direction (l, m, sqrt(1-l^2-m^2)-1)
    used by calculate_visibility_phasor (visibility/base.py:36-37).  Also returns whether the
    phasecentre attribute gets updated (imaging/base.py:71, :90)."""
    ra_i, dec_i = image_centre_radec(im)
    ra_v, dec_v = _radec(vis.phasecentre)
    if not separation(ra_v, dec_v, ra_i, dec_i) > 1e-15:
        return None, False
    l, m, nm1 = skycoord_to_lmn(ra_i, dec_i, ra_v, dec_v)
    if abs(nm1) < 1e-15:  # phaserotate_visibility: "no significant change" (visibility/base.py:78)
        return None, True
    return (l, m, float(numpy.sqrt(1.0 - l * l - m * m) - 1.0)), True


def _options(kwargs):
    return dict(
        epsilon=kwargs.get("epsilon", 1e-12),
        do_wstacking=kwargs.get("do_wstacking", True),
        verbosity=kwargs.get("verbosity", 0),
        precision=kwargs.get("precision", "double"),
        sumwt_on_device=kwargs.get("sumwt_on_device", None),
        shard=_shard_mode(kwargs.get("shard", None)),
        group=kwargs.get("group", None),
        src=kwargs.get("src", 0),
    )


def _shard_mode(shard):
    """Multi-GPU mode of one call (opt-in kwarg, never a positional; absent in the reference):
    None = this process alone; "channel" / "row" = ``bvis`` is this rank's own block of frequency
    channels / rows of the observation; "wplane" = every rank holds the whole observation and
    processes the samples of its own |w| interval (distributed.wplane_selection).  Needs an
    initialised ``torch.distributed`` process group (NCCL, one process per GPU)."""
    if shard in (None, False, "none"):
        return None
    if shard not in ("channel", "row", "wplane"):
        raise ValueError(f"unknown shard mode {shard!r} (channel, row, wplane or None)")
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        raise wgridder.WgridderError(
            f"shard={shard!r} needs an initialised torch.distributed process group")
    return shard


def _wplane_sel(D, freq, npix, pixsize, opt):
    """This rank's sample selection under w-plane sharding (identical thresholds on all ranks,
    computed without communication from the replicated uvw / freq)."""
    import torch.distributed as dist
    from . import distributed as b2dist
    world, rank = dist.get_world_size(opt["group"]), dist.get_rank(opt["group"])
    prec = "single" if wgridder._prec_code(opt["precision"]) == _lib.F32 else "double"
    return b2dist.wplane_selection(D.uvw, freq, world, rank, npix, pixsize, precision=prec)


def _pixsize(im):
    return float(numpy.abs(numpy.radians(im.image_acc.wcs.wcs.cdelt[0])))


def _vis_to_im(model, freq):
    return numpy.round(model.image_acc.wcs.sub([4]).wcs_world2pix(freq, 0)[0]).astype("int")


# ---- device plumbing --------------------------------------------------------------------------
def _torch():
    import torch
    if not torch.cuda.is_available():
        raise wgridder.WgridderError(
            "no CUDA device available: ska_sdp_func_python_b200 has no CPU path")
    return torch


_FLAG_CODES = {"uint8": _lib.FLAG_U8, "bool": _lib.FLAG_U8, "int32": _lib.FLAG_I32,
               "int64": _lib.FLAG_I64, "float64": _lib.FLAG_F64}


def _flags_for_device(flags):
    if flags is None:
        return None, _lib.FLAG_NONE
    name = flags.dtype.name
    if name not in _FLAG_CODES:
        flags = flags.astype(numpy.float64)
        name = "float64"
    if name == "bool":
        flags = flags.view(numpy.uint8)
    return flags, _FLAG_CODES[name]


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _dptr(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double)) if a is not None else None


class _Device:
    """Per-call device context: uvw (flipped for the gridder), f/c, stream."""

    def __init__(self, uvw_flipped, freq):
        torch = _torch()
        self.torch = torch
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.stream = torch.cuda.current_stream().cuda_stream
        self.uvw = torch.from_numpy(numpy.ascontiguousarray(uvw_flipped)).to(self.dev)
        self.fscale = torch.from_numpy(numpy.ascontiguousarray(freq, dtype=numpy.float64) / C_M_S).to(self.dev)

    def up(self, a):
        return self.torch.from_numpy(numpy.ascontiguousarray(a)).to(self.dev)

    def stager(self):
        from ._xfer import Stager
        return Stager(self.torch, self.dev)


def _rows_per_chunk(nchan, npol, bytes_per_elem):
    return max(1, int(_CHUNK_BYTES // max(1, nchan * npol * bytes_per_elem)))


def _prepare(D, vis, flags, weight, M, shift, dopsf, prec_code, npol_out):
    """Upload the Visibility cube in row chunks and run the prepare kernel.
    Returns planar device tensors ms [npol_out, nrow, nchan] complex, wgt (same, real), the
    [npol_out, nchan] mask of (pol, chan) planes holding a non-zero visibility (ng.py:238, :267)
    and the largest |imaging weight| if every one of them is an integer below 2^24 (the device
    sum of the flagged weights is then exact in any order and in either precision as long as
    count * max stays below 2^53), else None."""
    torch = D.torch
    nrow, nchan, npol_in = vis.shape
    cdt = torch.complex64 if prec_code == _lib.F32 else torch.complex128
    rdt = torch.float32 if prec_code == _lib.F32 else torch.float64
    ms = torch.empty((npol_out, nrow, nchan), device=D.dev, dtype=cdt)
    wg = torch.empty((npol_out, nrow, nchan), device=D.dev, dtype=rdt)
    if vis.dtype not in (numpy.complex64, numpy.complex128):
        vis = vis.astype(numpy.complex128)
    if weight.dtype not in (numpy.float32, numpy.float64):
        weight = weight.astype(numpy.float64)
    vcode = _lib.F32 if vis.dtype == numpy.complex64 else _lib.F64
    wcode = _lib.F32 if weight.dtype == numpy.float32 else _lib.F64
    flags, fcode = _flags_for_device(flags)
    Mh = numpy.ascontiguousarray(M, dtype=numpy.complex128)
    polmat = Mh.view(numpy.float64)
    lmn = None
    if shift is not None:
        # the device uvw is the flipped one (u -> -u, w -> -w): flip l and n-1 to compensate
        lmn = numpy.array([-shift[0], shift[1], -shift[2]], dtype=numpy.float64)
    L = _lib.lib()
    step = _rows_per_chunk(nchan, npol_in, vis.dtype.itemsize + weight.dtype.itemsize + 8)
    # staging through pinned buffers (_xfer.Stager); a single-precision plan takes its samples as
    # complex64 (converted during the staging copy: half the PCIe bytes); the weights travel in
    # their own type, so that the integrality test below sees the caller's values; integer flags
    # travel as uint8 when their values allow it and not at all when nothing is flagged
    stage = D.stager()
    vdt = None
    if prec_code == _lib.F32:
        vdt = torch.complex64
        vcode = _lib.F32
    fdt = None
    if flags is not None and flags.dtype.kind in "iub":
        lo, hi = (int(x) for x in torch.from_numpy(flags.reshape(-1)).aminmax())
        if lo == 0 and hi == 0:
            flags, fcode = None, _lib.FLAG_NONE
        elif flags.dtype.itemsize > 1 and lo >= 0 and hi <= 255:
            fdt, fcode = torch.uint8, _lib.FLAG_U8
    nonzero = torch.zeros((npol_out, nchan), device=D.dev, dtype=torch.bool)
    integral = torch.ones((), device=D.dev, dtype=torch.bool)
    wmax = torch.zeros((), device=D.dev, dtype=torch.float64)
    for r0 in range(0, nrow, step):
        r1 = min(nrow, r0 + step)
        dv = stage.put(vis[r0:r1], vdt)
        dw = stage.put(weight[r0:r1])
        df = stage.put(flags[r0:r1], fdt) if flags is not None else None
        _lib.check(L.b2g_vis_prepare(
            r1 - r0, nchan, npol_in, npol_out, _ptr(dv), vcode, _ptr(df), fcode, _ptr(dw), wcode,
            _dptr(polmat), _dptr(lmn), _ptr(D.uvw), _ptr(D.fscale), r0, nrow, int(bool(dopsf)),
            _ptr(ms), _ptr(wg), prec_code, ctypes.c_void_p(D.stream)))
        integral &= (dw == dw.round()).all()
        wmax = torch.maximum(wmax, dw.abs().amax().to(torch.float64)) if dw.numel() else wmax
        nonzero |= torch.view_as_real(ms[:, r0:r1]).ne(0).any(dim=3).any(dim=1)
        del dv, dw, df
    wmax = float(wmax.item())
    return ms, wg, nonzero.cpu().numpy(), (wmax if bool(integral.item()) and wmax < 2.0 ** 24 else None)


def _flagged_weight_host(bvis, nrow, vnchan, vnpol):
    """flagged_imaging_weight.astype('f8').reshape(nrow, nchan, npol) as the reference builds it
    (ng.py:203-204); used for the bit-exact sumwt."""
    acc = bvis.visibility_acc
    wgt = numpy.asarray(acc.flagged_imaging_weight).astype("f8")
    return wgt.reshape([nrow, vnchan, vnpol])


def _raw(bvis, name):
    try:
        return numpy.asarray(bvis[name].data)
    except (KeyError, AttributeError, TypeError):
        return None


def invert_ng(bvis, model, dopsf=False, normalise=True, **kwargs):
    """Invert using the B200 gridder.  Drop-in for ska_sdp_func_python.imaging.ng.invert_ng
    (ng.py:146-294): returns (Image like ``model`` holding the dirty image or PSF,
    sumwt[nchan, npol])."""
    return _invert(bvis, model, dopsf, normalise, False, kwargs)


def residual_ng(bvis, model, normalise=True, **kwargs):
    """Residual image of one major cycle: what
    ``invert_ng(subtract_visibility(bvis, predict_ng(bvis, model)), model)`` computes with the
    reference (ng.py:38-294, visibility/operations.py:86; the loop of
    sky_model/skymodel_imaging.py:70-235), in ONE pass per polarisation plane: the same plan
    (uv/w sort) serves the degridding of ``model`` and the gridding of the residual, the model is
    subtracted as it is degridded and the residual samples never leave the device
    (b2g_plan_residual).  Returns (residual Image, sumwt[nchan, npol])."""
    return _invert(bvis, model, False, normalise, True, kwargs)


def _invert(bvis, model, dopsf, normalise, subtract_model, kwargs):
    assert _is_image(model), model
    assert model.image_acc.is_canonical()
    opt = _options(kwargs)
    im = model.copy(deep=True)

    shift, _ = _shift_vector(bvis, im)
    freq = numpy.asarray(bvis.frequency.data)
    nrows, nbaselines, vnchan, vnpol = bvis.vis.shape
    nrow = nrows * nbaselines
    nchan, npol, ny, nx = im["pixels"].data.shape
    if nx != ny:
        raise ValueError("invert_ng: the reference supports square images only (npixdirty = nx)")
    pixsize = _pixsize(im)
    vis_to_im = _vis_to_im(model, freq)
    # (ng.py:228) -- a rank of a channel-sharded MFS inversion may own a single channel of the
    # band: it still images into the one MFS plane
    mfs = nchan == 1 and (vnchan > 1 or opt["shard"] == "channel")
    prec = wgridder._prec_code(opt["precision"])

    vis = _raw(bvis, "vis")
    flags = _raw(bvis, "flags")
    weight = _raw(bvis, "imaging_weight")
    if vis is None or weight is None:
        vis = numpy.asarray(bvis.visibility_acc.flagged_vis)
        weight = numpy.asarray(bvis.visibility_acc.flagged_imaging_weight)
        flags = None
    vis = vis.reshape(nrow, vnchan, vnpol)
    weight = weight.reshape(nrow, vnchan, vnpol)
    flags = None if flags is None else flags.reshape(nrow, vnchan, vnpol)

    uvw = numpy.array(bvis.uvw.data, dtype=numpy.float64).reshape(nrow, 3)
    uvw[:, 0] *= -1.0
    uvw[:, 2] *= -1.0

    M = pol_matrix(bvis.visibility_acc.polarisation_frame, im.image_acc.polarisation_frame)
    if M.shape != (npol, vnpol) and M.shape[0] < npol:
        raise ValueError("image has more polarisations than the converted visibilities")

    D = _Device(uvw, freq)
    torch = D.torch
    ms, wg, nonzero, wmax_int = _prepare(D, vis, flags, weight, M, shift, dopsf, prec, npol)
    mpix = None
    if subtract_model:
        assert npol == vnpol  # predict_ng's own requirement (ng.py:78)
        mpix = D.up(numpy.asarray(model["pixels"].data, dtype=numpy.float64))

    # ---- sum of weights: the reference's own numpy expression (bit-exact) ----------------
    sumwt = numpy.zeros([nchan, npol])
    on_device = opt["sumwt_on_device"]
    if on_device is None:
        # integer-valued weights (natural weighting: 0 / 1) below 2^24 are exact in either plan
        # precision and sum exactly in any order while the total stays below 2^53: the device sum
        # of the flagged weights then IS the reference's numpy sum, bit for bit, without a host
        # pass over the weight cube (the test runs on the caller's fp64 values, chunk by chunk)
        on_device = wmax_int is not None and wmax_int * nrow * vnchan < 2.0 ** 53
    if on_device:
        wsum = wg.sum(dim=1, dtype=torch.float64).cpu().numpy()  # [npol, vnchan]
        for pol in range(npol):
            if mfs:
                sumwt[0, pol] += wsum[pol].sum()
            else:
                for vchan in range(vnchan):
                    sumwt[vis_to_im[vchan], pol] += wsum[pol, vchan]
    else:
        wgtt = _flagged_weight_host(bvis, nrow, vnchan, vnpol).T
        for pol in range(npol):
            if mfs:
                sumwt[0, pol] += numpy.sum(wgtt[pol, :, :].T)
            else:
                for vchan in range(vnchan):
                    sumwt[vis_to_im[vchan], pol] += numpy.sum(wgtt[pol, vchan, :].T, axis=0)

    rdt = torch.float32 if prec == _lib.F32 else torch.float64
    pix = torch.zeros((nchan, npol, ny, nx), device=D.dev, dtype=torch.float64)
    if subtract_model:
        nonzero[...] = True  # a zero visibility still has a residual

    def run(plan, ichan, pol, m, w):
        if mpix is None:
            plan.invert(m, w, out=dirty)
        else:
            plan.residual(mpix[ichan, pol].t().contiguous().to(rdt), m, w, out=dirty)
    common = dict(epsilon=opt["epsilon"], do_wstacking=opt["do_wstacking"], precision=prec,
                  verbosity=opt["verbosity"], stream=D.stream)
    dirty = torch.empty((nx, ny), device=D.dev, dtype=rdt)
    if opt["shard"] == "wplane":
        # every rank holds the observation; it grids the samples of its own |w| interval only and
        # its plan spans only their w-planes (the per-plane FFT work divides with the samples)
        if not mfs:
            raise ValueError("shard='wplane' is for MFS imaging (image nchan == 1); shard a "
                             "spectral cube by 'channel'")
        sel = _wplane_sel(D, freq, nx, pixsize, opt)
        sumwt[...] = 0.0
        with wgridder.Plan(D.uvw, freq, nx, ny, pixsize, pixsize, sel=sel, **common) as plan:
            for pol in range(npol):
                w_sel = wg[pol].reshape(-1)[sel]
                sumwt[0, pol] = float(w_sel.sum(dtype=torch.float64).item())
                if nonzero[pol].any():
                    run(plan, 0, pol, ms[pol].reshape(-1)[sel], w_sel)
                    pix[0, pol] += dirty.t()
    elif mfs:
        with wgridder.Plan(D.uvw, freq, nx, ny, pixsize, pixsize, **common) as plan:
            for pol in range(npol):
                if nonzero[pol].any():
                    run(plan, 0, pol, ms[pol], wg[pol])
                    pix[0, pol] += dirty.t()
    else:
        # one gridding per (image channel, pol) over all visibility channels mapped to it
        # (algebraically the sum of the reference's per-channel ms2dirty calls, ng.py:260-289)
        for ichan in sorted(set(int(i) for i in vis_to_im)):
            if ichan < 0 or ichan >= nchan:
                raise IndexError(f"visibility channel maps to image channel {ichan} of {nchan}")
            sel = numpy.nonzero(vis_to_im == ichan)[0]
            if not nonzero[:, sel].any():
                continue
            tsel = torch.as_tensor(sel, device=D.dev)
            with wgridder.Plan(D.uvw, freq[sel], nx, ny, pixsize, pixsize, **common) as plan:
                for pol in range(npol):
                    if nonzero[pol, sel].any():
                        run(plan, ichan, pol, ms[pol].index_select(1, tsel).contiguous(),
                            wg[pol].index_select(1, tsel).contiguous())
                        pix[ichan, pol] += dirty.t()
    if opt["shard"] is not None:
        # the one exchange step of invert: un-normalised images and sumwt add over the ranks
        # (imaging_helpers.py:25-58 sum_invert_results), normalised once below
        from . import distributed as b2dist
        sw = torch.from_numpy(sumwt).to(D.dev)
        b2dist.allreduce_image_sumwt(pix, sw, opt["group"])
        sumwt = sw.cpu().numpy()
    if normalise:
        # normalise_sumwt, 2-D sumwt branch (imaging/base.py:117-128)
        for chan in range(nchan):
            for pol in range(npol):
                if sumwt[chan, pol] > 0.0:
                    pix[chan, pol] /= float(sumwt[chan, pol])
                else:
                    pix[chan, pol] = 0.0
    if kwargs.get("_device_result", False):
        # callers that keep accumulating on the device (skymodel_imaging.py): fp64 CUDA tensor
        return pix, sumwt
    dst = im["pixels"].data
    if (isinstance(dst, numpy.ndarray) and dst.flags.c_contiguous and dst.dtype == numpy.float64
            and dst.shape == tuple(pix.shape)):
        # through the pinned ring in 64 MB pieces: PCIe speed, and the copy into the image's own
        # (pageable) array runs while the next piece crosses (a plain .cpu() of an 8192^2 image
        # costs five times as much)
        flat_dev, flat_dst = pix.reshape(-1), dst.reshape(-1)
        stage = D.stager()
        step = 1 << 23
        for e0 in range(0, flat_dev.numel(), step):
            stage.get(flat_dev[e0:e0 + step], flat_dst[e0:e0 + step])
        stage.finish()
    else:
        dst[...] = pix.cpu().numpy()
    return im, sumwt


def predict_ng(bvis, model, **kwargs):
    """Predict using the B200 degridder.  Drop-in for
    ska_sdp_func_python.imaging.ng.predict_ng (ng.py:38-143)."""
    if model is None:
        return bvis
    assert _is_image(model), model
    assert model.image_acc.is_canonical()
    opt = _options(kwargs)
    # the result container (ng.py:63: a deep copy with a zeroed cube) is built by a helper thread
    # while the GPU degrids: copying the weights and flags is pure host-memory traffic
    import concurrent.futures
    pool = concurrent.futures.ThreadPoolExecutor(max_workers=1)
    newbvis_f = pool.submit(bvis.copy, deep=True, zero=True)
    pool.shutdown(wait=False)

    freq = numpy.asarray(bvis.frequency.data)
    nrows, nbaselines, vnchan, vnpol = bvis.vis.shape
    nrow = nrows * nbaselines
    uvw = numpy.array(bvis.uvw.data, dtype=numpy.float64).reshape(nrow, 3)
    if not numpy.isfinite(uvw).all():      # (nan_to_num itself is ten times the cost of the check)
        uvw = numpy.nan_to_num(uvw)
    m_nchan, m_npol, ny, nx = model["pixels"].data.shape
    assert m_npol == vnpol
    uvw[:, 0] *= -1.0
    uvw[:, 2] *= -1.0
    pixsize = _pixsize(model)
    vis_to_im = _vis_to_im(model, freq)
    mfs = m_nchan == 1
    prec = wgridder._prec_code(opt["precision"])

    D = _Device(uvw, freq)
    torch = D.torch
    cdt = torch.complex64 if prec == _lib.F32 else torch.complex128
    rdt = torch.float32 if prec == _lib.F32 else torch.float64
    pixels = D.up(numpy.asarray(model["pixels"].data, dtype=numpy.float64))
    if opt["shard"] is not None:
        # the one exchange step of predict: every rank degrids the same model image
        import torch.distributed as dist
        dist.broadcast(pixels, src=opt["src"], group=opt["group"])
    vist = torch.zeros((vnpol, nrow, vnchan), device=D.dev, dtype=cdt)
    common = dict(epsilon=opt["epsilon"], do_wstacking=opt["do_wstacking"], precision=prec,
                  verbosity=opt["verbosity"], stream=D.stream)
    if opt["shard"] == "wplane":
        # the rank predicts the samples of its own |w| interval; the others stay zero, so the
        # ranks' outputs are disjoint and add up to the full prediction (imaging_helpers.py:75-93)
        if not mfs:
            raise ValueError("shard='wplane' is for MFS imaging (image nchan == 1)")
        sel = _wplane_sel(D, freq, nx, pixsize, opt)
        with wgridder.Plan(D.uvw, freq, nx, ny, pixsize, pixsize, sel=sel, **common) as plan:
            for vpol in range(vnpol):
                vist[vpol].reshape(-1)[sel] = plan.predict(pixels[0, vpol].t().contiguous().to(rdt))
    elif mfs:
        with wgridder.Plan(D.uvw, freq, nx, ny, pixsize, pixsize, **common) as plan:
            for vpol in range(vnpol):
                plan.predict(pixels[0, vpol].t().contiguous().to(rdt), out=vist[vpol])
    else:
        for ichan in sorted(set(int(i) for i in vis_to_im)):
            if ichan < -m_nchan or ichan >= m_nchan:
                raise IndexError(f"visibility channel maps to image channel {ichan} of {m_nchan}")
            sel = numpy.nonzero(vis_to_im == ichan)[0]
            tsel = torch.as_tensor(sel, device=D.dev)
            with wgridder.Plan(D.uvw, freq[sel], nx, ny, pixsize, pixsize, **common) as plan:
                for vpol in range(vnpol):
                    out = plan.predict(pixels[ichan, vpol].t().contiguous().to(rdt))
                    vist[vpol].index_copy_(1, tsel, out)

    # image frame -> visibility frame, interleave, inverse phase rotation (ng.py:131-143)
    M = pol_matrix(model.image_acc.polarisation_frame, bvis.visibility_acc.polarisation_frame)
    # (the copy has bvis' phase centre; it is only collected once the samples are on the host, so
    # the helper thread's copying overlaps the download loop as well)
    shift, update_pc = _shift_vector(bvis, model)
    lmn = None
    if shift is not None:
        lmn = numpy.array([-shift[0], shift[1], -shift[2]], dtype=numpy.float64)
    Mh = numpy.ascontiguousarray(M, dtype=numpy.complex128)
    vis = numpy.empty((nrow, vnchan, vnpol), dtype=numpy.complex128)
    step = max(1, _rows_per_chunk(vnchan, vnpol, 16) // 2)
    L = _lib.lib()
    stage = D.stager()
    # a single-precision plan's samples cross PCIe as complex64 (half the bytes); the staging
    # ring widens them to the reference's complex128 on the way into the result cube
    odt, ocode = (torch.complex64, _lib.F32) if prec == _lib.F32 else (torch.complex128, _lib.F64)
    for r0 in range(0, nrow, step):
        r1 = min(nrow, r0 + step)
        chunk = torch.empty((r1 - r0, vnchan, vnpol), device=D.dev, dtype=odt)
        _lib.check(L.b2g_vis_finish(
            r1 - r0, vnchan, m_npol, vnpol, _ptr(vist), prec, _dptr(Mh.view(numpy.float64)),
            _dptr(lmn), _ptr(D.uvw), _ptr(D.fscale), r0, nrow, _ptr(chunk), ocode,
            ctypes.c_void_p(D.stream)))
        stage.get(chunk, vis[r0:r1])   # D2H into pinned memory while the next chunk is finished
    stage.finish()
    newbvis = newbvis_f.result()
    newbvis["vis"].data = vis.reshape([nrows, nbaselines, vnchan, vnpol])
    if update_pc:
        newbvis.attrs["phasecentre"] = model.image_acc.phasecentre
    return newbvis


def predict_visibility(vis, model, context="ng", gcfcf=None, **kwargs):
    """Dispatch as imaging/imaging.py:28-55 for the contexts this package implements."""
    if context == "2d":
        return predict_ng(vis, model, do_wstacking=False, **kwargs)
    if context == "ng":
        return predict_ng(vis, model, **kwargs)
    if context in ("awprojection", "wg"):
        raise ValueError(f"imaging context {context} is not provided by the B200 gridder; "
                         "use the reference implementation for it")
    raise ValueError(f"Unknown imaging context {context}")


def invert_visibility(vis, im, dopsf=False, normalise=True, context="ng", gcfcf=None, **kwargs):
    """Dispatch as imaging/imaging.py:58-105 for the contexts this package implements."""
    if context == "2d":
        return invert_ng(vis, im, dopsf=dopsf, normalise=normalise, do_wstacking=False, **kwargs)
    if context == "ng":
        return invert_ng(vis, im, dopsf=dopsf, normalise=normalise, **kwargs)
    if context in ("awprojection", "wg"):
        raise ValueError(f"imaging context {context} is not provided by the B200 gridder; "
                         "use the reference implementation for it")
    raise ValueError(f"Unknown imaging context {context}")
