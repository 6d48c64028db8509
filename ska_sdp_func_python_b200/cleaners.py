"""Drop-in CLEAN minor cycles on a B200 (the consumers of the dirty
image and PSF made by ``invert_ng``), mirroring
/root/reference/src/ska_sdp_func_python/image/cleaners.py function by function -- same arguments,
same assertions, same return values:

  hogbom (:23-133), hogbom_complex (:136-232), msclean (:279-470), msmfsclean (:686-898) and the
  module's helpers (overlapIndices, argmax, create_scalestack, convolve_scalestack,
  convolve_convolve_scalestack, spheroidal_function),

plus the image-list wrappers that call them (image/deconvolution.py: hogbom_kernel_list,
complex_hogbom_kernel_list, msclean_kernel_list, mmclean_kernel_list, common_arguments).

The iteration loops run on the device (csrc/clean_kernels.cu) in the reference's own fp64
operations and order: Hogbom and complex Hogbom are bit-exact; the multi-scale variants are
bit-exact given the same scale-convolved stacks, which are formed here with cuFFT instead of
numpy's FFT (agreement ~1e-12 end to end).  No CPU fallback.
"""
import ctypes
import logging

import numpy

from . import _lib
from .imaging import _ptr, _torch

log = logging.getLogger("func-python-logger")

__all__ = ["hogbom", "hogbom_complex", "hogbom_kernel_list", "complex_hogbom_kernel_list", "msclean", "msclean_kernel_list", "msmfsclean", "mmclean_kernel_list", "create_scalestack",
           "spheroidal_function", "common_arguments", "image_peak", "overlapIndices", "argmax",
           "convolve_scalestack", "convolve_convolve_scalestack"]


def _dev64(a, dev):
    torch = _torch()
    if isinstance(a, torch.Tensor):
        return a.to(device=dev, dtype=torch.float64).contiguous()
    return torch.from_numpy(numpy.ascontiguousarray(numpy.asarray(a), dtype=numpy.float64)).to(dev)


def image_peak(img, window=None, mode="absmax"):
    """(value, (x, y)) of max |img * window| (first occurrence, numpy.argmax order) or, with
    ``mode="max"``, (max img, index)."""
    torch = _torch()
    dev = img.device if isinstance(img, torch.Tensor) and img.is_cuda else torch.device(
        "cuda", torch.cuda.current_device())
    d = _dev64(img, dev)
    w = None if window is None else _dev64(window, dev)
    val, idx = ctypes.c_double(0.0), ctypes.c_int64(0)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().b2g_image_peak(_ptr(d), _ptr(w), d.shape[0], d.shape[1],
                                         1 if mode == "max" else 0, ctypes.byref(val),
                                         ctypes.byref(idx), stream))
    return val.value, (int(idx.value // d.shape[1]), int(idx.value % d.shape[1]))


def hogbom(dirty, psf, window, gain, thresh, niter, fracthresh, prefix=""):
    """Clean the point spread function from a dirty image (Hogbom 1974), cleaners.py:23-133.

    ``dirty``, ``psf``, ``window`` (or None): 2-D numpy arrays or CUDA torch tensors.  Returns
    (comps, res) as numpy arrays -- as CUDA tensors if ``dirty`` was one."""
    assert 0.0 < gain < 2.0
    assert niter > 0
    torch = _torch()
    on_device = isinstance(dirty, torch.Tensor) and dirty.is_cuda
    dev = dirty.device if on_device else torch.device("cuda", torch.cuda.current_device())
    d = _dev64(dirty, dev)
    p = _dev64(psf, dev)
    w = None if window is None else _dev64(window, dev)
    if d.dim() != 2 or p.dim() != 2 or (w is not None and tuple(w.shape) != tuple(d.shape)):
        raise ValueError("hogbom: dirty, psf and window must be 2-D, window shaped like dirty")
    pmax, _ = image_peak(p, mode="max")
    numpy.testing.assert_approx_equal(pmax, 1.0, err_msg=f"PSF does not have unit peak {pmax}")
    comps = torch.empty_like(d)
    res = torch.empty_like(d)
    iters = ctypes.c_int64(0)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().b2g_hogbom(_ptr(d), _ptr(p), _ptr(w), d.shape[0], d.shape[1], p.shape[0],
                                     p.shape[1], float(gain), float(pmax), float(thresh),
                                     float(fracthresh), int(niter), _ptr(comps), _ptr(res),
                                     ctypes.byref(iters), stream))
    log.info("hogbom %s End of minor cycle after %d iterations", prefix, iters.value)
    if on_device:
        return comps, res
    return comps.cpu().numpy(), res.cpu().numpy()


def common_arguments(**kwargs):
    """fracthresh, gain, niter, thresh, scales from kwargs, with the reference's defaults and
    range checks (image/deconvolution.py:514-540)."""
    gain = kwargs.get("gain", 0.1)
    if gain <= 0.0 or gain >= 2.0:
        raise ValueError("Loop gain must be between 0 and 2")
    thresh = kwargs.get("threshold", 0.0)
    if thresh < 0.0:
        raise ValueError("Threshold must be positive or zero")
    niter = kwargs.get("niter", 100)
    if niter < 0:
        raise ValueError("niter must be greater than zero")
    fracthresh = kwargs.get("fractional_threshold", 0.01)
    if fracthresh < 0.0 or fracthresh > 1.0:
        raise ValueError("Fractional threshold should be in range 0.0, 1.0")
    scales = kwargs.get("scales", [0, 3, 10, 30])
    return fracthresh, gain, niter, thresh, scales


def _new_image_like(im, data):
    """Image of ``im``'s class / WCS / polarisation frame holding ``data`` (Image.constructor of the
    datamodels when available, image/deconvolution.py:629-638)."""
    if hasattr(type(im), "constructor"):
        return type(im).constructor(data=data, polarisation_frame=im.image_acc.polarisation_frame,
                                    wcs=im.image_acc.wcs)
    return type(im)(data, im.image_acc.wcs, im.image_acc.polarisation_frame)


def hogbom_kernel_list(dirty_list, prefix, psf_list, window_list, **kwargs):
    """Hogbom CLEAN of lists of single-frequency images, each polarisation separately
    (image/deconvolution.py:543-643): returns (component image list, residual image list).
    Polarisations whose PSF plane is all zero are skipped (left zero), as in the reference."""
    fracthresh, gain, niter, thresh, _ = common_arguments(**kwargs)
    comp_images, residual_images = [], []
    for channel, dirty in enumerate(dirty_list):
        psf = psf_list[channel]
        dpix = dirty["pixels"].data
        comp_array = numpy.zeros(dpix.shape)
        residual_array = numpy.zeros(dpix.shape)
        for pol in range(dpix.shape[1]):
            if psf["pixels"].data[0, pol, :, :].max():
                window = None
                if window_list is not None and window_list[channel] is not None:
                    window = window_list[channel]["pixels"].data[0, pol, :, :]
                comp_array[0, pol, :, :], residual_array[0, pol, :, :] = hogbom(
                    dpix[0, pol, :, :], psf["pixels"].data[0, pol, :, :], window, gain, thresh,
                    niter, fracthresh, prefix)
            else:
                log.info("hogbom_kernel_list %s: Skipping pol %d, channel %d", prefix, pol, channel)
        comp_images.append(_new_image_like(dirty, comp_array))
        residual_images.append(_new_image_like(dirty, residual_array))
    return comp_images, residual_images


# ---- multi-scale CLEAN (cleaners.py:279-470) ---------------------------------------------------

_SPH_NUM = ((8.203343e-2, -3.644705e-1, 6.278660e-1, -5.335581e-1, 2.312756e-1),
            (4.028559e-3, -3.697768e-2, 1.021332e-1, -1.201436e-1, 6.412774e-2))
_SPH_DEN = ((1.0, 8.212018e-1, 2.078043e-1), (1.0, 9.599102e-1, 2.918724e-1))


def spheroidal_function(vnu):
    """Prolate spheroidal wave function, m = 6, alpha = 1 (Schwab 1984), the reference's rational
    approximation term by term (cleaners.py:613-683).  Scalar in, float out."""
    if 0.0 <= vnu < 0.75:
        part, end = 0, 0.75
    elif 0.75 <= vnu <= 1.0:
        part, end = 1, 1.0
    else:
        return 0.0
    delta = vnu ** 2 - end ** 2
    num, den = _SPH_NUM[part][0], _SPH_DEN[part][0]
    for k in range(1, 5):
        num += _SPH_NUM[part][k] * delta ** k
    for k in range(1, 3):
        den += _SPH_DEN[part][k] * delta ** k
    return max(num / den, 0.0) if den != 0.0 else 0.0


def _scale_patch(width):
    """The (2 h + 2)^2 patch of one scale shape around the centre pixel, un-normalised
    (cleaners.py:489-501)."""
    half = int(numpy.ceil(width / 2.0))
    inv = 1.0 / (float(width) / 2.0) ** 2
    offs = range(-half - 1, half + 1)
    patch = numpy.zeros((len(offs), len(offs)))
    for iy, oy in enumerate(offs):
        for ix, ox in enumerate(offs):
            r = numpy.sqrt(inv * (float(ox) * float(ox) + float(oy) * float(oy)))
            patch[ix, iy] = spheroidal_function(r) * (1.0 - r ** 2)
    patch[patch < 0.0] = 0.0
    return half, patch


def create_scalestack(scaleshape, scales, norm=True, device=None):
    """Cube of the scale shapes (cleaners.py:473-508).  The few non-zero pixels of each scale are
    evaluated on the host (a (scale + 2)^2 table); the cube itself is built on ``device`` -- a numpy
    array comes back when ``device`` is None, as in the reference."""
    assert scaleshape[0] == len(scales)
    torch = _torch()
    dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
    nx, ny = int(scaleshape[1]), int(scaleshape[2])
    basis = torch.zeros((len(scales), nx, ny), dtype=torch.float64, device=dev)
    xcen, ycen = int(numpy.ceil(float(nx) / 2.0)), int(numpy.ceil(float(ny) / 2.0))
    for iscale, width in enumerate(scales):
        if width > 0.0:
            half, patch = _scale_patch(width)
            if norm:
                patch = patch / numpy.sum(patch)
            x0, y0 = xcen - half - 1, ycen - half - 1
            if x0 < 0 or y0 < 0 or x0 + patch.shape[0] > nx or y0 + patch.shape[1] > ny:
                raise ValueError(f"create_scalestack: scale {width} does not fit a {nx} x {ny} image")
            basis[iscale, x0:x0 + patch.shape[0], y0:y0 + patch.shape[1]] = torch.from_numpy(patch).to(dev)
        else:
            basis[iscale, xcen, ycen] = 1.0
    return basis.cpu().numpy() if device is None else basis


def _spectrum(torch, a):
    # the reference shifts on both sides of every transform; the frequency-side shifts are
    # permutations common to all factors of the element-wise products and drop out
    return torch.fft.rfft2(torch.fft.fftshift(a, dim=(-2, -1)))


def _back(torch, spec, shape):
    return torch.fft.ifftshift(torch.fft.irfft2(spec, s=shape), dim=(-2, -1))


def _convolve_scalestack(torch, fscales, img):
    """cleaners.py:511-531 on the device (cuFFT, real-to-complex)."""
    return _back(torch, _spectrum(torch, img)[None] * torch.conj(fscales), tuple(img.shape))


def msclean(dirty, psf, window, sensitivity, gain, thresh, niter, scales, fracthresh, prefix=""):
    """Multi-scale CLEAN (Cornwell 2008), cleaners.py:279-470: same arguments and return value
    (component image, residual image).  The scale-convolved stacks are formed on the device with
    cuFFT; the minor cycle runs in csrc/clean_kernels.cu (``b2g_msclean``).  The loop arithmetic is
    the reference's own; the stacks differ from numpy's FFT by rounding, so results agree to
    ~1e-12 rather than bit for bit."""
    assert 0.0 < gain < 2.0
    assert niter > 0
    assert len(scales) > 0
    torch = _torch()
    on_device = isinstance(dirty, torch.Tensor) and dirty.is_cuda
    dev = dirty.device if on_device else torch.device("cuda", torch.cuda.current_device())
    d = _dev64(dirty, dev)
    p = _dev64(psf, dev)
    if d.dim() != 2 or p.dim() != 2:
        raise ValueError("msclean: dirty and psf must be 2-D")
    pmax, psfpeak = image_peak(p, mode="max")
    assert pmax > 0.0
    log.info("msclean %s: Peak of PSF = %s at %s", prefix, pmax, psfpeak)
    ns = len(scales)
    lpsf, ldirty = p / pmax, d / pmax
    scalestack = create_scalestack([ns, d.shape[0], d.shape[1]], scales, norm=True, device=dev)
    pscalestack = create_scalestack([ns, p.shape[0], p.shape[1]], scales, norm=True, device=dev)
    fscales = _spectrum(torch, scalestack)
    res_scalestack = _convolve_scalestack(torch, fscales, ldirty).contiguous()
    fp = _spectrum(torch, pscalestack)
    fpsf = _spectrum(torch, lpsf)
    psf_scalescalestack = torch.empty((ns, ns) + tuple(p.shape), dtype=torch.float64, device=dev)
    for s in range(ns):
        psf_scalescalestack[s] = _back(torch, (fpsf[None] * fp) * torch.conj(fp[s])[None], tuple(p.shape))
    coupling = psf_scalescalestack.amax(dim=(2, 3)).cpu().numpy()
    log.info("msclean %s: Coupling matrix =\n %s", prefix, coupling)
    windowstack = None
    if window is not None:
        w = _dev64(window, dev)
        windowstack = (_convolve_scalestack(torch, fscales, w) > 0.9).to(torch.float64).contiguous()
        assert float(windowstack.sum()) > 0
    del fscales, fp, fpsf
    sens = None if sensitivity is None else _dev64(sensitivity, dev)
    maxabs, _ = image_peak(res_scalestack[0])
    absolutethresh = max(thresh, fracthresh * maxabs)
    log.info("msclean %s: Max abs in dirty Image = %.6f Jy/beam; this minor cycle will stop at %d "
             "iterations or peak < %.6f (Jy/beam)", prefix, maxabs, niter, absolutethresh)
    comps = torch.empty_like(d)
    iters = ctypes.c_int64(0)
    cdiag = numpy.ascontiguousarray(numpy.diagonal(coupling), dtype=numpy.float64)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().b2g_msclean(
        _ptr(res_scalestack), _ptr(psf_scalescalestack), _ptr(pscalestack), _ptr(windowstack),
        _ptr(sens), cdiag.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), ns, d.shape[0], d.shape[1],
        p.shape[0], p.shape[1], float(gain), float(absolutethresh), int(niter), _ptr(comps),
        ctypes.byref(iters), stream))
    log.info("msclean %s: End of minor cycle after %d components", prefix, iters.value)
    res = pmax * res_scalestack[0]
    if on_device:
        return comps, res
    return comps.cpu().numpy(), res.cpu().numpy()


def msclean_kernel_list(dirty_list, prefix, psf_list, window_list, sensitivity_list=None, **kwargs):
    """Multi-scale CLEAN of lists of single-frequency images, each polarisation separately
    (image/deconvolution.py:827-930): returns (component image list, residual image list)."""
    fracthresh, gain, niter, thresh, scales = common_arguments(**kwargs)
    comp_images, residual_images = [], []
    for channel, dirty in enumerate(dirty_list):
        psf = psf_list[channel]
        dpix = dirty["pixels"].data
        comp_array = numpy.zeros_like(dpix)
        residual_array = numpy.zeros_like(dpix)
        for pol in range(dpix.shape[1]):
            sens = None
            if sensitivity_list is not None and sensitivity_list[channel] is not None:
                sens = sensitivity_list[channel]["pixels"].data[0, pol, :, :]
            if psf["pixels"].data[0, pol, :, :].max():
                window = None
                if window_list is not None and window_list[channel] is not None:
                    window = window_list[channel]["pixels"].data[0, pol, :, :]
                comp_array[0, pol, :, :], residual_array[0, pol, :, :] = msclean(
                    dpix[0, pol, :, :], psf["pixels"].data[0, pol, :, :], window, sens, gain, thresh,
                    niter, scales, fracthresh, prefix)
            else:
                log.info("msclean_kernel_list %s: Skipping pol %d, channel %d", prefix, pol, channel)
        comp_images.append(_new_image_like(dirty, comp_array))
        residual_images.append(_new_image_like(dirty, residual_array))
    return comp_images, residual_images


# ---- multi-scale multi-frequency CLEAN (cleaners.py:686-898) ------------------------------------

def msmfsclean(dirty, psf, window, sensitivity, gain, thresh, niter, scales, fracthresh,
               findpeak="RASCIL", prefix=""):
    """Image-plane multi-scale multi-frequency CLEAN (Rau & Cornwell 2011), cleaners.py:686-898:
    ``dirty`` [nmoment, nx, ny], ``psf`` [2 nmoment, px, py]; returns (moment model, moment
    residual), both [nmoment, nx, ny].  Setup (scale-convolved moment residuals, coupled PSFs,
    moment Hessians) with cuFFT on the device, the nmoment x nmoment Hessian inverses on the host
    (numpy.linalg.inv, as the reference), the minor cycle in csrc/clean_kernels.cu
    (``b2g_msmfsclean``).  ``findpeak``: "CASA" searches the chi^2 decrease (cleaners.py:936-960);
    "RASCIL", "Algorithm1" and anything else the moment-0 principal solution (:925-972)."""
    assert 0.0 < gain < 2.0
    assert niter > 0
    assert len(scales) > 0
    torch = _torch()
    on_device = isinstance(dirty, torch.Tensor) and dirty.is_cuda
    dev = dirty.device if on_device else torch.device("cuda", torch.cuda.current_device())
    d = _dev64(dirty, dev)
    p = _dev64(psf, dev)
    if d.dim() != 3 or p.dim() != 3:
        raise ValueError("msmfsclean: dirty and psf must be [nmoment, nx, ny]")
    nmoment, ns = int(d.shape[0]), len(scales)
    if nmoment > 1:
        assert p.shape[0] == 2 * nmoment
    if nmoment > 4 or ns > 8:
        raise ValueError("msmfsclean: at most 4 frequency moments and 8 scales")
    pmax = float(p.max())
    assert pmax > 0.0
    lpsf, ldirty = p / pmax, d / pmax
    shape, pshape = tuple(d.shape[1:]), tuple(p.shape[1:])
    scalestack = create_scalestack([ns, shape[0], shape[1]], scales, norm=True, device=dev)
    pscalestack = create_scalestack([ns, pshape[0], pshape[1]], scales, norm=True, device=dev)
    fscales = _spectrum(torch, scalestack)
    # smresidual [ns, nmoment, nx, ny]                                          cleaners.py:1034-1052
    smres = torch.empty((ns, nmoment) + shape, dtype=torch.float64, device=dev)
    for t in range(nmoment):
        smres[:, t] = _convolve_scalestack(torch, fscales, ldirty[t])
    # coupled PSFs by t + q: [a, b, k] = psf[k] * scale_b (x) scale_a          cleaners.py:1055-1078
    nmp = max(int(p.shape[0]) // 2, 1)
    if nmp != nmoment:
        raise ValueError("msmfsclean: psf must hold 2 x nmoment planes")
    nk = 2 * nmp - 1
    fp = _spectrum(torch, pscalestack)
    psfm = torch.empty((ns, ns, nk) + pshape, dtype=torch.float64, device=dev)
    for k in range(nk):
        fk = _spectrum(torch, lpsf[k])
        for a in range(ns):
            psfm[a, :, k] = _back(torch, (fk[None] * fp) * torch.conj(fp[a])[None], pshape)
    centre = psfm[:, :, :, pshape[0] // 2, pshape[1] // 2].cpu().numpy()       # [a, b, k]
    hessian = numpy.zeros((ns, nmoment, nmoment))
    for s in range(ns):
        for t in range(nmoment):
            for q in range(nmoment):
                hessian[s, t, q] = centre[s, s, t + q]
    ihessian = numpy.ascontiguousarray(numpy.array([numpy.linalg.inv(h) for h in hessian]))
    for s in range(ns):
        log.info("mmclean %s: Moment-moment coupling matrix[scale %d] =\n %s", prefix, s, hessian[s])
    windowstack = None
    if window is not None:
        w = _dev64(window, dev)
        windowstack = (_convolve_scalestack(torch, fscales, w) > 0.9).to(torch.float64).contiguous()
    del fscales, fp
    sens = None if sensitivity is None else _dev64(sensitivity, dev)
    if sens is not None and sens.dim() == 3:
        # a moment cube of sensitivities (what mmclean_kernel_list passes) broadcasts against the
        # 2-D solution in the reference; max over moments of |s_t r| = |r| max_t |s_t| exactly
        sens = sens.abs().amax(dim=0).contiguous()
    if sens is not None and tuple(sens.shape) != shape:
        raise ValueError("msmfsclean: sensitivity must be shaped like one moment plane (or a cube of them)")
    maxabs, _ = image_peak(smres[0, 0])
    absolutethresh = max(thresh, fracthresh * maxabs)
    log.info("mmclean %s: Max abs in dirty Image = %.6f; this minor cycle will stop at %d iterations "
             "or peak < %.6f (Jy/beam)", prefix, maxabs, niter, absolutethresh)
    model = torch.empty((nmoment,) + shape, dtype=torch.float64, device=dev)
    iters = ctypes.c_int64(0)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().b2g_msmfsclean(
        _ptr(smres), _ptr(psfm), _ptr(pscalestack), _ptr(windowstack), _ptr(sens),
        ihessian.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
        numpy.ascontiguousarray(hessian).ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
        1 if findpeak == "CASA" else 0, ns, nmoment, shape[0], shape[1],
        pshape[0], pshape[1], float(gain), float(absolutethresh), int(niter), _ptr(model),
        ctypes.byref(iters), stream))
    log.info("mmclean %s: End of minor cycle after %d components", prefix, iters.value)
    res = pmax * smres[0]
    if on_device:
        return model, res
    return model.cpu().numpy(), res.cpu().numpy()


def mmclean_kernel_list(dirty_list, prefix, psf_list, window_list, sensitivity_list=None, **kwargs):
    """MSMFS CLEAN of lists of single-frequency images (image/deconvolution.py:646-824): frequency
    moments of the dirty images and PSFs (2 nmoment of the latter), ``msmfsclean`` per
    polarisation with the Stokes-I PSF moments, and channel images rebuilt from the moment
    model / residual.  kwargs: nmoment (3), findpeak, gain (0.7), threshold, niter,
    fractional_threshold, scales."""
    from .taylor_terms import (calculate_image_list_frequency_moments,
                               calculate_image_list_from_frequency_taylor_terms)
    findpeak = kwargs.get("findpeak", "RASCIL")
    nmoment = kwargs.get("nmoment", 3)
    if not nmoment >= 1:
        raise ValueError("Number of frequency moments must be greater than or equal to one")
    fracthresh, gain, niter, thresh, scales = common_arguments(**kwargs)
    gain = kwargs.get("gain", 0.7)
    if not 0.0 < gain < 2.0:
        raise ValueError("Loop gain must be between 0 and 2")
    nchan = len(dirty_list)
    if not nchan > 2 * (nmoment - 1):
        raise ValueError(f"Requires `nchan > 2 * (nmoment - 1)` ({nchan} > {2 * (nmoment - 1)})")
    dirty_taylor = calculate_image_list_frequency_moments(dirty_list, nmoment=nmoment)
    sensitivity_taylor = None
    if sensitivity_list is not None:
        sensitivity_taylor = calculate_image_list_frequency_moments(sensitivity_list, nmoment=nmoment)
        sensitivity_taylor["pixels"].data /= nchan
    window_taylor = None
    if window_list is not None:
        window_taylor = calculate_image_list_frequency_moments(window_list, nmoment=nmoment)
        window_taylor["pixels"].data /= nchan
    psf_taylor = calculate_image_list_frequency_moments(psf_list, nmoment=2 * nmoment if nmoment > 1 else 1)
    psf_peak = numpy.max(psf_taylor["pixels"].data)
    dirty_taylor["pixels"].data /= psf_peak
    psf_taylor["pixels"].data /= psf_peak
    dpix = dirty_taylor["pixels"].data
    comp_array = numpy.zeros(dpix.shape)
    residual_array = numpy.zeros(dpix.shape)
    for pol in range(dpix.shape[1]):
        sens = sensitivity_taylor["pixels"].data[:, pol, :, :] if sensitivity_taylor is not None else None
        if psf_taylor["pixels"].data[0, 0, :, :].max():
            window = window_taylor["pixels"].data[0, pol, :, :] if window_taylor is not None else None
            ppix = psf_taylor["pixels"].data[:, 0, :, :]
            comp_array[:, pol, :, :], residual_array[:, pol, :, :] = msmfsclean(
                dpix[:, pol, :, :], ppix, window, sens, gain, thresh, niter, scales, fracthresh,
                findpeak, prefix)
        else:
            log.info("deconvolve_cube %s: Skipping pol %d", prefix, pol)
    comp_taylor = _new_image_like(dirty_taylor, comp_array)
    residual_taylor = _new_image_like(dirty_taylor, residual_array)
    return (calculate_image_list_from_frequency_taylor_terms(dirty_list, comp_taylor),
            calculate_image_list_from_frequency_taylor_terms(dirty_list, residual_taylor))


# ---- complex Hogbom CLEAN of Q + iU (cleaners.py:136-232) ----------------------------------------

def hogbom_complex(dirty_q, dirty_u, psf_q, psf_u, window, gain, thresh, niter, fracthresh):
    """Complex Hogbom CLEAN of polarised data (Pratley & Johnston-Hollitt 2016), cleaners.py:136-232:
    same arguments and return value (comps.real, comps.imag, res.real, res.imag).  The loop runs on
    the device (``b2g_hogbom_complex``) in numpy's own complex arithmetic: bit-exact."""
    assert 0.0 < gain < 2.0
    assert niter > 0
    torch = _torch()
    on_device = isinstance(dirty_q, torch.Tensor) and dirty_q.is_cuda
    dev = dirty_q.device if on_device else torch.device("cuda", torch.cuda.current_device())
    dq, du = _dev64(dirty_q, dev), _dev64(dirty_u, dev)
    pq, pu = _dev64(psf_q, dev), _dev64(psf_u, dev)
    w = None if window is None else _dev64(window, dev)
    if dq.dim() != 2 or tuple(du.shape) != tuple(dq.shape) or pq.dim() != 2:
        raise ValueError("hogbom_complex: dirty_q / dirty_u must be 2-D images of one shape")
    assert tuple(pq.shape) == tuple(pu.shape) and bool(torch.equal(pq, pu))      # cleaners.py:190
    pmax, _ = image_peak(pq, mode="max")
    assert pmax > 0.0
    outs = [torch.empty_like(dq) for _ in range(4)]
    iters = ctypes.c_int64(0)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().b2g_hogbom_complex(
        _ptr(dq), _ptr(du), _ptr(pq), _ptr(w), dq.shape[0], dq.shape[1], pq.shape[0], pq.shape[1],
        float(gain), float(pmax), float(thresh), float(fracthresh), int(niter), _ptr(outs[0]),
        _ptr(outs[1]), _ptr(outs[2]), _ptr(outs[3]), ctypes.byref(iters), stream))
    log.info("hogbom: End of minor cycle after %d iterations", iters.value)
    if on_device:
        return tuple(outs)
    return tuple(o.cpu().numpy() for o in outs)


def complex_hogbom_kernel_list(dirty_list, psf_list, window_list, **kwargs):
    """Complex Hogbom CLEAN of lists of single-frequency stokesIQUV images
    (image/deconvolution.py:398-511): I and V by ``hogbom``, Q + iU by ``hogbom_complex``."""
    fracthresh, gain, niter, thresh, _ = common_arguments(**kwargs)
    comp_images, residual_images = [], []
    for channel, dirty in enumerate(dirty_list):
        psf = psf_list[channel]
        window = window_list[channel] if window_list is not None else None
        dpix, ppix = dirty["pixels"].data, psf["pixels"].data
        if dpix.shape[1] != 4:
            raise ValueError("complex_hogbom_kernel_list: needs stokesIQUV images")
        comp_array = numpy.zeros(dpix.shape)
        residual_array = numpy.zeros(dpix.shape)
        for pol in (0, 3):
            if ppix[0, pol, :, :].max():
                wd = window["pixels"].data[0, pol, :, :] if window is not None else None
                comp_array[0, pol], residual_array[0, pol] = hogbom(
                    dpix[0, pol, :, :], ppix[0, pol, :, :], wd, gain, thresh, niter, fracthresh)
            else:
                log.info("complex_hogbom_kernel_list: Skipping pol %d, channel %d", pol, channel)
        if ppix[0, 1:2, :, :].max():
            wd = window["pixels"].data[0, 1, :, :] if window is not None else None
            (comp_array[0, 1], comp_array[0, 2], residual_array[0, 1], residual_array[0, 2]) = hogbom_complex(
                dpix[0, 1, :, :], dpix[0, 2, :, :], ppix[0, 1, :, :], ppix[0, 2, :, :], wd, gain, thresh,
                niter, fracthresh)
        else:
            log.info("complex_hogbom_kernel_list: Skipping Q/U, channel %d", channel)
        comp_images.append(_new_image_like(dirty, comp_array))
        residual_images.append(_new_image_like(dirty, residual_array))
    return comp_images, residual_images


# ---- small public helpers of the reference module ---------------------------------------------

def overlapIndices(res, psf, peakx, peaky):
    """Index ranges where the PSF centred on (peakx, peaky) overlaps the image (cleaners.py:235-267):
    ((x0, x1, y0, y1) in the image, (x0, x1, y0, y1) in the PSF).  The kernels evaluate the same
    clipping on the device; this is the host form for callers that slice arrays themselves."""
    nx, ny = res.shape[0], res.shape[1]
    hx, hy = psf.shape[0] // 2, psf.shape[1] // 2
    x0, y0 = max(0, peakx - hx), max(0, peaky - hy)
    x1, y1 = min(nx, peakx + hx), min(peaky + hy, ny)
    return ((x0, x1, y0, y1),
            (max(0, hx + (x0 - peakx)), min(psf.shape[0], hx + (x1 - peakx)),
             max(0, hy + (y0 - peaky)), min(hy + (y1 - peaky), psf.shape[1])))


def argmax(a):
    """Unravelled index of the maximum (cleaners.py:270-276), on the device for CUDA tensors."""
    torch = _torch()
    if isinstance(a, torch.Tensor):
        return tuple(int(v) for v in numpy.unravel_index(int(a.reshape(-1).argmax()), tuple(a.shape)))
    return numpy.unravel_index(a.argmax(), a.shape)


def convolve_scalestack(scalestack, img):
    """img correlated with every plane of scalestack (cleaners.py:511-531), with cuFFT; numpy in,
    numpy out (CUDA tensors in, CUDA tensor out)."""
    torch = _torch()
    on_device = isinstance(img, torch.Tensor) and img.is_cuda
    dev = img.device if on_device else torch.device("cuda", torch.cuda.current_device())
    st, im = _dev64(scalestack, dev), _dev64(img, dev)
    out = _convolve_scalestack(torch, _spectrum(torch, st), im)
    return out if on_device else out.cpu().numpy()


def convolve_convolve_scalestack(scalestack, img):
    """[s, p] = img convolved with scale p and correlated with scale s (cleaners.py:534-562)."""
    torch = _torch()
    on_device = isinstance(img, torch.Tensor) and img.is_cuda
    dev = img.device if on_device else torch.device("cuda", torch.cuda.current_device())
    st, im = _dev64(scalestack, dev), _dev64(img, dev)
    fs, fi = _spectrum(torch, st), _spectrum(torch, im)
    out = torch.stack([_back(torch, (fi[None] * fs) * torch.conj(fs[s])[None], tuple(im.shape))
                       for s in range(st.shape[0])])
    return out if on_device else out.cpu().numpy()
