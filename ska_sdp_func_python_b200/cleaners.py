"""Drop-in Hogbom CLEAN on a B200 (the minor cycle that consumes the
dirty image and PSF made by ``invert_ng``).

``hogbom`` mirrors /root/reference/src/ska_sdp_func_python/image/cleaners.py:23-133 -- same
arguments, same assertions, same return value (component image, residual image) -- with the
iteration loop on the device (csrc/clean_kernels.cu): one pass over the L2-resident residual per
iteration subtracts the shifted PSF and finds the next peak.  All arithmetic is the reference's own
fp64 operations in the reference's order, so the results are bit-exact.  No CPU fallback.
"""
import ctypes
import logging

import numpy

from . import _lib
from .imaging import _ptr, _torch

log = logging.getLogger("func-python-logger")

__all__ = ["hogbom", "hogbom_kernel_list", "common_arguments", "image_peak"]


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
