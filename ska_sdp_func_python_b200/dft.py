"""Drop-in direct Fourier transform of sky components on a B200.

Mirrors /root/reference/src/ska_sdp_func_python/imaging/dft.py: ``dft_skycomponent_visibility``
(:32-56), ``extract_direction_and_flux`` (:59-118) and ``dft_kernel`` (:121-183, replacing its
``cpu_looped`` / ``gpu_cupy_raw`` / ``proc_func`` back ends with csrc/dft_kernels.cu).  Same names,
argument meaning and return types; there is no CPU fallback.
"""
import collections.abc
import ctypes

import numpy

from . import _lib
from .imaging import C_M_S, _flags_for_device, _ptr, _radec, _torch, skycoord_to_lmn
from .polarisation import pol_matrix

__all__ = ["dft_skycomponent_visibility", "extract_direction_and_flux", "dft_kernel",
           "dft_predict_device", "idft_visibility_skycomponent", "dft_invert_device"]


def dft_predict_device(uvw, kscale, direction_cosines, vfluxes, nchan, out=None, precision="double",
                       accumulate=False):
    """Array-level entry on CUDA tensors.  ``uvw`` [nrow, 3] metres with ``kscale`` [nchan] = f/c,
    or ``kscale=None`` and ``uvw`` [nrow*nchan, 3] in wavelengths; ``direction_cosines``
    [ncomp, 3]; ``vfluxes`` [ncomp, nchan or 1, npol] complex128.  Returns ``out``
    [nrow, nchan, npol] complex128 (CUDA tensor)."""
    torch = _torch()
    dev = uvw.device
    ncomp, fchan, npol = vfluxes.shape
    nrow = uvw.shape[0] if kscale is not None else uvw.shape[0] // nchan
    if out is None:
        out = torch.empty((nrow, nchan, npol), device=dev, dtype=torch.complex128)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().b2g_dft_predict(
        nrow, nchan, npol, ncomp, _ptr(uvw), _ptr(kscale), _ptr(direction_cosines), _ptr(vfluxes),
        fchan, _ptr(out), _lib.F32 if precision == "single" else _lib.F64, int(bool(accumulate)),
        stream))
    return out


def dft_kernel(direction_cosines, vfluxes, uvw_lambda, dft_compute_kernel=None):
    """DFT of components (dft.py:121-183).  ``dft_compute_kernel`` is accepted for compatibility
    (``None``, "cpu_looped", "gpu_cupy_raw", "proc_func" all run the B200 kernel; anything else
    raises ValueError as the reference does).

    :param direction_cosines: [ncomp, 3]
    :param vfluxes: [ncomp, nchan, npol] complex
    :param uvw_lambda: [ntimes, nbaselines, nchan, 3] wavelengths
    :return: vis [ntimes, nbaselines, nchan, npol] complex128 (numpy)
    """
    if dft_compute_kernel not in (None, "cpu_looped", "gpu_cupy_raw", "proc_func"):
        raise ValueError(f"dft_compute_kernel {dft_compute_kernel} not known")
    torch = _torch()
    dev = torch.device("cuda", torch.cuda.current_device())
    lam = numpy.ascontiguousarray(numpy.asarray(uvw_lambda), dtype=numpy.float64)
    ntimes, nbaselines, nchan, _ = lam.shape
    fl = numpy.ascontiguousarray(numpy.asarray(vfluxes), dtype=numpy.complex128)
    if fl.shape[1] not in (1, nchan):
        raise ValueError(f"flux has {fl.shape[1]} channels, visibility {nchan}")
    dc = numpy.ascontiguousarray(direction_cosines, dtype=numpy.float64)
    out = dft_predict_device(torch.from_numpy(lam.reshape(-1, 3)).to(dev), None,
                             torch.from_numpy(dc).to(dev), torch.from_numpy(fl).to(dev), nchan)
    return out.cpu().numpy().reshape(ntimes, nbaselines, nchan, fl.shape[2])


def extract_direction_and_flux(sc, vis):
    """Component direction cosines [ncomp, 3] and fluxes [ncomp, nchan, npol] scaled to the
    polarisation frame and channels of ``vis`` (dft.py:59-118)."""
    if not isinstance(sc, collections.abc.Iterable):
        sc = [sc]
    vpf = vis.visibility_acc.polarisation_frame
    vfreq = numpy.asarray(vis.frequency.data)
    ra0, dec0 = _radec(vis.phasecentre)
    vfluxes, direction_cosines = [], []
    for comp in sc:
        flux = numpy.asarray(comp.flux)
        if comp.polarisation_frame != vpf:
            M = pol_matrix(comp.polarisation_frame, vpf)
            flux = flux @ M.T
        cfreq = numpy.atleast_1d(numpy.asarray(comp.frequency))
        if len(cfreq) == len(vfreq) and numpy.allclose(cfreq, vfreq, rtol=1e-15):
            vflux = flux
        else:
            nchan, npol = flux.shape
            if nchan > 1:
                from scipy import interpolate
                vflux = numpy.zeros([len(vfreq), npol])
                for pol in range(npol):
                    fint = interpolate.interp1d(cfreq, numpy.asarray(comp.flux)[:, pol], kind="cubic")
                    vflux[:, pol] = fint(vfreq)
            else:
                vflux = flux
        vfluxes.append(vflux)
        ra, dec = _radec(comp.direction)
        l, m, _ = skycoord_to_lmn(ra, dec, ra0, dec0)
        direction_cosines.append(numpy.array([l, m, numpy.sqrt(1 - l ** 2 - m ** 2) - 1.0]))
    return numpy.array(direction_cosines), numpy.array(vfluxes).astype("complex")


def dft_skycomponent_visibility(vis, sc, dft_compute_kernel=None):
    """DFT to get the visibility from a SkyComponent (or list), written into ``vis`` (dft.py:32-56).
    Uses uvw [metres] x f/c on the device instead of materialising uvw_lambda."""
    if sc is None or (isinstance(sc, list) and len(sc) == 0):
        return vis
    if dft_compute_kernel not in (None, "cpu_looped", "gpu_cupy_raw", "proc_func"):
        raise ValueError(f"dft_compute_kernel {dft_compute_kernel} not known")
    direction_cosines, vfluxes = extract_direction_and_flux(sc, vis)
    torch = _torch()
    dev = torch.device("cuda", torch.cuda.current_device())
    shape = vis["vis"].data.shape
    nchan = shape[2]
    uvw = torch.from_numpy(numpy.ascontiguousarray(vis.uvw.data, dtype=numpy.float64).reshape(-1, 3)).to(dev)
    kscale = torch.from_numpy(numpy.asarray(vis.frequency.data, dtype=numpy.float64) / C_M_S).to(dev)
    out = dft_predict_device(uvw, kscale, torch.from_numpy(numpy.ascontiguousarray(direction_cosines)).to(dev),
                             torch.from_numpy(numpy.ascontiguousarray(vfluxes)).to(dev), nchan)
    vis["vis"].data = out.cpu().numpy().reshape(shape[0], shape[1], nchan, vfluxes.shape[2])
    return vis


def dft_invert_device(uvw, kscale, direction_cosines, vis, weight=None, flags=None, flag_code=_lib.FLAG_NONE):
    """Array-level adjoint on CUDA tensors: (flux [ncomp, nchan, npol] complex128, wsum
    [nchan, npol] float64) with flux = sum_rows weight (1-flags) vis (1-flags) conj(phasor)."""
    torch = _torch()
    nrow, nchan, npol = vis.shape
    ncomp = direction_cosines.shape[0]
    flux = torch.empty((ncomp, nchan, npol), device=vis.device, dtype=torch.complex128)
    wsum = torch.empty((nchan, npol), device=vis.device, dtype=torch.float64)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().b2g_dft_invert(nrow, nchan, npol, ncomp, _ptr(uvw), _ptr(kscale),
                                         _ptr(direction_cosines), _ptr(vis), _ptr(weight), _ptr(flags),
                                         flag_code, _ptr(flux), _ptr(wsum), stream))
    return flux, wsum


def idft_visibility_skycomponent(vis, sc):
    """Inverse DFT a SkyComponent (or list) from Visibility (dft.py:334-384): returns (list of
    components with the weighted-mean flux at their direction, list of weight arrays)."""
    if sc is None:
        return sc
    if not isinstance(sc, collections.abc.Iterable):
        sc = [sc]
    sc = list(sc)
    torch = _torch()
    dev = torch.device("cuda", torch.cuda.current_device())
    shape = vis["vis"].data.shape
    nchan, npol = shape[2], shape[3]
    ra0, dec0 = _radec(vis.phasecentre)
    dcs = []
    for comp in sc:
        ra, dec = _radec(comp.direction)
        l, m, _ = skycoord_to_lmn(ra, dec, ra0, dec0)
        dcs.append([l, m, numpy.sqrt(1 - l ** 2 - m ** 2) - 1.0])   # visibility/base.py:36-37
    up = lambda a, dt=None: torch.from_numpy(numpy.ascontiguousarray(a)).to(dev) if dt is None else \
        torch.from_numpy(numpy.ascontiguousarray(a)).to(dev).to(dt)  # noqa: E731
    fl, fcode = _flags_for_device(numpy.asarray(vis["flags"].data).reshape(-1, nchan, npol))
    flux, wsum = dft_invert_device(
        up(numpy.asarray(vis.uvw.data, dtype=numpy.float64).reshape(-1, 3)),
        up(numpy.asarray(vis.frequency.data, dtype=numpy.float64) / C_M_S),
        up(numpy.array(dcs, dtype=numpy.float64).reshape(-1, 3)),
        up(numpy.asarray(vis["vis"].data).reshape(-1, nchan, npol), torch.complex128),
        up(numpy.asarray(vis["weight"].data).reshape(-1, nchan, npol), torch.float64),
        up(fl), fcode)
    flux, weight = flux.cpu().numpy(), wsum.cpu().numpy()
    vpf = vis.visibility_acc.polarisation_frame
    newsc, weights_list = [], []
    for k, comp in enumerate(sc):
        newcomp = comp.copy()
        f = flux[k].copy()
        f[weight > 0.0] = f[weight > 0.0] / weight[weight > 0.0]
        f[weight <= 0.0] = 0.0
        if comp.polarisation_frame != vpf:
            f = f @ pol_matrix(vpf, comp.polarisation_frame).T
        newcomp.flux = f
        newsc.append(newcomp)
        weights_list.append(weight.copy())
    return newsc, weights_list
