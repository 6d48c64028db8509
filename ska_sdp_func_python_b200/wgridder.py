"""Array-level drop-in for ``ducc0.wgridder`` on a B200 (L1 boundary of SURVEY.md section 8b).

``ms2dirty`` / ``dirty2ms`` take the positional arguments exactly as the reference passes
them (/root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256 and :99-112), so this
module can be bound to the name ``ng`` in ``ska_sdp_func_python.imaging.ng`` (or injected as
``sys.modules["ducc0.wgridder"]``, the precedent being docs/src/conf.py:28-37 of the
reference).  All arithmetic happens in the CUDA library behind include/b200_wgridder.h; there
is no CPU path.

Inputs may be numpy arrays (staged through the C ABI, host pointers) or CUDA ``torch``
tensors (zero copy, device pointers; outputs are then CUDA tensors too).
"""
import ctypes

import numpy as np

from . import _lib
from ._lib import ALGO_AUTO, ALGO_SIMPLE, ALGO_WARPRUN, DEVICE, F32, F64, HOST, WgridderError

__all__ = ["Plan", "ms2dirty", "dirty2ms", "sumwt", "WgridderError", "ALGO_AUTO",
           "ALGO_SIMPLE", "ALGO_WARPRUN"]

STAGES = ("h2d", "prep", "vis", "fft", "screen", "correct", "d2h", "total")


def _is_torch(a):
    return type(a).__module__.startswith("torch")


def _torch():
    import torch
    return torch


def _prec_code(precision, ms_dtype=None):
    if precision is None:
        if ms_dtype is not None and np.dtype(ms_dtype) == np.complex64:
            return F32
        return F64
    if isinstance(precision, str):
        if precision in ("single", "f32", "float32"):
            return F32
        if precision in ("double", "f64", "float64"):
            return F64
    elif isinstance(precision, int) and not isinstance(precision, bool) and precision in (F32, F64):
        return precision
    raise ValueError(f"unknown precision {precision!r}")


class _Arr:
    """A view of a caller array as (pointer, dtype code, space) + keep-alive reference."""

    def __init__(self, a, kind, allow_none=False):
        self.keep = None
        self.ptr = None
        self.code = F64
        self.space = HOST
        self.is_torch = False
        if a is None:
            if not allow_none:
                raise ValueError("array must not be None")
            return
        if _is_torch(a):
            torch = _torch()
            if not a.is_cuda:
                a = a.detach().cpu().numpy()
            else:
                want = {"real": (torch.float32, torch.float64),
                        "complex": (torch.complex64, torch.complex128)}[kind]
                if a.dtype not in want:
                    a = a.to(want[1])
                a = a.contiguous()
                self.keep = a
                self.ptr = ctypes.c_void_p(a.data_ptr())
                self.code = F32 if a.dtype == want[0] else F64
                self.space = DEVICE
                self.is_torch = True
                self.shape = tuple(a.shape)
                return
        want = {"real": (np.float32, np.float64), "complex": (np.complex64, np.complex128)}[kind]
        a = np.asarray(a)
        if a.dtype not in want:
            a = a.astype(want[1])
        a = np.ascontiguousarray(a)
        self.keep = a
        self.ptr = ctypes.c_void_p(a.ctypes.data)
        self.code = F32 if a.dtype == want[0] else F64
        self.shape = a.shape


def _f64_host_or_device(a):
    """uvw / freq: float64, numpy (host) or CUDA tensor (device)."""
    if _is_torch(a) and a.is_cuda:
        torch = _torch()
        a = a.to(torch.float64).contiguous()
        return a, ctypes.c_void_p(a.data_ptr()), DEVICE
    if _is_torch(a):
        a = a.detach().cpu().numpy()
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    return a, ctypes.c_void_p(a.ctypes.data), HOST


class Plan:
    """Everything that depends only on (uvw, freq, image geometry, epsilon): kernel choice,
    w-plane layout, correction tables and the sorted visibility index.  Re-usable across
    polarisations and major cycles (the reference re-does this inside every ducc0 call,
    ng.py:235-289)."""

    def __init__(self, uvw, freq, npix_x, npix_y, pixsize_x, pixsize_y, epsilon,
                 do_wstacking=True, precision="double", nu=0, nv=0, algo=ALGO_AUTO,
                 verbosity=0, stream=None, sel=None):
        L = _lib.lib()
        self._L = L
        self._h = ctypes.c_void_p()
        uvw_a, uvw_p, space_u = _f64_host_or_device(uvw)
        freq_a, freq_p, space_f = _f64_host_or_device(freq)
        if space_u != space_f:
            # bring freq to where uvw lives (it is tiny)
            if space_u == DEVICE:
                torch = _torch()
                freq_a = torch.as_tensor(np.asarray(freq_a), device=uvw_a.device)
                freq_p = ctypes.c_void_p(freq_a.data_ptr())
            else:
                freq_a = freq_a.detach().cpu().numpy()
                freq_p = ctypes.c_void_p(freq_a.ctypes.data)
        if len(uvw_a.shape) != 2 or uvw_a.shape[1] != 3:
            raise ValueError("uvw must have shape [nrow, 3]")
        self.nrow = int(uvw_a.shape[0])
        self.nchan = int(freq_a.shape[0])
        self.npix_x, self.npix_y = int(npix_x), int(npix_y)
        self.precision = _prec_code(precision)
        # sel: optional sample selection (w-plane sharding); ms / wgt are then compact [nsel]
        sel_a, sel_p, self.nsel = None, ctypes.c_void_p(), None
        if sel is not None:
            if space_u == DEVICE:
                torch = _torch()
                sel_a = torch.as_tensor(sel, device=uvw_a.device).to(torch.int64)
                sel_a = sel_a.to(torch.uint32) if hasattr(torch, "uint32") else sel_a
                sel_a = sel_a.contiguous()
                sel_p = ctypes.c_void_p(sel_a.data_ptr())
            else:
                if _is_torch(sel):
                    sel = sel.detach().cpu().numpy()
                sel_a = np.ascontiguousarray(np.asarray(sel).astype(np.uint32))
                sel_p = ctypes.c_void_p(sel_a.ctypes.data)
            self.nsel = int(sel_a.shape[0])
        self.vis_shape = (self.nrow, self.nchan) if sel is None else (self.nsel,)
        self._keep = (uvw_a, freq_a, sel_a)  # device arrays are referenced, not copied
        self.stream = stream
        sp = ctypes.c_void_p(int(stream) if stream else 0)
        _lib.check(L.b2g_plan_create_sel(
            ctypes.byref(self._h), self.nrow, self.nchan, uvw_p, freq_p, space_u, sel_p,
            self.nsel or 0, self.npix_x, self.npix_y, float(pixsize_x), float(pixsize_y), int(nu), int(nv),
            float(epsilon), int(bool(do_wstacking)), self.precision, int(algo),
            int(verbosity), sp))

    # -- lifetime -------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._L.b2g_plan_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- execution ------------------------------------------------------------------
    def invert(self, ms, wgt=None, out=None, out_dtype=None):
        """ms2dirty on this plan.  Returns dirty [npix_x, npix_y]."""
        m = _Arr(ms, "complex")
        if tuple(m.shape) != self.vis_shape:
            raise ValueError(f"ms has shape {tuple(m.shape)}, expected {self.vis_shape}")
        w = _Arr(wgt, "real", allow_none=True)
        if w.ptr is not None:
            if tuple(w.shape) != self.vis_shape:
                raise ValueError("wgt shape does not match ms")
            if w.space != m.space or w.code != m.code:
                w = _Arr(_coerce_like(wgt, m), "real")
        if out_dtype is None:
            out_dtype = np.float64 if m.code == F64 else np.float32
        out_code = F64 if np.dtype(out_dtype) == np.float64 else F32
        if out is None:
            if m.is_torch:
                torch = _torch()
                out = torch.empty((self.npix_x, self.npix_y), device=m.keep.device,
                                  dtype=torch.float64 if out_code == F64 else torch.float32)
            else:
                out = np.empty((self.npix_x, self.npix_y), dtype=out_dtype)
        o = _Arr(out, "real")
        if o.keep is not out:
            raise ValueError("out must be a contiguous float32/float64 array")
        _lib.check(self._L.b2g_plan_invert(self._h, m.ptr, w.ptr, m.code, m.space, o.ptr,
                                           o.code, o.space))
        return out

    def predict(self, dirty, wgt=None, out=None, out_dtype=None):
        """dirty2ms on this plan.  Returns ms [nrow, nchan]."""
        d = _Arr(dirty, "real")
        if tuple(d.shape) != (self.npix_x, self.npix_y):
            raise ValueError(f"dirty has shape {tuple(d.shape)}, expected "
                             f"{(self.npix_x, self.npix_y)}")
        if out_dtype is None:
            out_dtype = np.complex128 if d.code == F64 else np.complex64
        out_code = F64 if np.dtype(out_dtype) == np.complex128 else F32
        if out is None:
            if d.is_torch:
                torch = _torch()
                out = torch.empty(self.vis_shape, device=d.keep.device,
                                  dtype=torch.complex128 if out_code == F64 else torch.complex64)
            else:
                out = np.empty(self.vis_shape, dtype=out_dtype)
        o = _Arr(out, "complex")
        if o.keep is not out:
            raise ValueError("out must be a contiguous complex64/complex128 array")
        w = _Arr(wgt, "real", allow_none=True)
        if w.ptr is not None and (w.space != o.space or w.code != o.code):
            w = _Arr(_coerce_like(wgt, o), "real")
        _lib.check(self._L.b2g_plan_predict(self._h, d.ptr, d.code, d.space, w.ptr, o.ptr,
                                            o.code, o.space))
        return out

    def residual(self, model, ms, wgt=None, out=None, ms_resid=None):
        """Residual image ms2dirty(ms - dirty2ms(model), wgt) in one call (b2g_plan_residual):
        the model is subtracted as it is degridded and the residual samples are gridded without
        leaving the device.  ``ms_resid`` (optional, same type / place as ``ms``; may be ``ms``
        itself) receives the residual visibilities.  Returns the residual image."""
        d = _Arr(model, "real")
        if tuple(d.shape) != (self.npix_x, self.npix_y):
            raise ValueError(f"model has shape {tuple(d.shape)}, expected "
                             f"{(self.npix_x, self.npix_y)}")
        m = _Arr(ms, "complex")
        if tuple(m.shape) != self.vis_shape:
            raise ValueError(f"ms has shape {tuple(m.shape)}, expected {self.vis_shape}")
        w = _Arr(wgt, "real", allow_none=True)
        if w.ptr is not None and (w.space != m.space or w.code != m.code):
            w = _Arr(_coerce_like(wgt, m), "real")
        r = _Arr(ms_resid, "complex", allow_none=True)
        if r.ptr is not None and (r.keep is not ms_resid or r.space != m.space or r.code != m.code):
            raise ValueError("ms_resid must be a contiguous array of the type and place of ms")
        if out is None:
            if d.is_torch:
                out = _torch().empty_like(d.keep)
            else:
                out = np.empty_like(d.keep)
        o = _Arr(out, "real")
        if o.keep is not out or o.space != d.space or o.code != d.code:
            raise ValueError("out must be a contiguous array of the type and place of model")
        _lib.check(self._L.b2g_plan_residual(self._h, d.ptr, d.code, d.space, m.ptr, w.ptr, m.code,
                                             m.space, r.ptr, o.ptr))
        return out

    # -- introspection --------------------------------------------------------------
    def info(self):
        ii = (ctypes.c_int64 * 12)()
        dd = (ctypes.c_double * 9)()
        _lib.check(self._L.b2g_plan_info(self._h, ii, 12, dd, 9))
        keys_i = ("W", "nu", "nv", "nplanes", "nslabs", "planes_per_slab", "algo",
                  "poly_degree", "nvis", "key_bits", "launches_last_call", "hand_fft")
        keys_d = ("sigma", "eps_kernel", "dw", "wmin", "wmax", "beta", "e0", "t_plan_ms",
                  "t_last_ms")
        res = {k: int(v) for k, v in zip(keys_i, ii)}
        res.update({k: float(v) for k, v in zip(keys_d, dd)})
        return res

    def timings(self):
        t = (ctypes.c_double * len(STAGES))()
        _lib.check(self._L.b2g_plan_timings(self._h, t, len(STAGES)))
        return dict(zip(STAGES, (float(x) for x in t)))


def _coerce_like(a, ref):
    """Bring real array `a` to the memory space / component type of _Arr `ref`."""
    if ref.is_torch:
        torch = _torch()
        dt = torch.float32 if ref.code == F32 else torch.float64
        return torch.as_tensor(np.asarray(a) if not _is_torch(a) else a,
                               device=ref.keep.device).to(dt)
    if _is_torch(a):
        a = a.detach().cpu().numpy()
    return np.asarray(a, dtype=np.float32 if ref.code == F32 else np.float64)


def ms2dirty(uvw, freq, ms, wgt, npix_x, npix_y, pixsize_x, pixsize_y, nu, nv, epsilon,
             do_wstacking, nthreads=1, verbosity=0, mask=None,
             double_precision_accumulation=False, *, precision=None, algo=ALGO_AUTO):
    """Drop-in for ``ducc0.wgridder.ms2dirty`` (positional order as at ng.py:240-256).

    ``nthreads`` and ``double_precision_accumulation`` are accepted and ignored (the GPU
    has no thread count; accumulation precision follows ``precision``).  ``precision``
    defaults to the component type of ``ms`` as in ducc0 (complex128 -> double)."""
    if mask is not None:
        m = np.asarray(mask) != 0
        wgt = m.astype(np.float64) if wgt is None else np.asarray(wgt) * m
    ms_dtype = ms.dtype if not _is_torch(ms) else (
        np.complex64 if str(ms.dtype) == "torch.complex64" else np.complex128)
    with Plan(uvw, freq, npix_x, npix_y, pixsize_x, pixsize_y, epsilon, do_wstacking,
              precision=_prec_code(precision, ms_dtype), nu=nu, nv=nv, algo=algo,
              verbosity=verbosity) as plan:
        return plan.invert(ms, wgt)


def dirty2ms(uvw, freq, dirty, wgt, pixsize_x, pixsize_y, nu, nv, epsilon, do_wstacking,
             nthreads=1, verbosity=0, mask=None, *, precision=None, algo=ALGO_AUTO):
    """Drop-in for ``ducc0.wgridder.dirty2ms`` (positional order as at ng.py:99-112)."""
    if mask is not None:
        m = np.asarray(mask) != 0
        wgt = m.astype(np.float64) if wgt is None else np.asarray(wgt) * m
    d_dtype = np.float32 if str(dirty.dtype) in ("float32", "torch.float32") else np.float64
    prec = _prec_code(precision) if precision is not None else (
        F32 if d_dtype == np.float32 else F64)
    shape = tuple(dirty.shape)
    with Plan(uvw, freq, shape[0], shape[1], pixsize_x, pixsize_y, epsilon, do_wstacking,
              precision=prec, nu=nu, nv=nv, algo=algo, verbosity=verbosity) as plan:
        return plan.predict(dirty, wgt)


def sumwt(wgt):
    """Deterministic fp64 sum of weights on the device (mirrors numpy.sum at ng.py:258)."""
    w = _Arr(wgt, "real")
    out = ctypes.c_double(0.0)
    n = int(np.prod(w.shape))
    _lib.check(_lib.lib().b2g_sumwt(w.ptr, w.code, w.space, n, ctypes.byref(out), None))
    return out.value
