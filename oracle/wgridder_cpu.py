"""CPU restatement of the w-stacking ES-kernel gridder -- TEST INFRASTRUCTURE / CPU BASELINE ONLY.

Driver for oracle/wgridder_cpu.c (see the header there for the algorithm and provenance:
the reference runs this stage inside ducc0.wgridder.ms2dirty / dirty2ms, call sites
/root/reference/src/ska_sdp_func_python/imaging/ng.py:99-112, :240-256; ducc0 0.27.0 is not in
/root/reference and not installable here, so this follows Arras et al. 2021, A&A 646 A58).
Always fp64, as the reference forces (ng.py:198, :203).  The 2-D FFTs go through scipy.fft
(pocketfft, the FFT family ducc0 itself uses) with ``workers=nthreads``.

Signatures mirror ducc0.wgridder: ``ms2dirty(uvw, freq, ms, wgt, npix_x, npix_y, pixsize_x,
pixsize_y, nu, nv, epsilon, do_wstacking, nthreads=1, ...)``.

Parity status: validated against the direct-DFT oracle (tests/test_oracle.py); unpinned
against ducc0 output itself.
"""
import ctypes
import json
import os
import subprocess

import numpy as np
import scipy.fft

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
C_M_S = 299792458.0
_DB = None


def build(force=False):
    so = os.path.join(_HERE, "libwgridder_cpu.so")
    src = os.path.join(_HERE, "wgridder_cpu.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "libwgridder_cpu.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build())
        i64, dbl, vp, ci = ctypes.c_int64, ctypes.c_double, ctypes.c_void_p, ctypes.c_int
        L.wg_create.restype = vp
        L.wg_create.argtypes = [i64, i64, vp, vp, i64, i64, dbl, dbl, ci, ci, vp, ci, ci, dbl,
                                dbl, ci]
        L.wg_destroy.argtypes = [vp]
        L.wg_destroy.restype = None
        L.wg_bucket.argtypes = [vp, vp, ci]
        L.wg_bucket.restype = ci
        L.wg_nkept.argtypes = [vp]
        L.wg_nkept.restype = i64
        L.wg_grid_plane.argtypes = [vp, ci, vp, vp, vp, ci]
        L.wg_grid_plane.restype = None
        L.wg_degrid_plane.argtypes = [vp, ci, vp, vp, ci]
        L.wg_degrid_plane.restype = None
        L.wg_finish_ms.argtypes = [vp, vp, vp, ci]
        L.wg_finish_ms.restype = None
        L.wg_screen_accumulate.argtypes = [vp, vp, i64, i64, dbl, vp, ci]
        L.wg_screen_accumulate.restype = None
        L.wg_screen_apply.argtypes = [vp, vp, i64, i64, dbl, vp, ci]
        L.wg_screen_apply.restype = None
        L.wg_correction.argtypes = [vp, i64, i64, vp, vp, vp, ci, dbl, vp, ci]
        L.wg_correction.restype = None
        _LIB = L
    return _LIB


# ---- kernel mathematics (numpy; independent of the product's es_kernel.h) ---------------

def es_phi(t, W, beta, e0):
    a = np.clip(1.0 - t * t, 0.0, None)
    return np.exp(beta * W * (a ** e0 - 1.0))


def poly_coefficients(W, beta, e0, deg=None):
    """[W][deg+1] monomial coefficients (highest power first) of the per-cell polynomial
    pieces; piece s covers t in [-1+2s/W, -1+2(s+1)/W], local variable xi in [-1,1]."""
    deg = deg if deg is not None else W + 3
    out = np.zeros((W, deg + 1))
    k = np.arange(deg + 1)
    xi = np.cos(np.pi * (k + 0.5) / (deg + 1))
    for s in range(W):
        t = -1.0 + (2.0 * s + 1.0 + xi) / W
        cheb = np.polynomial.chebyshev.chebfit(xi, es_phi(t, W, beta, e0), deg)
        mono = np.polynomial.chebyshev.cheb2poly(cheb)
        mono = np.pad(mono, (0, deg + 1 - len(mono)))
        out[s] = mono[::-1]
    return out


def kernel_ft(coef, x):
    """psi-hat at x (cycles per cell) of the piecewise-polynomial kernel (support W cells)."""
    W, n = coef.shape
    gx, gw = np.polynomial.legendre.leggauss(40)
    x = np.atleast_1d(np.asarray(x, dtype=np.float64))
    res = np.zeros_like(x)
    for s in range(W):
        vals = np.polyval(coef[s], gx)
        g = -0.5 * W + s + 0.5 * (1.0 + gx)
        res += 0.5 * (np.cos(2.0 * np.pi * np.outer(x, g)) @ (gw * vals))
    return res


def good_size_even(n):
    m = max(1, (int(n) + 1) // 2)
    while True:
        r = m
        for p in (2, 3, 5, 7):
            while r % p == 0:
                r //= p
        if r == 1:
            return 2 * m
        m += 1


def kernel_db():
    global _DB
    if _DB is None:
        path = os.path.join(os.path.dirname(_HERE), "ska_sdp_func_python_b200", "kernel_db.json")
        with open(path) as fh:
            _DB = json.load(fh)
    return _DB


def choose_kernel(epsilon, do_wstacking, sigma_max=2.0):
    """Smallest support W (then smallest oversampling) whose tabulated aliasing error meets
    epsilon, from the (W, sigma, beta, e0) table shared with the product (parameters only)."""
    eps_dim = epsilon / (1.5 * np.sqrt(3.0 if do_wstacking else 2.0))
    rows = [r for r in kernel_db() if r["eps"] <= eps_dim and 1.2 <= r["sigma"] <= sigma_max]
    if not rows:
        raise ValueError(f"no kernel reaches epsilon={epsilon}")
    rows.sort(key=lambda r: (r["W"], r["sigma"]))
    return rows[0]


class CpuPlan:
    def __init__(self, uvw, freq, npix_x, npix_y, pixsize_x, pixsize_y, epsilon, do_wstacking,
                 nthreads=None, kernel=None, nu=0, nv=0):
        self.L = lib()
        self.nthreads = int(nthreads or os.cpu_count() or 1)
        self.uvw = np.ascontiguousarray(uvw, dtype=np.float64)
        self.freq = np.ascontiguousarray(freq, dtype=np.float64)
        self.nrow, self.nchan = self.uvw.shape[0], self.freq.shape[0]
        self.npix_x, self.npix_y = int(npix_x), int(npix_y)
        self.psx, self.psy = float(pixsize_x), float(pixsize_y)
        self.wstack = bool(do_wstacking)
        k = kernel or choose_kernel(epsilon, do_wstacking)
        self.W, self.sigma, self.beta, self.e0 = int(k["W"]), k["sigma"], k["beta"], k["e0"]
        self.nu = int(nu) or max(good_size_even(np.ceil(self.sigma * npix_x)), 32)
        self.nv = int(nv) or max(good_size_even(np.ceil(self.sigma * npix_y)), 32)
        self.coef = np.ascontiguousarray(poly_coefficients(self.W, self.beta, self.e0))
        xm, ym = self.psx * (npix_x // 2), self.psy * (npix_y // 2)
        r2 = xm * xm + ym * ym
        self.nm1min = -1.0 if r2 >= 1 else -r2 / (np.sqrt(1 - r2) + 1)
        fs = self.freq / C_M_S
        if self.wstack and self.nrow:
            aw = np.abs(self.uvw[:, 2])
            wmin, wmax = aw.min() * fs.min(), aw.max() * fs.max()
            self.dw = 0.5 / (self.sigma * max(abs(self.nm1min), 1e-12))
            self.nplanes = int(np.ceil((wmax - wmin) / self.dw)) + self.W
            self.w0 = 0.5 * (wmin + wmax) - 0.5 * (self.nplanes - 1) * self.dw
        else:
            self.dw, self.nplanes, self.w0 = 1.0, 1, 0.0
        self.ctx = self.L.wg_create(
            self.nrow, self.nchan, self.uvw.ctypes.data, self.freq.ctypes.data, self.nu, self.nv,
            self.psx, self.psy, self.W, self.coef.shape[1] - 1, self.coef.ctypes.data,
            int(self.wstack), self.nplanes, self.w0, self.dw, self.nthreads)
        if not self.ctx:
            raise RuntimeError("wg_create failed")
        self._bucket_keep = "unset"
        self._corr = None

    def close(self):
        if self.ctx:
            self.L.wg_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _bucket(self, keep):
        kp = None if keep is None else np.ascontiguousarray(keep, dtype=np.uint8)
        rc = self.L.wg_bucket(self.ctx, None if kp is None else kp.ctypes.data, self.nthreads)
        if rc:
            raise RuntimeError(f"wg_bucket failed ({rc})")

    def correction(self):
        """1 / (psihat_u psihat_v psihat_w n) on the image [npix_x, npix_y]."""
        if self._corr is not None:
            return self._corr
        cu = 1.0 / kernel_ft(self.coef, (np.arange(self.npix_x) - self.npix_x // 2) / self.nu)
        cv = 1.0 / kernel_ft(self.coef, (np.arange(self.npix_y) - self.npix_y // 2) / self.nv)
        corr = np.empty((self.npix_x, self.npix_y))
        lo = min(self.nm1min, -1e-12)
        if self.wstack:
            fit = np.polynomial.chebyshev.Chebyshev.interpolate(
                lambda t: 1.0 / kernel_ft(self.coef, t * self.dw), 64, domain=[lo, 0.0])
            cheb = np.ascontiguousarray(fit.coef)
        else:
            cheb = np.ones(1)
        cu, cv = np.ascontiguousarray(cu), np.ascontiguousarray(cv)
        self.L.wg_correction(self.ctx, self.npix_x, self.npix_y, cu.ctypes.data, cv.ctypes.data,
                             cheb.ctypes.data, len(cheb), lo, corr.ctypes.data, self.nthreads)
        self._corr = corr
        return corr

    def invert(self, ms, wgt=None, planes=None):
        """``planes``: iterable of w-plane indices to process (default all).  A subset gives a
        partial image; bench.py uses it to time a bounded sample of a many-plane workload."""
        ms = np.ascontiguousarray(ms, dtype=np.complex128)
        wgt = None if wgt is None else np.ascontiguousarray(wgt, dtype=np.float64)
        keep = (ms != 0) if wgt is None else ((ms != 0) & (wgt != 0))
        self._bucket(keep)
        nt = self.nthreads
        grid = np.empty((self.nu, self.nv), dtype=np.complex128)
        acc = np.zeros((self.npix_x, self.npix_y))
        for p in (range(self.nplanes) if planes is None else planes):
            self.L.wg_grid_plane(self.ctx, p, ms.ctypes.data,
                                 None if wgt is None else wgt.ctypes.data, grid.ctypes.data, nt)
            f = scipy.fft.ifft2(grid, norm="forward", workers=nt, overwrite_x=True)
            self.L.wg_screen_accumulate(self.ctx, f.ctypes.data, self.npix_x, self.npix_y,
                                        self.w0 + p * self.dw, acc.ctypes.data, nt)
        return acc * self.correction()

    def predict(self, dirty, wgt=None, planes=None):
        dirty = np.ascontiguousarray(dirty, dtype=np.float64)
        wgt = None if wgt is None else np.ascontiguousarray(wgt, dtype=np.float64)
        self._bucket(None if wgt is None else (wgt != 0))
        nt = self.nthreads
        pre = np.ascontiguousarray(dirty * self.correction())
        grid = np.empty((self.nu, self.nv), dtype=np.complex128)
        out = np.zeros((self.nrow, self.nchan), dtype=np.complex128)
        for p in (range(self.nplanes) if planes is None else planes):
            self.L.wg_screen_apply(self.ctx, pre.ctypes.data, self.npix_x, self.npix_y,
                                   self.w0 + p * self.dw, grid.ctypes.data, nt)
            f = scipy.fft.fft2(grid, norm="backward", workers=nt, overwrite_x=True)
            self.L.wg_degrid_plane(self.ctx, p, f.ctypes.data, out.ctypes.data, nt)
        self.L.wg_finish_ms(self.ctx, None if wgt is None else wgt.ctypes.data, out.ctypes.data, nt)
        return out


def ms2dirty(uvw, freq, ms, wgt, npix_x, npix_y, pixsize_x, pixsize_y, nu, nv, epsilon,
             do_wstacking, nthreads=1, verbosity=0, mask=None,
             double_precision_accumulation=False, kernel=None):
    plan = CpuPlan(uvw, freq, npix_x, npix_y, pixsize_x, pixsize_y, epsilon, do_wstacking,
                   nthreads=nthreads, kernel=kernel, nu=nu, nv=nv)
    try:
        return plan.invert(ms, wgt)
    finally:
        plan.close()


def dirty2ms(uvw, freq, dirty, wgt, pixsize_x, pixsize_y, nu, nv, epsilon, do_wstacking,
             nthreads=1, verbosity=0, mask=None, kernel=None):
    plan = CpuPlan(uvw, freq, dirty.shape[0], dirty.shape[1], pixsize_x, pixsize_y, epsilon,
                   do_wstacking, nthreads=nthreads, kernel=kernel, nu=nu, nv=nv)
    try:
        return plan.predict(dirty, wgt)
    finally:
        plan.close()
