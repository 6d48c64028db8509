"""Direct-DFT oracle (fp64) for ms2dirty / dirty2ms -- TEST INFRASTRUCTURE ONLY.

Restates the measurement equation the reference evaluates through ducc0.wgridder
(call sites /root/reference/src/ska_sdp_func_python/imaging/ng.py:99-112, :240-256);
formula and provenance in oracle/dft_oracle.c.  Parity status: the sign/axis convention
is pinned to the reference's own dft_cpu_looped
(/root/reference/src/ska_sdp_func_python/imaging/dft.py:265-285) via tests/golden/;
it is unpinned against ducc0 output itself (ducc0 is not installable here).
"""
import concurrent.futures
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
C_M_S = 299792458.0


def build(force=False):
    """Compile liboracle.so with gcc (recipe: oracle/Makefile)."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("dft_oracle.c",)]
    if (not force and os.path.exists(so)
            and all(os.path.getmtime(so) >= os.path.getmtime(s) for s in srcs)):
        return so
    subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
    return _LIB


def _threaded(work, n, nthreads=None):
    """Run work(lo, hi) over [0, n) in chunks on a thread pool (ctypes drops the GIL)."""
    nthreads = nthreads or os.cpu_count() or 1
    if n <= 0:
        return
    nchunk = min(n, nthreads * 4)
    edges = [n * i // nchunk for i in range(nchunk + 1)]
    with concurrent.futures.ThreadPoolExecutor(max_workers=nthreads) as ex:
        list(ex.map(lambda i: work(edges[i], edges[i + 1]), range(nchunk)))


def _p(a, t=ctypes.c_double):
    return None if a is None else a.ctypes.data_as(ctypes.POINTER(t))


def ms2dirty_points(uvw, freq, ms, wgt, npix_x, npix_y, pixsize_x, pixsize_y,
                    do_wstacking, px, py):
    """dirty[px[k], py[k]] by direct summation (ducc0 array convention)."""
    uvw = np.ascontiguousarray(uvw, dtype=np.float64)
    freq = np.ascontiguousarray(freq, dtype=np.float64)
    ms = np.ascontiguousarray(ms, dtype=np.complex128)
    wgt = None if wgt is None else np.ascontiguousarray(wgt, dtype=np.float64)
    px = np.ascontiguousarray(px, dtype=np.int64)
    py = np.ascontiguousarray(py, dtype=np.int64)
    out = np.zeros(px.shape[0], dtype=np.float64)
    nrow, nchan = ms.shape
    L = lib()

    def work(lo, hi):
        L.oracle_ms2dirty_points(
            ctypes.c_int64(nrow), ctypes.c_int64(nchan), _p(uvw), _p(freq), _p(ms), _p(wgt),
            ctypes.c_int64(npix_x), ctypes.c_int64(npix_y), ctypes.c_double(pixsize_x),
            ctypes.c_double(pixsize_y), ctypes.c_int(int(do_wstacking)),
            ctypes.c_int64(hi - lo), _p(px[lo:hi], ctypes.c_int64), _p(py[lo:hi], ctypes.c_int64),
            _p(out[lo:hi]))

    _threaded(work, px.shape[0])
    return out


def ms2dirty(uvw, freq, ms, wgt, npix_x, npix_y, pixsize_x, pixsize_y, do_wstacking):
    """Full dirty image [npix_x, npix_y] by direct summation."""
    ix, iy = np.meshgrid(np.arange(npix_x), np.arange(npix_y), indexing="ij")
    out = ms2dirty_points(uvw, freq, ms, wgt, npix_x, npix_y, pixsize_x, pixsize_y,
                          do_wstacking, ix.ravel(), iy.ravel())
    return out.reshape(npix_x, npix_y)


def dirty2ms_samples(uvw, freq, dirty, wgt, pixsize_x, pixsize_y, do_wstacking, rows, chans):
    """ms[rows[k], chans[k]] by direct summation."""
    uvw = np.ascontiguousarray(uvw, dtype=np.float64)
    freq = np.ascontiguousarray(freq, dtype=np.float64)
    dirty = np.ascontiguousarray(dirty, dtype=np.float64)
    wgt = None if wgt is None else np.ascontiguousarray(wgt, dtype=np.float64)
    rows = np.ascontiguousarray(rows, dtype=np.int64)
    chans = np.ascontiguousarray(chans, dtype=np.int64)
    out = np.zeros(rows.shape[0], dtype=np.complex128)
    L = lib()

    def work(lo, hi):
        L.oracle_dirty2ms_samples(
            ctypes.c_int64(freq.shape[0]), _p(uvw), _p(freq), _p(dirty), _p(wgt),
            ctypes.c_int64(dirty.shape[0]), ctypes.c_int64(dirty.shape[1]),
            ctypes.c_double(pixsize_x), ctypes.c_double(pixsize_y),
            ctypes.c_int(int(do_wstacking)), ctypes.c_int64(hi - lo),
            _p(rows[lo:hi], ctypes.c_int64), _p(chans[lo:hi], ctypes.c_int64), _p(out[lo:hi]))

    _threaded(work, rows.shape[0])
    return out


def dirty2ms_samples_sparse(uvw, freq, px, py, flux, npix_x, npix_y, pixsize_x, pixsize_y,
                             do_wstacking, rows, chans):
    """ms[rows[k], chans[k]] of a model image that is zero except dirty[px[j], py[j]] += flux[j]
    (same sum as oracle_dirty2ms_samples restricted to the non-zero pixels; for images too large
    to scan per sample).  numpy, fp64, phase reduced to [-1/2, 1/2] turns."""
    uvw = np.asarray(uvw, dtype=np.float64)
    fl = np.asarray(freq, dtype=np.float64)[np.asarray(chans)] / C_M_S
    x = (np.asarray(px, dtype=np.float64) - npix_x // 2) * pixsize_x
    y = (np.asarray(py, dtype=np.float64) - npix_y // 2) * pixsize_y
    r2 = x * x + y * y
    if do_wstacking:
        nm1 = -r2 / (np.sqrt(1.0 - r2) + 1.0)
    else:
        nm1 = np.zeros_like(r2)
    n = nm1 + 1.0
    u, v, w = (uvw[np.asarray(rows), i] * fl for i in range(3))
    ph = u[:, None] * x[None, :] + v[:, None] * y[None, :] - w[:, None] * nm1[None, :]
    ph -= np.rint(ph)
    return np.sum((np.asarray(flux) / n)[None, :] * np.exp(-2j * np.pi * ph), axis=1)


def dirty2ms(uvw, freq, dirty, wgt, pixsize_x, pixsize_y, do_wstacking):
    """Full ms [nrow, nchan] by direct summation."""
    nrow, nchan = uvw.shape[0], np.asarray(freq).shape[0]
    r, c = np.meshgrid(np.arange(nrow), np.arange(nchan), indexing="ij")
    out = dirty2ms_samples(uvw, freq, dirty, wgt, pixsize_x, pixsize_y, do_wstacking,
                           r.ravel(), c.ravel())
    return out.reshape(nrow, nchan)


# ---- pure-numpy statement of the same sums (small cases; cross-checks the C) ----------

def ms2dirty_numpy(uvw, freq, ms, wgt, npix_x, npix_y, pixsize_x, pixsize_y, do_wstacking):
    x = (np.arange(npix_x) - npix_x // 2) * pixsize_x
    y = (np.arange(npix_y) - npix_y // 2) * pixsize_y
    xx, yy = np.meshgrid(x, y, indexing="ij")
    r2 = xx * xx + yy * yy
    if do_wstacking:
        nm1 = -r2 / (np.sqrt(1.0 - r2) + 1.0)
        n = nm1 + 1.0
    else:
        nm1 = np.zeros_like(r2)
        n = np.ones_like(r2)
    res = np.zeros((npix_x, npix_y))
    v = ms if wgt is None else ms * wgt
    for r in range(uvw.shape[0]):
        for c in range(len(freq)):
            if v[r, c] == 0:
                continue
            ph = freq[c] / C_M_S * (xx * uvw[r, 0] + yy * uvw[r, 1] - uvw[r, 2] * nm1)
            res += (v[r, c] * np.exp(2j * np.pi * ph)).real
    return res / n


def dirty2ms_numpy(uvw, freq, dirty, wgt, pixsize_x, pixsize_y, do_wstacking):
    npix_x, npix_y = dirty.shape
    x = (np.arange(npix_x) - npix_x // 2) * pixsize_x
    y = (np.arange(npix_y) - npix_y // 2) * pixsize_y
    xx, yy = np.meshgrid(x, y, indexing="ij")
    r2 = xx * xx + yy * yy
    if do_wstacking:
        nm1 = -r2 / (np.sqrt(1.0 - r2) + 1.0)
        n = nm1 + 1.0
    else:
        nm1 = np.zeros_like(r2)
        n = np.ones_like(r2)
    out = np.zeros((uvw.shape[0], len(freq)), dtype=np.complex128)
    dn = dirty / n
    for r in range(uvw.shape[0]):
        for c in range(len(freq)):
            ph = freq[c] / C_M_S * (xx * uvw[r, 0] + yy * uvw[r, 1] - uvw[r, 2] * nm1)
            out[r, c] = np.sum(dn * np.exp(-2j * np.pi * ph))
    return out if wgt is None else out * wgt


def dft_components(direction_cosines, uvw_lambda, vfluxes):
    """Component DFT, restating dft_cpu_looped
    (/root/reference/src/ska_sdp_func_python/imaging/dft.py:265-285) with the phase reduced to
    [-1/2, 1/2] turns before the exponential (exact in fp64; the reference loses ~|phase|*1e-16).
    direction_cosines [ncomp, 3], uvw_lambda [..., nchan, 3], vfluxes [ncomp, nchan or 1, npol]
    -> vis [..., nchan, npol] complex128.  Pinned by tests/golden/dft_comp_case.npz (outputs of
    the reference function itself)."""
    import numpy as np
    npol = vfluxes.shape[-1]
    vis = np.zeros(uvw_lambda.shape[:-1] + (npol,), dtype=complex)
    for icomp in range(direction_cosines.shape[0]):
        phase = np.sum(uvw_lambda * direction_cosines[icomp, :], axis=-1)
        phase = phase - np.rint(phase)
        phasor = np.exp(-2j * np.pi * phase)
        for pol in range(npol):
            vis[..., pol] += vfluxes[icomp, :, pol] * phasor
    return vis


def idft_components(direction_cosines, uvw_lambda, vis, weight, flags):
    """Adjoint of dft_components: the sums of idft_visibility_skycomponent
    (/root/reference/src/ska_sdp_func_python/imaging/dft.py:334-384).  uvw_lambda [nrow, nchan, 3],
    vis / weight / flags [nrow, nchan, npol] -> (flux [ncomp, nchan, npol] already divided by the
    weight sum where it is positive, wsum [nchan, npol]).  Pinned by tests/golden/idft_case.npz."""
    import numpy as np
    fw = weight * (1 - flags)
    fv = vis * (1 - flags)
    wsum = np.sum(fw, axis=0)
    out = []
    for dc in direction_cosines:
        phase = np.sum(uvw_lambda * dc, axis=-1)
        phasor = np.exp(2j * np.pi * (phase - np.rint(phase)))[..., None]
        f = np.sum(fw * fv * phasor, axis=0)
        f[wsum > 0.0] = f[wsum > 0.0] / wsum[wsum > 0.0]
        f[wsum <= 0.0] = 0.0
        out.append(f)
    return np.array(out), wsum
