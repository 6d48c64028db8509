"""Run-length statistics of the sorted samples of C2 (equal footprint origin = one accumulation run of
the warp-run kernels): input to the next-round plan in DESIGN.md section 4.  Host only, ~1 min.
    python tools/probe/run_lengths.py"""
import sys, os, time, numpy as np
sys.path.insert(0, os.getcwd())
from ska_sdp_func_python_b200 import simulate
C = 299792458.0
obs = simulate.observation("C2")
uvw = simulate.flip_for_gridder(obs["uvw"]); freq = obs["freq"]
npix, cell = obs["npix"], obs["cellsize"]
W, nu, nv, nplanes, dw = 7, 3200, 3200, 9, 4774.05
fs = freq / C
w_all = np.abs(uvw[:, 2:3] * fs[None, :])
wmin, wmax = 0.0, float(w_all.max())
w0 = 0.5 * (wmin + wmax) - 0.5 * (nplanes - 1) * dw
t = time.time()
keys = np.empty(uvw.shape[0] * len(freq), dtype=np.int64)
nch = len(freq)
step = 200000
for r0 in range(0, uvw.shape[0], step):
    r1 = min(uvw.shape[0], r0 + step)
    u = uvw[r0:r1, 0:1] * fs[None, :]; v = uvw[r0:r1, 1:2] * fs[None, :]; w = uvw[r0:r1, 2:3] * fs[None, :]
    neg = w < 0
    u = np.where(neg, -u, u); v = np.where(neg, -v, v); w = np.abs(w)
    def origin(c, n):
        f = c - np.rint(c)
        s = f * n - 0.5 * W
        i0 = np.ceil(s).astype(np.int64)
        return np.where(i0 < 0, i0 + n, np.where(i0 >= n, i0 - n, i0))
    iu = origin(u * cell, nu); iv = origin(v * cell, nv)
    p0 = np.clip(np.ceil(w / dw - w0 / dw - 0.5 * W).astype(np.int64), 0, nplanes - W)
    keys[r0 * nch:r1 * nch] = ((p0 * nu + iu) * nv + iv).ravel()
print("keys", time.time() - t)
keys.sort()
print("sorted", time.time() - t)
b = np.flatnonzero(np.diff(keys)) + 1
runs = np.diff(np.concatenate([[0], b, [keys.size]]))
n = keys.size
print("samples", n, "runs", runs.size, "mean run", n / runs.size)
for k in (2, 8, 16, 64, 256):
    print(f"fraction of samples in runs >= {k}: {runs[runs >= k].sum() / n:.4f}")
# padding waste if every run is rounded up to a multiple of 8 samples (one K = 8 MMA step)
print("K=8 padding overhead:", (np.ceil(runs / 8) * 8).sum() / n - 1)
