#!/usr/bin/env python
"""Development check on a GPU box: every (precision, w-stacking, kernel family) against the
direct-DFT oracle, plus adjointness.  Prints one line per case; exits non-zero on failure.
(The pytest -m gpu suite is the judged parity suite; this is the quick table.)"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import dft as odft  # noqa: E402
from ska_sdp_func_python_b200 import wgridder  # noqa: E402

C = 299792458.0


def make_case(rng, nrow, nchan, npx, npy, psx, psy, wfrac):
    freq = np.linspace(100e6, 112e6, nchan) if nchan > 1 else np.array([100e6])
    lam_max = freq[-1] / C
    uvw = rng.uniform(-1, 1, (nrow, 3))
    uvw[:, 0] *= 0.48 / psx / lam_max
    uvw[:, 1] *= 0.48 / psy / lam_max
    uvw[:, 2] *= wfrac * 0.48 / max(psx, psy) / lam_max
    # a dense clump (many samples per cell) + exact repeats + zero-weight rows
    k = nrow // 4
    uvw[:k, :2] *= 0.02
    uvw[k:k + 10] = uvw[k]
    ms = rng.normal(size=(nrow, nchan)) + 1j * rng.normal(size=(nrow, nchan))
    wgt = rng.uniform(0.5, 1.5, (nrow, nchan))
    wgt[::17] = 0.0
    return uvw, freq, ms, wgt


def main():
    rng = np.random.default_rng(1805550721)
    fails = 0
    cases = []
    for prec, eps in (("double", 1e-12), ("double", 1e-7), ("single", 1e-5), ("single", 2e-3)):
        for ws in (False, True):
            for algo in (wgridder.ALGO_SIMPLE, wgridder.ALGO_WARPRUN):
                cases.append((prec, eps, ws, algo))
    geos = [(3000, 4, 256, 192, 2.0e-4, 2.5e-4, 0.05), (2000, 1, 128, 128, 2e-3, 2e-3, 0.3)]
    ref_cache = {}
    for gi, geo in enumerate(geos):
        nrow, nchan, npx, npy, psx, psy, wfrac = geo
        uvw, freq, ms, wgt = make_case(np.random.default_rng(100 + gi), *geo)
        model = np.zeros((npx, npy))
        model[rng.integers(0, npx, 40), rng.integers(0, npy, 40)] = rng.uniform(1, 2, 40)
        model[0, 0] = 1.0
        model[npx - 1, npy - 1] = -1.0
        for ws in (False, True):
            t0 = time.time()
            ref_cache[(gi, ws, "d")] = odft.ms2dirty(uvw, freq, ms, wgt, npx, npy, psx, psy, ws)
            ref_cache[(gi, ws, "v")] = odft.dirty2ms(uvw, freq, model, wgt, psx, psy, ws)
            print(f"# oracle geo{gi} ws={int(ws)}: {time.time() - t0:.1f} s", flush=True)
        keep = None
        for prec, eps, ws, algo in cases:
            if algo == wgridder.ALGO_SIMPLE:
                keep = None
            tag = f"geo{gi} {prec:6s} eps={eps:7.0e} ws={int(ws)} algo={algo}"
            try:
                plan = wgridder.Plan(uvw, freq, npx, npy, psx, psy, eps, ws, precision=prec,
                                     algo=algo)
            except wgridder.WgridderError as exc:
                if "do not support" in str(exc):
                    print(f"{tag}: n/a ({exc})")
                    continue
                print(f"{tag}: PLAN FAILED {exc}")
                fails += 1
                continue
            try:
                info = plan.info()
                cdt = np.complex128 if prec == "double" else np.complex64
                rdt = np.float64 if prec == "double" else np.float32
                d = plan.invert(ms.astype(cdt), wgt.astype(rdt))
                ref = ref_cache[(gi, ws, "d")]
                e_d = np.linalg.norm(d - ref) / np.linalg.norm(ref)
                m_d = np.abs(d - ref).max() / np.abs(ref).max()
                v = plan.predict(model.astype(rdt), wgt.astype(rdt))
                vref = ref_cache[(gi, ws, "v")]
                e_v = np.linalg.norm(v - vref) / np.linalg.norm(vref)
                m_v = np.abs(v - vref).max() / np.abs(vref).max()
                # adjointness <A x, y> = <x, A^H y>  (no weights)
                x = rng.normal(size=(npx, npy)).astype(rdt)
                y = (rng.normal(size=ms.shape) + 1j * rng.normal(size=ms.shape)).astype(cdt)
                Ax = plan.predict(x)
                Ahy = plan.invert(y)
                lhs = np.vdot(y.astype(np.complex128), Ax.astype(np.complex128)).real
                rhs = np.vdot(Ahy.astype(np.float64), x.astype(np.float64))
                adj = abs(lhs - rhs) / max(abs(lhs), abs(rhs))
                xs = ""
                if algo == wgridder.ALGO_SIMPLE:
                    keep = (d, v, Ax, Ahy)
                elif keep is not None:
                    xs = " x-algo " + " ".join(
                        f"{np.abs(a - b).max() / np.abs(b).max():.1e}"
                        for a, b in zip((d, v, Ax, Ahy), keep))
                ok = e_d < eps and e_v < eps and m_d < 10 * eps and m_v < 10 * eps
                adj_tol = 1e-11 if prec == "double" else 2e-5
                ok = ok and adj < adj_tol
                fails += 0 if ok else 1
                print(f"{tag}: W={info['W']} sig={info['sigma']:.2f} planes={info['nplanes']} "
                      f"L2 d={e_d:.1e} v={e_v:.1e} max d={m_d:.1e} v={m_v:.1e} adj={adj:.1e} "
                      f"{'ok' if ok else 'FAIL'}{xs}", flush=True)
            except Exception as exc:  # noqa: BLE001
                print(f"{tag}: EXEC FAILED {type(exc).__name__}: {exc}")
                fails += 1
            finally:
                plan.close()
    print("FAILS", fails)
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
