#!/usr/bin/env python
"""Timing of the Hogbom minor cycle (row N4): iterations per second on an npix^2 image with a
same-size PSF, device-resident, against the numpy oracle (= the reference's loop) on the host.

    python tools/clean_perf.py --npix 2048 --niter 2000 [--json out.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--npix", type=int, default=2048)
    ap.add_argument("--niter", type=int, default=2000)
    ap.add_argument("--cpu-niter", type=int, default=50)
    ap.add_argument("--json", default=None)
    ap.add_argument("--msmfs", action="store_true", help="time msmfsclean (3 moments, scales 0,3,10,30)")
    ap.add_argument("--msclean", action="store_true", help="time multi-scale CLEAN (scales 0,3,10,30) instead")
    a = ap.parse_args()
    import torch
    from ska_sdp_func_python_b200 import cleaners
    from oracle import clean as oc
    n = a.npix
    rng = np.random.default_rng(4)
    yy, xx = np.meshgrid(np.arange(n) - n // 2, np.arange(n) - n // 2)
    r = np.hypot(xx, yy)
    psf = np.exp(-(r / 2.0) ** 2) + 0.05 * np.cos(r / 3.1) * np.exp(-r / 60.0) * (r > 0)
    dirty = 0.01 * rng.normal(size=(n, n))
    dirty[rng.integers(0, n, 500), rng.integers(0, n, 500)] += rng.uniform(1, 5, 500)
    d, p = torch.from_numpy(dirty).cuda(), torch.from_numpy(psf).cuda()
    if a.msclean:
        return msclean_perf(a, cleaners, oc, torch, dirty, psf, d, p)
    if a.msmfs:
        return msmfs_perf(a, cleaners, oc, torch, dirty, psf)
    cleaners.hogbom(d, p, None, 0.1, 0.0, 64, 0.0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    comps, res = cleaners.hogbom(d, p, None, 0.1, 0.0, a.niter, 0.0)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t0 = time.perf_counter()
    oc_c, oc_r, _ = oc.hogbom(dirty, psf, None, 0.1, 0.0, a.cpu_niter, 0.0)
    dtc = time.perf_counter() - t0
    chk_c, chk_r = cleaners.hogbom(d, p, None, 0.1, 0.0, a.cpu_niter, 0.0)
    exact = bool(np.array_equal(chk_c.cpu().numpy(), oc_c) and np.array_equal(chk_r.cpu().numpy(), oc_r))
    # algorithmic bytes per iteration: residual read + written on the overlap, PSF read, all fp64
    byt = n * n * 8 * 3
    out = {"npix": n, "niter": a.niter, "gpu_us_per_iter": 1e6 * dt / a.niter,
           "gpu_alg_GBs": byt / (dt / a.niter) / 1e9,
           "cpu_us_per_iter": 1e6 * dtc / a.cpu_niter, "cpu_kind": "port (numpy, the reference's loop)",
           "bit_exact_vs_oracle": exact}
    print(json.dumps(out))
    if a.json:
        with open(a.json, "w") as fh:
            fh.write(json.dumps(out) + "\n")


def msclean_perf(a, cleaners, oc, torch, dirty, psf, d, p):
    """Multi-scale CLEAN: setup (cuFFT stacks) and minor cycle timed separately by differencing a
    short and a long run; CPU leg = the numpy oracle (the reference's loop and FFT setup)."""
    n = a.npix
    scales = [0, 3, 10, 30]
    run = lambda k: cleaners.msclean(d, p, None, None, 0.7, 0.0, k, scales, 0.0)
    run(8)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); run(8); torch.cuda.synchronize(); t_short = time.perf_counter() - t0
    t0 = time.perf_counter(); comps, res = run(a.niter); torch.cuda.synchronize(); t_long = time.perf_counter() - t0
    per_iter = (t_long - t_short) / (a.niter - 8)
    t0 = time.perf_counter()
    pmax, pstack, rs, pss, coupling, wstack = oc.msclean_setup(dirty, psf, None, scales)
    t_cpu_setup = time.perf_counter() - t0
    t0 = time.perf_counter()
    o_comps, done = oc.msclean_loop(rs, pss, pstack, coupling, None, None, psf.shape, 0.7, 0.0, a.cpu_niter)
    t_cpu_loop = time.perf_counter() - t0
    chk_c, chk_r = run(a.cpu_niter)
    err = float(max(np.abs(chk_c.cpu().numpy() - o_comps).max(), np.abs(chk_r.cpu().numpy() - pmax * rs[0]).max()))
    ns = len(scales)
    byt = n * n * 8 * 3 * ns      # every scale: residual read + written, coupled PSF read
    out = {"algorithm": "msclean", "scales": scales, "npix": n, "niter": a.niter,
           "gpu_setup_ms": 1e3 * (t_short - 8 * per_iter), "gpu_us_per_iter": 1e6 * per_iter,
           "gpu_alg_GBs": byt / per_iter / 1e9, "cpu_setup_ms": 1e3 * t_cpu_setup,
           "cpu_us_per_iter": 1e6 * t_cpu_loop / max(done, 1),
           "cpu_kind": "port (numpy, the reference's loop)", "max_abs_err_vs_oracle": err}
    print(json.dumps(out))
    if a.json:
        with open(a.json, "w") as fh:
            fh.write(json.dumps(out) + "\n")


def msmfs_perf(a, cleaners, oc, torch, dirty, psf):
    """MSMFS CLEAN, 3 moments x scales [0, 3, 10, 30]; CPU leg = the numpy oracle (the reference's
    setup and loop) on a bounded number of iterations."""
    n = a.npix
    scales = [0, 3, 10, 30]
    wc = np.linspace(-0.2, 0.2, 5)
    dm = np.array([dirty * np.mean(wc ** t * (1.0 + 0.5 * wc)) for t in range(3)])
    dm[1:] += 0.002 * np.random.default_rng(9).normal(size=(2, n, n))
    pm = np.array([psf * np.mean(wc ** k) for k in range(6)])
    d, p = torch.from_numpy(dm).cuda(), torch.from_numpy(pm).cuda()
    run = lambda k: cleaners.msmfsclean(d, p, None, None, 0.7, 0.0, k, scales, 0.0)
    run(8)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); run(8); torch.cuda.synchronize(); t_short = time.perf_counter() - t0
    t0 = time.perf_counter(); run(a.niter); torch.cuda.synchronize(); t_long = time.perf_counter() - t0
    per_iter = (t_long - t_short) / (a.niter - 8)
    out = {"algorithm": "msmfsclean", "scales": scales, "nmoment": 3, "npix": n, "niter": a.niter,
           "gpu_setup_ms": 1e3 * (t_short - 8 * per_iter), "gpu_us_per_iter": 1e6 * per_iter,
           # every (scale, moment) residual plane read + written, 5 coupled PSF moment planes per scale read
           "gpu_alg_GBs": n * n * 8 * (2 * 12 + 20) / per_iter / 1e9}
    if a.cpu_niter > 0:
        t0 = time.perf_counter()
        pmax, pstack, smres, ssmm, _, ihess, _ = oc.msmfs_setup(dm, pm, None, scales)
        out["cpu_setup_ms"] = 1e3 * (time.perf_counter() - t0)
        t0 = time.perf_counter()
        o_model, done = oc.msmfs_loop(smres, ssmm, pstack, ihess, None, None, 0.7, 0.0, a.cpu_niter)
        out["cpu_us_per_iter"] = 1e6 * (time.perf_counter() - t0) / max(done, 1)
        out["cpu_kind"] = "port (numpy, the reference's loop)"
        chk_m, chk_r = run(a.cpu_niter)
        out["max_abs_err_vs_oracle"] = float(max(np.abs(chk_m.cpu().numpy() - o_model).max(),
                                                 np.abs(chk_r.cpu().numpy() - pmax * smres[0]).max()))
    print(json.dumps(out))
    if a.json:
        with open(a.json, "w") as fh:
            fh.write(json.dumps(out) + "\n")


if __name__ == "__main__":
    main()
