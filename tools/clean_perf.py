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


if __name__ == "__main__":
    main()
