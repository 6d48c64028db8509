#!/usr/bin/env python
"""Timing of the component-DFT kernel (row N3): CUDA events around b2g_dft_predict with inputs
resident in HBM, on the rows of a bench configuration; the numpy oracle (= dft_cpu_looped with
range reduction) on a row subsample as the CPU baseline.

    python tools/dft_perf.py --config C2 --ncomp 100 --npol 1 [--rows 400000] [--json out.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
C = 299792458.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2")
    ap.add_argument("--ncomp", type=int, default=100)
    ap.add_argument("--npol", type=int, default=1)
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--cpu-rows", type=int, default=2000)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    import torch
    from ska_sdp_func_python_b200 import simulate, dft
    obs = simulate.observation(args.config)
    uvw, freq = obs["uvw"], obs["freq"]
    if args.rows:
        uvw = uvw[:: max(1, uvw.shape[0] // args.rows)]
    nrow, nchan = uvw.shape[0], len(freq)
    rng = np.random.default_rng(1)
    fov = obs["npix"] * obs["cellsize"] / 2
    dc = rng.uniform(-fov, fov, (args.ncomp, 3))
    dc[:, 2] = np.sqrt(1 - dc[:, 0] ** 2 - dc[:, 1] ** 2) - 1
    fl = (rng.uniform(0.1, 1, (args.ncomp, nchan, args.npol)) + 0j)
    dev = torch.device("cuda")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    uvw_d, k_d, dc_d, fl_d = t(uvw), t(freq / C), t(dc), t(fl)
    out = torch.empty((nrow, nchan, args.npol), device=dev, dtype=torch.complex128)
    res = {"config": args.config, "nrow": nrow, "nchan": nchan, "ncomp": args.ncomp,
           "npol": args.npol, "sample_comps": nrow * nchan * args.ncomp}
    for prec in ("double", "single"):
        best = None
        for _ in range(args.reps + 1):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            dft.dft_predict_device(uvw_d, k_d, dc_d, fl_d, nchan, out=out, precision=prec)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        res[prec] = {"ms": best, "sample_comps_per_s": nrow * nchan * args.ncomp / (best * 1e-3),
                     "vis_per_s": nrow * nchan / (best * 1e-3)}
    from oracle import dft as odft
    k = min(args.cpu_rows, nrow)
    lam = np.einsum("rs,k->rks", uvw[:k], freq / C)
    t0 = time.perf_counter()
    ref = odft.dft_components(dc, lam, fl)
    dt = time.perf_counter() - t0
    res["cpu_oracle"] = {"rows": k, "seconds": dt, "sample_comps_per_s": k * nchan * args.ncomp / dt,
                         "kind": "port (numpy dft_cpu_looped restatement, 1 thread)"}
    chk = dft.dft_predict_device(uvw_d[:k].contiguous(), k_d, dc_d, fl_d, nchan).cpu().numpy()
    res["max_err_vs_oracle"] = float(np.abs(chk - ref).max() / np.abs(fl).sum(axis=0).max())
    line = json.dumps(res)
    print(line)
    if args.json:
        with open(args.json, "w") as fh:
            fh.write(line + "\n")


if __name__ == "__main__":
    main()
