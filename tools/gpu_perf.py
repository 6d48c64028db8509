#!/usr/bin/env python
"""Development timing on a GPU box: per-stage device times of invert / predict for one of the
BASELINE configurations with device-resident inputs (not the judged bench; see bench.py)."""
import argparse
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ska_sdp_func_python_b200 import simulate, wgridder  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2")
    ap.add_argument("--nchan", type=int, default=None)
    ap.add_argument("--ntimes", type=int, default=None)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--algo", type=int, default=-1)
    ap.add_argument("--eps", type=float, default=None)
    ap.add_argument("--precision", default=None)
    ap.add_argument("--nowstack", action="store_true")
    ap.add_argument("--nu", type=int, default=0)
    ap.add_argument("--wscale", type=float, default=1.0)
    a = ap.parse_args()
    obs = simulate.observation(a.config, ntimes=a.ntimes, nchan=a.nchan)
    uvw = simulate.flip_for_gridder(obs["uvw"])
    uvw[:, 2] *= a.wscale
    freq = obs["freq"]
    nrow, nchan = uvw.shape[0], len(freq)
    prec = a.precision or obs["precision"]
    eps = a.eps or obs["epsilon"]
    dev = torch.device("cuda:0")
    cdt = torch.complex64 if prec == "single" else torch.complex128
    rdt = torch.float32 if prec == "single" else torch.float64
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    ms = torch.view_as_complex(torch.randn((nrow, nchan, 2), device=dev, dtype=rdt, generator=g))
    wgt = torch.ones((nrow, nchan), device=dev, dtype=rdt)
    wgt[torch.as_tensor(obs["auto"], device=dev)] = 0
    d_uvw = torch.as_tensor(uvw, device=dev)
    nvis = int((wgt != 0).sum().item())
    t0 = time.time()
    plan = wgridder.Plan(d_uvw, freq, obs["npix"], obs["npix"], obs["cellsize"], obs["cellsize"],
                         eps, not a.nowstack, precision=prec, algo=a.algo, verbosity=1,
                         nu=a.nu, nv=a.nu)
    torch.cuda.synchronize()
    print(f"plan: {time.time() - t0:.3f} s wall; info {plan.info()}")
    print(f"nrow={nrow} nchan={nchan} nvis(nonzero wgt)={nvis:.4g}")
    for r in range(a.reps):
        dirty = plan.invert(ms, wgt)
        ti = plan.timings()
        vis = plan.predict(dirty, None)
        tp = plan.timings()
        print(f"rep {r}: invert {ti['total']:.2f} ms ({nvis / ti['total'] / 1e6:.1f} Gvis/s)  "
              + " ".join(f"{k}={v:.2f}" for k, v in ti.items() if k != "total"))
        print(f"       predict {tp['total']:.2f} ms ({nrow * nchan / tp['total'] / 1e6:.1f} Gvis/s)  "
              + " ".join(f"{k}={v:.2f}" for k, v in tp.items() if k != "total"))
    print("dirty stats", float(dirty.abs().max()), "vis[:1000] stats", float(vis[:1000].abs().max()))


if __name__ == "__main__":
    main()
