#!/usr/bin/env python
"""Gridding / degridding of the SPARSE part of a configuration (the samples above a |w| threshold:
long baselines, short runs of equal footprint origin) with each kernel family: warp-run (with and
without the tensor-core degridder) and the one-thread-per-sample kernels.

    python tools/sparse_perf.py --config C4 --nchan 256 --world 2 --rank 1
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ska_sdp_func_python_b200 import _lib, distributed as b2dist, simulate, wgridder  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C4")
    ap.add_argument("--nchan", type=int, default=256)
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--rank", type=int, default=1)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    obs = simulate.observation(a.config, nchan=a.nchan)
    uvw = torch.as_tensor(simulate.flip_for_gridder(obs["uvw"]), device=dev)
    freq, npix, cell, eps = obs["freq"], obs["npix"], obs["cellsize"], obs["epsilon"]
    thr = b2dist.wplane_thresholds(uvw, freq, a.world, npix, cell, precision="single")
    sel = b2dist.wplane_selection(uvw, freq, a.world, a.rank, npix, cell, precision="single", thresholds=thr)
    n = int(sel.numel())
    gen = torch.Generator(device=dev)
    gen.manual_seed(5)
    ms = torch.view_as_complex(torch.randn((n, 2), device=dev, dtype=torch.float32, generator=gen))
    model = torch.rand((npix, npix), device=dev, dtype=torch.float32, generator=gen)
    out = torch.empty((n,), device=dev, dtype=torch.complex64)
    dirty = torch.empty((npix, npix), device=dev, dtype=torch.float32)
    L = _lib.lib()
    line = {"config": a.config, "nchan": len(freq), "samples": n, "w_threshold": thr[a.rank]}
    for name, algo, tc in (("warprun", wgridder.ALGO_WARPRUN, 0), ("warprun_tc", wgridder.ALGO_WARPRUN, 1),
                           ("simple", wgridder.ALGO_SIMPLE, 0)):
        L.b2g_set_tc_mode(tc)
        with wgridder.Plan(uvw, freq, npix, npix, cell, cell, eps, True, precision="single", sel=sel,
                           algo=algo) as p:
            for _ in range(2):
                p.invert(ms, out=dirty)
                ti = p.timings()["vis"]
                p.predict(model, out=out)
                tp = p.timings()["vis"]
            line[name] = {"grid_ms": ti, "degrid_ms": tp, "nplanes": p.info()["nplanes"]}
    L.b2g_set_tc_mode(1)
    print(json.dumps(line))


if __name__ == "__main__":
    main()
