#!/usr/bin/env python
"""Development timing on a GPU box: wall-clock split of the host-pointer path
(Plan(host uvw) / invert(host) / predict(host)) that bench.py reports as `e2e`."""
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
    ap.add_argument("--reps", type=int, default=4)
    ap.add_argument("--pageable", action="store_true")
    ap.add_argument("--verbosity", type=int, default=0)
    ap.add_argument("--chunk-log2", type=int, default=0,
                    help="samples per host-streaming chunk = 2^this (default: the library's 2^24)")
    a = ap.parse_args()
    if a.chunk_log2:
        from ska_sdp_func_python_b200 import _lib
        _lib.check(_lib.lib().b2g_set_host_chunk_samples(1 << a.chunk_log2))
    obs = simulate.observation(a.config, nchan=a.nchan)
    uvw = simulate.flip_for_gridder(obs["uvw"])
    freq = obs["freq"]
    nrow, nchan, npix, cell = uvw.shape[0], len(freq), obs["npix"], obs["cellsize"]
    prec = obs["precision"]
    cdt = torch.complex64 if prec == "single" else torch.complex128
    rdt = torch.float32 if prec == "single" else torch.float64
    pin = not a.pageable
    ms_h = torch.empty((nrow, nchan), dtype=cdt, pin_memory=pin)
    ms_h.copy_(torch.view_as_complex(torch.randn((nrow, nchan, 2), dtype=rdt)))
    wgt_h = torch.ones((nrow, nchan), dtype=rdt, pin_memory=pin)
    dirty_h = torch.empty((npix, npix), dtype=rdt, pin_memory=pin)
    vis_h = torch.empty((nrow, nchan), dtype=cdt, pin_memory=pin)
    ms, wgt, dirty, vis = ms_h.numpy(), wgt_h.numpy(), dirty_h.numpy(), vis_h.numpy()
    for r in range(a.reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        plan = wgridder.Plan(uvw, freq, npix, npix, cell, cell, obs["epsilon"], True, precision=prec,
                             verbosity=a.verbosity)
        t1 = time.perf_counter()
        plan.invert(ms, wgt, out=dirty)
        t2 = time.perf_counter()
        ti = plan.timings()
        plan.predict(dirty, out=vis)
        t3 = time.perf_counter()
        tp = plan.timings()
        info = plan.info()
        plan.close()
        t4 = time.perf_counter()
        print(f"rep {r}: plan {1e3 * (t1 - t0):.1f} (device {info['t_plan_ms']:.1f})  invert "
              f"{1e3 * (t2 - t1):.1f} (device total {ti['total']:.1f}, vis {ti['vis']:.1f})  predict "
              f"{1e3 * (t3 - t2):.1f} (device total {tp['total']:.1f}, vis {tp['vis']:.1f})  close "
              f"{1e3 * (t4 - t3):.1f}  sum {1e3 * (t4 - t0):.1f} ms")
    gb_in = (ms.nbytes + wgt.nbytes) / 1e9
    print(f"H2D {gb_in:.2f} GB, D2H {vis.nbytes / 1e9:.2f} GB")


if __name__ == "__main__":
    main()
