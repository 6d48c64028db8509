#!/usr/bin/env python
"""Degridding on C2 (or another config) with the warp-run kernel and with the tensor-core kernel
(b2g_set_tc_mode 0 / 1 / 2): stage times from the plan's CUDA-event timers, and the error of each
against a double-precision plan of the same geometry on a sample subset.

    python tools/tc_perf.py --config C2 [--nchan 64] [--steps 3]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ska_sdp_func_python_b200 import _lib, simulate, wgridder  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2")
    ap.add_argument("--nchan", type=int, default=None)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--modes", default="0,1,2")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    obs = simulate.observation(a.config, nchan=a.nchan)
    uvw = torch.as_tensor(simulate.flip_for_gridder(obs["uvw"]), device=dev)
    freq, npix, cell, eps = obs["freq"], obs["npix"], obs["cellsize"], obs["epsilon"]
    nrow, nchan = uvw.shape[0], len(freq)
    gen = torch.Generator(device=dev)
    gen.manual_seed(7)
    model = torch.zeros((npix, npix), device=dev, dtype=torch.float32)
    ix = torch.randint(0, npix, (2000,), device=dev, generator=gen)
    iy = torch.randint(0, npix, (2000,), device=dev, generator=gen)
    model[ix, iy] = torch.rand(2000, device=dev, generator=gen) + 0.5
    L = _lib.lib()
    res, line = {}, {"config": a.config, "nrow": nrow, "nchan": nchan, "npix": npix, "epsilon": eps}
    for mode in [int(m) for m in a.modes.split(",")]:
        L.b2g_set_tc_mode(mode)
        with wgridder.Plan(uvw, freq, npix, npix, cell, cell, eps, True, precision="single") as p:
            out = torch.empty((nrow, nchan), device=dev, dtype=torch.complex64)
            for _ in range(2):
                p.predict(model, out=out)
            vis_ms = []
            for _ in range(a.steps):
                p.predict(model, out=out)
                vis_ms.append(p.timings()["vis"])
            info = p.info()
        res[mode] = out[:: max(1, nrow // 20000)].cpu().numpy().astype(np.complex128)
        line[f"mode{mode}_degrid_ms"] = float(np.median(vis_ms))
        line["W"], line["nplanes"] = info["W"], info["nplanes"]
    L.b2g_set_tc_mode(0)
    # double-precision plan with the SAME kernel parameters is not selectable; compare with an
    # fp64 plan at a 100x tighter epsilon on the row subset instead (kernel error << fp32 error)
    sub = slice(None, None, max(1, nrow // 20000))
    uvw_s = uvw[sub].contiguous()
    with wgridder.Plan(uvw_s, freq, npix, npix, cell, cell, eps / 100, True, precision="double") as p:
        ref = p.predict(model.double()).cpu().numpy()
    for mode, v in res.items():
        line[f"mode{mode}_rel_l2_vs_fp64"] = float(np.linalg.norm(v - ref) / np.linalg.norm(ref))
    print(json.dumps(line))


if __name__ == "__main__":
    main()
