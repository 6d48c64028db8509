#!/usr/bin/env python
"""Wall-clock of the data-model level drop-ins (invert_ng / predict_ng / weight_visibility /
residual_ng on Visibility / Image objects in host memory) for a bench configuration: what a user of
the reference's API sees, including every host copy and PCIe transfer.

    python tools/l2_perf.py --config C2 [--nchan 64] [--json out.json]
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
    ap.add_argument("--config", default="C2")
    ap.add_argument("--nchan", type=int, default=None)
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    import torch
    from ska_sdp_func_python_b200 import imaging, simulate, weighting
    from ska_sdp_func_python_b200 import minimal_models as mm
    obs = simulate.observation(a.config, nchan=a.nchan)
    nt, nbl = obs["ntimes"], obs["nbaselines"]
    freq, npix, cell = obs["freq"], obs["npix"], obs["cellsize"]
    nchan = len(freq)
    rng = np.random.default_rng(3)
    shape = (nt, nbl, nchan, 1)
    vis = np.empty(shape, dtype=np.complex128)
    vis.real = rng.standard_normal(shape, dtype=np.float32)
    vis.imag = rng.standard_normal(shape, dtype=np.float32)
    weight = np.ones(shape)
    weight[obs["auto"].reshape(nt, nbl)] = 0.0
    pc = mm.SkyCoord.from_deg(15.0, -35.0)
    bvis = mm.Visibility(vis, obs["uvw"].reshape(nt, nbl, 3), freq, pc, mm.PolarisationFrame("stokesI"),
                         weight=weight, flags=np.zeros(shape, dtype=np.int8))
    im = mm.create_image(npix, cell, pc, frequency=[freq[0]], nchan=1,
                         channel_bandwidth=2 * (freq[-1] - freq[0]) + 1e6)
    kw = dict(epsilon=obs["epsilon"], precision=obs["precision"])
    res = {"config": a.config, "samples": int(nt * nbl * nchan), "host_GB": {
        "vis": vis.nbytes / 1e9, "weight": weight.nbytes / 1e9}}

    def timed(name, fn, reps=2):
        best = None
        out = None
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = fn()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        res[name] = {"s": best, "vis_per_s": res["samples"] / best}
        return out

    timed("weight_visibility_uniform", lambda: weighting.weight_visibility(bvis, im, "uniform"))
    bvis["imaging_weight"].data[...] = weight
    dirty, sumwt = timed("invert_ng", lambda: imaging.invert_ng(bvis, im, **kw))
    model = im.copy(deep=True)
    model["pixels"].data[0, 0, rng.integers(0, npix, 100), rng.integers(0, npix, 100)] = 1.0
    timed("predict_ng", lambda: imaging.predict_ng(bvis, model, **kw))
    timed("residual_ng", lambda: imaging.residual_ng(bvis, model, **kw))
    print(json.dumps(res))
    if a.json:
        with open(a.json, "w") as fh:
            fh.write(json.dumps(res) + "\n")


if __name__ == "__main__":
    main()
