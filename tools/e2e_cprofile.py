#!/usr/bin/env python
"""cProfile of the drop-in calls (invert_ng + predict_ng on host objects) for a bench configuration:
where the host side of `e2e` spends its time.

    python tools/e2e_cprofile.py --config C4 --nchan 256
"""
import argparse
import cProfile
import io
import os
import pstats
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C4")
    ap.add_argument("--nchan", type=int, default=256)
    a = ap.parse_args()
    import torch
    from ska_sdp_func_python_b200 import imaging, simulate
    from ska_sdp_func_python_b200 import minimal_models as mm
    obs = simulate.observation(a.config, nchan=a.nchan)
    nt, nbl = obs["ntimes"], obs["nbaselines"]
    freq, npix, cell = obs["freq"], obs["npix"], obs["cellsize"]
    nchan = len(freq)
    shape = (nt, nbl, nchan, 1)
    rng = np.random.default_rng(3)
    vis = np.empty(shape, dtype=np.complex128)
    one = np.empty((nbl, nchan, 1), dtype=np.complex128)
    one.real = rng.standard_normal(one.shape, dtype=np.float32)
    one.imag = rng.standard_normal(one.shape, dtype=np.float32)
    for t in range(nt):
        vis[t] = one
    weight = np.ones(shape)
    flags = np.zeros(shape, dtype=np.int64)
    pc = mm.SkyCoord.from_deg(15.0, -35.0)
    bvis = mm.Visibility(vis, obs["uvw"].reshape(nt, nbl, 3), freq, pc, mm.PolarisationFrame("stokesI"),
                         weight=weight, imaging_weight=weight, flags=flags)
    model = mm.create_image(npix, cell, pc, frequency=[float(freq.mean())], nchan=1,
                            channel_bandwidth=float(freq[-1] - freq[0]) or 1e6)
    kw = dict(epsilon=obs["epsilon"], do_wstacking=True, precision=obs["precision"])

    def step():
        dirty, sumwt = imaging.invert_ng(bvis, model, dopsf=False, normalise=True, **kw)
        out = imaging.predict_ng(bvis, dirty, **kw)
        return float(out["vis"].data[0, 1, 0, 0].real)

    step()
    torch.cuda.synchronize()
    for name, fn in (("invert_ng", lambda: imaging.invert_ng(bvis, model, normalise=True, **kw)),
                     ("predict_ng", lambda: imaging.predict_ng(bvis, model, **kw))):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        print(f"{name}: {time.perf_counter() - t0:.3f} s for {nt * nbl * nchan:.3e} samples")
    pr = cProfile.Profile()
    pr.enable()
    step()
    torch.cuda.synchronize()
    pr.disable()
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(28)
    print(s.getvalue()[:6000])


if __name__ == "__main__":
    main()
