#!/usr/bin/env python
"""Strong scaling of ONE observation across the GPUs of a box by w-plane sharding (SURVEY.md
section 8e, option 2; DESIGN.md "Multi-GPU"): every rank owns the samples of one |w| interval
(cost-balanced thresholds, distributed.wplane_thresholds), builds a plan over that selection only
(b2g_plan_create_sel: its own w-plane range, compact sample arrays) and the step is
    invert : local gridding of the rank's samples -> ONE all-reduce(sum) of the dirty image
    predict: local degridding of the (replicated) model image into the rank's own samples
so both the per-sample work and the per-plane FFT work divide by the number of ranks.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/wshard_bench.py --config C4 [--nchan 1024] [--steps 3] [--json out.json]
    python tools/wshard_bench.py --config C4            # N = 1: plain plan over all samples

Timed with CUDA events on the launching stream between barriers, max over ranks; value =
samples of the whole observation / time per (invert + predict) step.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ska_sdp_func_python_b200 import distributed as b2dist  # noqa: E402
from ska_sdp_func_python_b200 import simulate, wgridder  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C4")
    ap.add_argument("--nchan", type=int, default=None)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--json", default=None)
    ap.add_argument("--rebalance", type=int, default=1,
                    help="rounds of measured re-balancing of the |w| thresholds (N > 1)")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    obs = simulate.observation(a.config, nchan=a.nchan)
    uvw = simulate.flip_for_gridder(obs["uvw"])
    freq, npix, cell = obs["freq"], obs["npix"], obs["cellsize"]
    nrow, nchan = uvw.shape[0], len(freq)
    prec, eps = obs["precision"], obs["epsilon"]
    cdt = torch.complex64 if prec == "single" else torch.complex128
    rdt = torch.float32 if prec == "single" else torch.float64
    uvw_d = torch.as_tensor(uvw, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    t0 = time.time()
    sel = None
    if world > 1:
        thr = b2dist.wplane_thresholds(uvw_d, freq, world, npix, cell, precision=prec)
        sel = b2dist.wplane_selection(uvw_d, freq, world, rank, npix, cell, precision=prec,
                                      thresholds=thr)
        nloc = int(sel.numel())
        vis_shape = (nloc,)
    else:
        thr = [0.0, float("inf")]
        nloc = nrow * nchan
        vis_shape = (nrow, nchan)
    torch.cuda.synchronize()
    t_sel = time.time() - t0
    gen = torch.Generator(device=dev)
    gen.manual_seed(100 + rank)
    ms = torch.view_as_complex(torch.randn(vis_shape + (2,), device=dev, dtype=rdt, generator=gen))
    wgt = torch.ones(vis_shape, device=dev, dtype=rdt)
    t0 = time.time()
    plan = wgridder.Plan(uvw_d, freq, npix, npix, cell, cell, eps, True, precision=prec,
                         stream=stream, sel=sel)
    torch.cuda.synchronize()
    t_plan = time.time() - t0
    del sel   # the plan keeps its own uint32 copy
    torch.cuda.empty_cache()
    info = plan.info()
    dirty = torch.empty((npix, npix), device=dev, dtype=rdt)
    vis = torch.empty(vis_shape, device=dev, dtype=cdt)
    stage = {}

    def step(record):
        plan.invert(ms, wgt, out=dirty)
        ti = plan.timings()
        if world > 1:
            dist.all_reduce(dirty, op=dist.ReduceOp.SUM)
        plan.predict(dirty, out=vis)
        tp = plan.timings()
        if record:
            for k in ti:
                stage[k] = stage.get(k, 0.0) + ti[k] + tp[k]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(a.warmup):
        step(False)
    barrier()
    # measured rebalancing: one timed step under the a-priori thresholds, new thresholds from the
    # measured per-sample and per-plane costs, new selection + plan (amortised over the major
    # cycles that reuse the plan)
    t_rebalance = 0.0
    for _ in range(a.rebalance if world > 1 else 0):
        stage.clear()
        step(True)
        barrier()
        t0 = time.time()
        loc = {"n": nloc, "vis": stage["vis"], "planes": info["nplanes"],
               "img": stage["fft"] + stage["screen"] + stage["prep"]}
        allm = [None] * world
        dist.all_gather_object(allm, loc)
        plane_ms = sum(m["img"] for m in allm) / max(1, sum(m["planes"] for m in allm))
        thr = b2dist.wplane_rebalance(uvw_d, freq, world, npix, cell, thr,
                                      [m["n"] for m in allm], [m["vis"] for m in allm], plane_ms)
        plan.close()
        del ms, wgt, vis
        torch.cuda.empty_cache()
        sel = b2dist.wplane_selection(uvw_d, freq, world, rank, npix, cell, precision=prec,
                                      thresholds=thr)
        nloc = int(sel.numel())
        ms = torch.view_as_complex(torch.randn((nloc, 2), device=dev, dtype=rdt, generator=gen))
        wgt = torch.ones((nloc,), device=dev, dtype=rdt)
        vis = torch.empty((nloc,), device=dev, dtype=cdt)
        plan = wgridder.Plan(uvw_d, freq, npix, npix, cell, cell, eps, True, precision=prec,
                             stream=stream, sel=sel)
        del sel
        torch.cuda.empty_cache()
        info = plan.info()
        step(False)
        barrier()
        t_rebalance += time.time() - t0
    stage.clear()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step(True)
    e1.record()
    barrier()
    ms_step = e0.elapsed_time(e1) / a.steps
    peak_mem = torch.cuda.max_memory_allocated(dev) / 2 ** 30
    free, total = torch.cuda.mem_get_info(dev)
    mine = {"rank": rank, "nsamples": nloc, "ms_per_step": ms_step, "nplanes": info["nplanes"],
            "W": info["W"], "nu": info["nu"], "plan_s": t_plan, "select_s": t_sel, "rebalance_s": t_rebalance,
            "wlo": thr[rank], "whi": thr[rank + 1] if np.isfinite(thr[rank + 1]) else None,
            "hbm_used_GiB": (total - free) / 2 ** 30, "torch_peak_GiB": peak_mem,
            "stage_ms": {k: v / a.steps for k, v in stage.items()}}
    if world > 1:
        allr = [None] * world
        dist.all_gather_object(allr, mine)
    else:
        allr = [mine]
    if rank == 0:
        tmax = max(r["ms_per_step"] for r in allr)
        nsamp = nrow * nchan
        line = {"tool": "wshard_bench", "config": a.config, "nchan": nchan, "nrow": nrow,
                "npix": npix, "n_gpus": world, "samples": nsamp, "ms_per_step": tmax,
                "value_vis_per_s": nsamp / (tmax * 1e-3), "scaling": "strong", "precision": prec,
                "epsilon": eps, "ranks": allr}
        print(json.dumps(line), flush=True)
        if a.json:
            with open(a.json, "w") as fh:
                fh.write(json.dumps(line) + "\n")
    plan.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
