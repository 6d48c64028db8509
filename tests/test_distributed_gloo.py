"""World-size-2 check of the multi-GPU partitioning logic on CPU (gloo): channel shards, ONE
all-reduce of the un-normalised image + sumwt, one normalisation; broadcast for predict.  The
local gridding step is played by the CPU oracle here (test infrastructure); on the GPU box the
same functions wrap Plan.invert / Plan.predict (bench.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import make_case

from ska_sdp_func_python_b200 import distributed as b2dist

NPX, NPY, PSX, PSY = 32, 24, 2e-3, 2.5e-3


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, tmpdir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from oracle import dft as odft
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    uvw, freq, ms, wgt = make_case(21, 80, 4, NPX, NPY, PSX, PSY, 0.2)
    lo, hi = b2dist.shard_bounds(len(freq), world, rank)
    f, m, w = freq[lo:hi], np.ascontiguousarray(ms[:, lo:hi]), np.ascontiguousarray(wgt[:, lo:hi])

    def local_invert(ms_l, wgt_l):
        return torch.from_numpy(odft.ms2dirty(uvw, f, ms_l, wgt_l, NPX, NPY, PSX, PSY, True))

    dirty, sumwt = b2dist.invert_sharded(local_invert, m, w, w.sum(), normalise_result=True)

    model = torch.zeros((NPX, NPY), dtype=torch.float64)
    if rank == 0:
        model[5, 7], model[20, 3] = 1.5, -0.5   # only the source rank holds the model

    def local_predict(im):
        return odft.dirty2ms(uvw, f, im.numpy(), None, PSX, PSY, True)

    vis = b2dist.predict_sharded(local_predict, model, src=0)
    np.savez(os.path.join(tmpdir, f"rank{rank}.npz"), dirty=dirty.numpy(), sumwt=sumwt, vis=vis,
             lo=lo, hi=hi)
    dist.destroy_process_group()


def test_shard_bounds_cover_everything():
    for n in (1, 7, 64, 1000):
        for world in (1, 2, 3, 8):
            b = b2dist.channel_shards(n, world)
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(300)
def test_two_rank_invert_and_predict(tmp_path):
    from oracle import dft as odft
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    uvw, freq, ms, wgt = make_case(21, 80, 4, NPX, NPY, PSX, PSY, 0.2)
    ref = odft.ms2dirty(uvw, freq, ms, wgt, NPX, NPY, PSX, PSY, True) / wgt.sum()
    model = np.zeros((NPX, NPY))
    model[5, 7], model[20, 3] = 1.5, -0.5
    vref = odft.dirty2ms(uvw, freq, model, None, PSX, PSY, True)
    outs = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    for o in outs:
        assert np.abs(o["dirty"] - ref).max() < 1e-13 * np.abs(ref).max()
        assert abs(float(o["sumwt"][0]) - wgt.sum()) < 1e-12 * wgt.sum()
        lo, hi = int(o["lo"]), int(o["hi"])
        assert np.abs(o["vis"] - vref[:, lo:hi]).max() < 1e-12 * np.abs(vref).max()
    assert np.array_equal(outs[0]["dirty"], outs[1]["dirty"])  # identical on every rank


def test_normalise_zero_sumwt():
    d = torch.ones((4, 4), dtype=torch.float64)
    assert float(b2dist.normalise(d.clone(), 2.0)[0, 0]) == 0.5
    assert float(b2dist.normalise(d.clone(), 0.0).abs().max()) == 0.0   # imaging/base.py:127-128


def test_wplane_selection_partitions_and_balances():
    """w-plane sharding: the ranks' selections partition the samples, each rank's |w| range is an
    interval, and sample counts are balanced when planes are cheap."""
    import torch
    from ska_sdp_func_python_b200 import distributed as b2d
    rng = np.random.default_rng(5)
    nrow, nchan, world = 3000, 7, 4
    uvw = rng.normal(size=(nrow, 3)) * np.array([300.0, 300.0, 40.0])
    freq = np.linspace(100e6, 130e6, nchan)
    th = b2d.wplane_thresholds(uvw, freq, world, npix=64, pixsize=1e-3, block_rows=700)
    assert len(th) == world + 1 and th[0] == 0.0 and th[-1] == float("inf")
    assert all(a <= b for a, b in zip(th, th[1:]))
    sels = [b2d.wplane_selection(uvw, freq, world, r, 64, 1e-3, thresholds=th, block_rows=700)
            for r in range(world)]
    allsel = torch.cat(sels).numpy()
    assert len(allsel) == nrow * nchan and len(np.unique(allsel)) == nrow * nchan
    wl = np.abs(uvw[:, 2:3]) * freq[None, :] / 299792458.0
    for r, s in enumerate(sels):
        s = s.numpy()
        assert np.all(np.diff(s) > 0)
        ws = wl.reshape(-1)[s]
        assert ws.min() >= th[r] and ws.max() < th[r + 1]
        assert 0.5 * nrow * nchan / world < len(s) < 2.0 * nrow * nchan / world
    ms = rng.normal(size=(nrow, nchan))
    back = np.zeros_like(ms)
    for s in sels:
        b2d.scatter_back(back, s, b2d.compact(ms, s))
    assert np.array_equal(back, ms)


def test_wplane_rebalance_follows_measured_costs():
    """Measured re-balancing: equal measured costs reproduce a sample-balanced cut; when the outer
    ranks report dearer samples their intervals shrink (thresholds move outwards less far)."""
    from ska_sdp_func_python_b200 import distributed as b2d
    rng = np.random.default_rng(6)
    nrow, nchan, world = 4000, 5, 4
    uvw = rng.normal(size=(nrow, 3)) * np.array([300.0, 300.0, 40.0])
    freq = np.linspace(100e6, 130e6, nchan)
    th0 = b2d.wplane_thresholds(uvw, freq, world, npix=64, pixsize=1e-3)
    sels = [b2d.wplane_selection(uvw, freq, world, r, 64, 1e-3, thresholds=th0) for r in range(world)]
    n = [int(s.numel()) for s in sels]
    flat = b2d.wplane_rebalance(uvw, freq, world, 64, 1e-3, th0, n, [1e-3 * k for k in n], 0.0)
    cnt = [int(b2d.wplane_selection(uvw, freq, world, r, 64, 1e-3, thresholds=flat).numel())
           for r in range(world)]
    assert max(cnt) - min(cnt) < 0.05 * nrow * nchan          # uniform cost -> equal counts
    dear = b2d.wplane_rebalance(uvw, freq, world, 64, 1e-3, th0, n,
                                [1e-3 * k * (1 + 2 * r) for r, k in enumerate(n)], 0.0)
    assert dear[0] == 0.0 and dear[-1] == float("inf")
    assert all(a <= b for a, b in zip(dear, dear[1:]))
    cnt2 = [int(b2d.wplane_selection(uvw, freq, world, r, 64, 1e-3, thresholds=dear).numel())
            for r in range(world)]
    assert sum(cnt2) == nrow * nchan
    assert cnt2[0] > cnt[0] and cnt2[-1] < cnt[-1]            # cheap core rank takes more samples
