"""Multi-GPU entry of the drop-in API: invert_ng / predict_ng with ``shard="channel"`` (every rank
owns a block of frequency channels) and ``shard="wplane"`` (every rank holds the observation and
works on its own |w| interval), one process per rank, against the reference-generated golden
(tests/golden/ng_case_a.npz).  World size = min(2, visible GPUs) over NCCL; on a one-GPU box the
same code runs as a single-rank process group, and as two ranks sharing the GPU over gloo.
Reference semantics of the exchange step: imaging_helpers.py:25-58 (sum un-normalised images and
sumwt, normalise once), :75-93 (disjoint predictions need no reduction)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import load_ng_case, rel_max

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _launch_once(mode, world, out):
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "_shard_worker.py"), mode,
                                       str(out)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    logs = [p.communicate(timeout=600)[0] for p in procs]
    return [f"--- rank {r} (exit {p.returncode}) ---\n{log[-2500:]}"
            for r, (p, log) in enumerate(zip(procs, logs)) if p.returncode != 0]


def _launch(mode, world, out):
    """The ranks' processes must exit 0.  One retry if a process dies: seen once, in the very first
    launch on a fresh two-GPU box, the log was not captured and the next run of the same build
    passed (a rendezvous race is possible: _free_port releases the port before rank 0 binds it).
    The first attempt's log is written to stderr.  The numerical assertions are not retried."""
    failed = _launch_once(mode, world, out)
    if failed:
        sys.stderr.write("first launch failed, retrying once:\n" + "\n".join(failed) + "\n")
        failed = _launch_once(mode, world, out)
    assert not failed, "\n".join(failed)


@pytest.mark.timeout(900)
@pytest.mark.parametrize("mode", ["channel", "wplane"])
@pytest.mark.parametrize("world", [1, 2])
def test_sharded_invert_predict(built, tmp_path, mode, world):
    g, vis, im = load_ng_case("ng_case_a.npz")
    _launch(mode, world, tmp_path)
    outs = [np.load(tmp_path / f"{mode}_{r}.npz") for r in range(world)]
    pred = np.zeros_like(g["predicted"])
    for o in outs:
        # every rank ends with the same, fully normalised image and the total sumwt
        assert rel_max(o["dirty"], g["dirty_norm"]) < 1e-9
        if world == 1 and mode == "channel":
            assert np.array_equal(o["sumwt"], g["sumwt"])     # same sum, same order: bit-exact
        else:   # partial sums added by the all-reduce: the same value up to summation order
            assert np.allclose(o["sumwt"], g["sumwt"], rtol=1e-13, atol=0)
        if mode == "channel":
            pred[:, :, int(o["lo"]):int(o["hi"])] = o["vis"]
        else:
            pred += o["vis"]      # disjoint supports: the ranks' predictions add up
    assert rel_max(pred, g["predicted"]) < 1e-9
