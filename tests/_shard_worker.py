"""Worker of tests/test_gpu_shard.py: one rank of invert_ng / predict_ng with the ``shard`` kwarg.
Launched with RANK / WORLD_SIZE / MASTER_* in the environment; rank r uses cuda:(r % ngpu)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    from conftest import load_ng_case
    from ska_sdp_func_python_b200 import imaging
    from ska_sdp_func_python_b200 import minimal_models as mm
    from ska_sdp_func_python_b200.distributed import shard_bounds
    mode, out = sys.argv[1], sys.argv[2]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    ngpu = torch.cuda.device_count()
    torch.cuda.set_device(rank % ngpu)
    # NCCL needs one GPU per rank; with fewer GPUs than ranks fall back to gloo (CUDA tensors ok)
    backend = "nccl" if ngpu >= world else "gloo"
    dist.init_process_group(backend, rank=rank, world_size=world)
    g, vis, im = load_ng_case("ng_case_a.npz")
    nchan = vis["vis"].data.shape[2]
    if mode == "channel":
        lo, hi = shard_bounds(nchan, world, rank)
        sl = slice(lo, hi)
        mine = mm.Visibility(vis["vis"].data[:, :, sl].copy(), vis["uvw"].data, vis["frequency"].data[sl],
                             vis.phasecentre, vis.visibility_acc.polarisation_frame,
                             weight=vis["weight"].data[:, :, sl].copy(),
                             imaging_weight=vis["imaging_weight"].data[:, :, sl].copy(),
                             flags=vis["flags"].data[:, :, sl].copy())
    else:
        lo, hi = 0, nchan
        mine = vis
    d, sw = imaging.invert_ng(mine, im, normalise=True, epsilon=1e-10, shard=mode)
    model = im.copy(deep=True)
    if rank == 0:   # only the source rank holds the model; the others receive it by broadcast
        model["pixels"].data[...] = g["model"]
    p = imaging.predict_ng(mine, model, epsilon=1e-10, shard=mode)
    np.savez(os.path.join(out, f"{mode}_{rank}.npz"), dirty=d["pixels"].data, sumwt=sw,
             vis=p["vis"].data, lo=lo, hi=hi)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
