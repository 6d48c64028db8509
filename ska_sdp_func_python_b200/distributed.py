"""Multi-GPU partitioning of invert / predict: one process per GPU, ``torch.distributed``
(NCCL over NVLink / NVSwitch) for the single exchange step each direction has.

The reference has no distributed code; its host-side list reductions define the semantics
that the collective must reproduce (/root/reference/src/ska_sdp_func_python/imaging/
imaging_helpers.py:25-58 ``sum_invert_results``: partial dirty images combine as the plain
sum of the UN-normalised images and of sumwt, normalised once at the end; :75-93
``sum_predict_results``: predicted visibilities of disjoint sample sets need no reduction).

Sharding: every rank owns a disjoint set of samples (a block of frequency channels, or a
block of rows).  invert  = local gridding -> ONE all-reduce(sum) of the un-normalised dirty
image and of sumwt -> normalise.  predict = broadcast of the model image from ``src`` ->
local degridding of the rank's own samples (no collective on the output).
"""
import numpy as np


def _dist():
    import torch.distributed as dist
    return dist


def shard_bounds(n, world_size, rank):
    """[lo, hi) of the contiguous block of n items owned by ``rank`` (sizes differ by <= 1)."""
    base, rem = divmod(int(n), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def channel_shards(nchan, world_size):
    """List of (lo, hi) channel blocks, one per rank."""
    return [shard_bounds(nchan, world_size, r) for r in range(world_size)]


def allreduce_image_sumwt(dirty, sumwt, group=None):
    """In-place sum over ranks of the un-normalised dirty image (tensor on the rank's device)
    and of sumwt (float64 tensor).  One collective each; the image dominates the bytes."""
    dist = _dist()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(dirty, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(sumwt, op=dist.ReduceOp.SUM, group=group)
    return dirty, sumwt


def normalise(dirty, sumwt_scalar):
    """normalise_sumwt for one (chan, pol) plane (imaging/base.py:121-128): divide by sumwt if
    it is positive, else zero the plane."""
    if float(sumwt_scalar) > 0.0:
        dirty /= float(sumwt_scalar)
    else:
        dirty.zero_()
    return dirty


def invert_sharded(local_invert, ms, wgt, sumwt_local, normalise_result=True, group=None):
    """``local_invert(ms, wgt)`` grids the rank's own samples into an un-normalised dirty image
    (a torch tensor); the partial images and sumwt are summed over ranks and normalised.
    Returns (dirty, sumwt) identical on every rank."""
    import torch
    dirty = local_invert(ms, wgt)
    sw = torch.as_tensor(np.atleast_1d(np.asarray(sumwt_local, dtype=np.float64)),
                         device=dirty.device)
    allreduce_image_sumwt(dirty, sw, group)
    total = float(sw.sum().item()) if sw.numel() == 1 else None
    if normalise_result and total is not None:
        normalise(dirty, total)
    return dirty, sw.cpu().numpy()


def predict_sharded(local_predict, model, src=0, group=None):
    """Broadcast the model image from rank ``src`` and degrid the rank's own samples with
    ``local_predict(model)``."""
    dist = _dist()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.broadcast(model, src=src, group=group)
    return local_predict(model)
