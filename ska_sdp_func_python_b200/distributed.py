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
    sw = sw.reshape(np.shape(sumwt_local)) if np.ndim(sumwt_local) > 1 else sw
    allreduce_image_sumwt(dirty, sw, group)
    if normalise_result:
        if sw.numel() == 1:
            normalise(dirty, float(sw.item()))
        elif dirty.dim() > 2 and sw.numel() == int(np.prod(dirty.shape[:-2])):
            # one sumwt per leading (chan, pol) plane: normalise_sumwt's 2-D branch
            planes = dirty.reshape((-1,) + tuple(dirty.shape[-2:]))
            for k, w in enumerate(sw.reshape(-1).tolist()):
                normalise(planes[k], w)
        else:
            raise ValueError(f"normalise_result=True needs one sumwt per image plane: image "
                             f"{tuple(dirty.shape)}, sumwt {tuple(sw.shape)}")
    return dirty, sw.cpu().numpy()


def predict_sharded(local_predict, model, src=0, group=None):
    """Broadcast the model image from rank ``src`` and degrid the rank's own samples with
    ``local_predict(model)``."""
    dist = _dist()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.broadcast(model, src=src, group=group)
    return local_predict(model)


# ----------------------------------------------------------------------------------------
# w-plane sharding (option 2): rank r owns the samples whose |w| (in
# wavelengths) lies in [t_r, t_{r+1}).  Its plan then spans only the w-planes of that range (plus
# the W-1 planes of kernel support), so the per-plane FFT / screen work divides by the number
# of ranks together with the gridding work.  No exchange between ranks: invert ends with the
# same single all-reduce of the image, predict outputs are disjoint by construction.
# ----------------------------------------------------------------------------------------
C_M_S = 299792458.0


def _plane_cost_in_samples(npix, precision):
    """Cost of one more w-plane (cuFFT passes + screen + memset, es_kernel.h cost model) in units
    of the cost of one sample."""
    csize = 8 if precision in ("single", 0) else 16
    nu = 1.5 * npix
    t_plane = nu * nu * csize * (4.0 / 3.3e12 + 1.5 / 6.0e12) + npix * npix * csize / 6.0e12
    t_sample = 65e-12 if csize == 8 else 1.5e-9
    return t_plane / t_sample


def _w_histogram(uvw, freq, nbins, block_rows):
    """Histogram of |w| in wavelengths over all (row, chan) samples: (hist [nbins] f64, wmax)."""
    import torch
    uvw_t = torch.as_tensor(uvw)
    dev = uvw_t.device
    w = uvw_t[:, 2].abs().to(torch.float64)
    fs = torch.as_tensor(np.asarray(freq, dtype=np.float64) / C_M_S, device=dev)
    wmax = float(w.max().item()) * float(fs.max().item()) if w.numel() else 0.0
    hist = torch.zeros(nbins, dtype=torch.float64, device=dev)
    if wmax <= 0.0:
        return hist, 0.0
    scale = nbins / (wmax * (1.0 + 1e-12))
    for r0 in range(0, w.numel(), block_rows):
        b = (w[r0:r0 + block_rows, None] * fs[None, :] * scale).to(torch.int64).clamp_(0, nbins - 1)
        hist += torch.bincount(b.reshape(-1), minlength=nbins).to(torch.float64)
    return hist, wmax


def _plane_spacing(npix, pixsize, sigma):
    """dw = 1 / (2 sigma max|n-1|) of the reference model."""
    xm = pixsize * npix / 2.0
    r2 = min(2.0 * xm * xm, 0.999)
    nm1 = r2 / (np.sqrt(1.0 - r2) + 1.0)
    return 0.5 / (sigma * max(nm1, 1e-12))


def _cut(cost, wmax, world_size):
    import torch
    nbins = cost.numel()
    total = float(cost[-1].item())
    targets = torch.tensor([total * r / world_size for r in range(1, world_size)],
                           dtype=torch.float64, device=cost.device)
    cut = torch.searchsorted(cost, targets).clamp_(0, nbins - 1).cpu().numpy()
    edges = [0.0] + [float((c + 1) * wmax / nbins) for c in cut] + [float("inf")]
    for i in range(1, len(edges)):  # monotone even for degenerate histograms
        edges[i] = max(edges[i], edges[i - 1])
    return edges


def wplane_thresholds(uvw, freq, world_size, npix, pixsize, precision="single", sigma=1.5,
                      nbins=8192, block_rows=1 << 16):
    """|w| thresholds t_0 = 0 < t_1 < ... < t_world = inf that balance (samples + planes) cost.
    uvw: [nrow, 3] metres (numpy or torch, any device), freq: [nchan] Hz.  Deterministic: every
    rank computes the same thresholds from the same inputs, no communication."""
    import torch
    hist, wmax = _w_histogram(uvw, freq, nbins, block_rows)
    if world_size <= 1 or wmax <= 0.0:
        return [0.0] + [float("inf")] * max(1, world_size)
    planes_per_bin = (wmax / nbins) / _plane_spacing(npix, pixsize, sigma)
    cost = torch.cumsum(hist + planes_per_bin * _plane_cost_in_samples(npix, precision), 0)
    return _cut(cost, wmax, world_size)


def wplane_rebalance(uvw, freq, world_size, npix, pixsize, thresholds, samples_per_rank,
                     vis_ms_per_rank, plane_ms, sigma=1.5, nbins=8192, block_rows=1 << 16):
    """New thresholds from MEASURED costs of a first step under ``thresholds``: the samples of
    rank r's old |w| interval cost vis_ms_per_rank[r] / samples_per_rank[r] each (sparse long
    baselines are several times dearer per sample than the dense core, which an a-priori model
    does not know), every w-plane costs ``plane_ms`` (FFT + screen + clear, both directions).  All
    ranks must pass the same (all-gathered) numbers; no communication here."""
    import torch
    hist, wmax = _w_histogram(uvw, freq, nbins, block_rows)
    if world_size <= 1 or wmax <= 0.0:
        return [0.0] + [float("inf")] * max(1, world_size)
    centres = (torch.arange(nbins, dtype=torch.float64, device=hist.device) + 0.5) * (wmax / nbins)
    edges = torch.tensor([t if np.isfinite(t) else 4.0 * wmax for t in thresholds[1:-1]],
                         dtype=torch.float64, device=hist.device)
    owner = torch.searchsorted(edges, centres, right=True)          # old rank of every bin
    per = torch.tensor([v / max(n, 1) for v, n in zip(vis_ms_per_rank, samples_per_rank)],
                       dtype=torch.float64, device=hist.device)
    # ranks that held too few samples to time reliably inherit their neighbour's cost
    few = 1e-3 * float(sum(samples_per_rank))
    for r in range(1, world_size):
        if samples_per_rank[r] < few:
            per[r] = per[r - 1]
    planes_per_bin = (wmax / nbins) / _plane_spacing(npix, pixsize, sigma)
    cost = torch.cumsum(hist * per[owner] + planes_per_bin * float(plane_ms), 0)
    return _cut(cost, wmax, world_size)


def wplane_selection(uvw, freq, world_size, rank, npix, pixsize, precision="single", sigma=1.5,
                     thresholds=None, block_rows=1 << 16):
    """Indices (row*nchan + chan, ascending, int64 tensor on uvw's device) of the samples owned
    by ``rank`` under w-plane sharding.  The selections of all ranks partition the samples."""
    import torch
    if thresholds is None:
        thresholds = wplane_thresholds(uvw, freq, world_size, npix, pixsize, precision, sigma,
                                       block_rows=block_rows)
    lo, hi = thresholds[rank], thresholds[rank + 1]
    uvw_t = torch.as_tensor(uvw)
    dev = uvw_t.device
    w = uvw_t[:, 2].abs().to(torch.float64)
    fs = torch.as_tensor(np.asarray(freq, dtype=np.float64) / C_M_S, device=dev)
    nchan = fs.numel()
    parts = []
    for r0 in range(0, w.numel(), block_rows):
        wl = w[r0:r0 + block_rows, None] * fs[None, :]
        m = (wl >= lo) & (wl < hi)
        parts.append(torch.nonzero(m.reshape(-1)).reshape(-1) + r0 * nchan)
    if not parts:
        return torch.zeros(0, dtype=torch.int64, device=dev)
    return torch.cat(parts)


def compact(samples, sel):
    """The rank's own entries of a full [nrow, nchan] sample array (torch tensor or numpy)."""
    import torch
    if isinstance(samples, np.ndarray):
        return samples.reshape(-1)[np.asarray(sel.cpu() if hasattr(sel, "cpu") else sel)]
    return samples.reshape(-1)[torch.as_tensor(sel, device=samples.device)]


def scatter_back(full, sel, values):
    """Write the rank's predicted samples into a full [nrow, nchan] array (in place)."""
    flat = full.reshape(-1)
    flat[sel if not isinstance(full, np.ndarray) else np.asarray(sel.cpu() if hasattr(sel, "cpu") else sel)] = values
    return full
