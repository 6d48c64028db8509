"""Tensor-core degridder (csrc/vis_tc.cu: tcgen05 / TMEM, 3 x TF32 split) against the exact DFT
oracle and against the warp-run kernel it replaces on dense runs.

Mode 2 sends every chunk of 1024 sorted samples through the tensor-core kernel (sparse test
geometries then exercise tiles of one or two samples, run boundaries inside and across chunks,
zero-weight rows, flipped samples, the ragged last chunk); mode 1 is the production setting: a
classification pass picks the chunks with long runs, the warp-run kernel takes the rest, and every
sample must be written exactly once.  The degridded samples replace what the reference gets from
ducc0.wgridder.dirty2ms (/root/reference/src/ska_sdp_func_python/imaging/ng.py:99-112); bound:
relative L2 error <= epsilon against the exact DFT (north_star: fp32, max-abs <= 10 epsilon)."""
import numpy as np
import pytest

from conftest import make_case, rel_l2, rel_max

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wg(built):
    import torch
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    from ska_sdp_func_python_b200 import wgridder
    return wgridder


@pytest.fixture()
def tc_mode(built):
    def set_mode(m):
        assert built.b2g_set_tc_mode(int(m)) == 0
    yield set_mode
    built.b2g_set_tc_mode(1)     # the library default


@pytest.fixture()
def slab_cap(built):
    def set_cap(n):
        assert built.b2g_set_max_slab_planes(int(n)) == 0
    yield set_cap
    built.b2g_set_max_slab_planes(0)


def _dense_case(seed, nrow, nchan, npx, npy, psx, psy, wfrac, nclump):
    """Samples concentrated on a few uv-cells: long runs of equal footprint origin (what the dense
    core of an SKA-LOW observation looks like after the sort), plus a sparse remainder."""
    uvw, freq, ms, wgt = make_case(seed, nrow, nchan, npx, npy, psx, psy, wfrac, clump=False)
    rng = np.random.default_rng(seed + 1)
    centres = uvw[rng.integers(0, nrow, nclump)]
    k = (3 * nrow) // 4
    which = rng.integers(0, nclump, k)
    # within a third of a cell of the clump centre (cell = 1 / (npix * pixsize * sigma) wavelengths)
    lam = freq[-1] / 299792458.0
    cell_u, cell_v = 1.0 / (npx * psx * 2.0) / lam, 1.0 / (npy * psy * 2.0) / lam
    uvw[:k] = centres[which]
    uvw[:k, 0] += rng.uniform(-0.3, 0.3, k) * cell_u
    uvw[:k, 1] += rng.uniform(-0.3, 0.3, k) * cell_v
    uvw[:k, 2] *= 1.0 + rng.uniform(-1e-4, 1e-4, k)
    return uvw, freq, ms, wgt


@pytest.mark.parametrize("eps", [1e-5, 1e-4, 2e-3])
@pytest.mark.parametrize("geom", [(1500, 3, 128, 112, 2.0e-3, 2.2e-3, 1.0), (2500, 2, 96, 160, 3.0e-4, 2.0e-4, 0.1)])
def test_tc_all_chunks_against_dft(wg, tc_mode, slab_cap, eps, geom):
    from oracle import dft as odft
    nrow, nchan, npx, npy, psx, psy, wfrac = geom
    uvw, freq, ms, wgt = make_case(31, nrow, nchan, npx, npy, psx, psy, wfrac)
    rng = np.random.default_rng(9)
    model = np.zeros((npx, npy), dtype=np.float32)
    model[rng.integers(0, npx, 50), rng.integers(0, npy, 50)] = rng.uniform(1, 2, 50)
    vref = odft.dirty2ms(uvw, freq, model.astype(np.float64), wgt, psx, psy, True)
    res = {}
    for name, mode, cap in (("warprun", 0, 0), ("tc", 2, 0), ("tc_slabs", 2, 2), ("tc_auto", 1, 0)):
        tc_mode(mode)
        slab_cap(cap)
        with wg.Plan(uvw, freq, npx, npy, psx, psy, eps, True, precision="single") as p:
            assert p.info()["W"] <= 7
            v = p.predict(model, wgt.astype(np.float32))
            r = p.residual(model, ms.astype(np.complex64), wgt.astype(np.float32))
            r0 = p.invert((ms - v / np.where(wgt == 0, 1, wgt)).astype(np.complex64), wgt.astype(np.float32))
        res[name] = (v, r, r0)
    for name, (v, r, r0) in res.items():
        assert np.isfinite(v).all(), name
        assert rel_l2(v, vref) < eps and rel_max(v, vref) < 10 * eps, name
        # the fused residual call degrids with the same kernel (sub_from path)
        assert rel_l2(r, r0) < 20 * eps, name
    for name in ("tc", "tc_slabs", "tc_auto"):
        # same kernel, different fp32 arithmetic (3 x TF32 split: ~1e-6 per product): far inside epsilon
        assert rel_l2(res[name][0], res["warprun"][0]) < max(4e-6, eps / 20), name
        assert np.array_equal(res[name][0] == 0, res["warprun"][0] == 0), name   # zero-weight samples


@pytest.mark.parametrize("mode", [1, 2])
def test_tc_dense_runs(wg, tc_mode, mode):
    """Long runs (mean > 100 samples per origin): the production case.  In mode 1 the classified
    chunks and the warp-run remainder must cover every sample exactly once."""
    from oracle import dft as odft
    nrow, nchan, npx, npy, psx, psy, wfrac = 6000, 4, 256, 256, 1.0e-3, 1.0e-3, 0.5
    uvw, freq, ms, wgt = _dense_case(5, nrow, nchan, npx, npy, psx, psy, wfrac, nclump=12)
    rng = np.random.default_rng(2)
    model = np.zeros((npx, npy), dtype=np.float32)
    model[rng.integers(0, npx, 80), rng.integers(0, npy, 80)] = rng.uniform(1, 2, 80)
    eps = 1e-5
    rows, chans = rng.integers(0, nrow, 400), rng.integers(0, nchan, 400)
    vref = odft.dirty2ms_samples(uvw, freq, model.astype(np.float64), wgt, psx, psy, True, rows, chans)
    import torch
    dev = torch.device("cuda", 0)
    uvw_d, model_d = torch.as_tensor(uvw, device=dev), torch.as_tensor(model, device=dev)
    wgt_d = torch.as_tensor(wgt.astype(np.float32), device=dev)
    out = {}
    for m in (0, mode):
        tc_mode(m)
        with wg.Plan(uvw_d, freq, npx, npy, psx, psy, eps, True, precision="single") as p:
            # device-resident output pre-filled with NaN: a sample nobody writes stays NaN
            sentinel = torch.full((nrow, nchan), float("nan"), device=dev, dtype=torch.complex64)
            out[m] = p.predict(model_d, wgt_d, out=sentinel).cpu().numpy()
    assert np.isfinite(out[mode]).all()
    assert rel_l2(out[mode][rows, chans], vref) < eps
    assert rel_l2(out[mode], out[0]) < 4e-6
