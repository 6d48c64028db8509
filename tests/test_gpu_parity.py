"""Parity of the CUDA gridder / degridder (through the C ABI) against the oracle.

Tolerances (BASELINE.json north_star): relative L2 error <= epsilon against the exact DFT (the
accuracy contract ducc0 itself states), max-abs error <= 10*epsilon of the image / visibility
peak in fp32 mode and <= 1e-10 of the peak in fp64 mode at epsilon = 1e-12; sumwt bit-exact.
"""
import ctypes

import numpy as np
import pytest

from conftest import make_case, rel_l2, rel_max

pytestmark = pytest.mark.gpu

GEOS = {
    # name: nrow, nchan, npix_x, npix_y, pixsize_x, pixsize_y, wfrac
    "narrow": (3000, 4, 256, 192, 2.0e-4, 2.5e-4, 0.05),
    "wide": (2000, 1, 128, 128, 2.0e-3, 2.0e-3, 0.3),
}
_CACHE = {}


def _oracle(geo, ws):
    from oracle import dft as odft
    key = (geo, ws)
    if key not in _CACHE:
        nrow, nchan, npx, npy, psx, psy, wfrac = GEOS[geo]
        uvw, freq, ms, wgt = make_case(100 + len(geo), nrow, nchan, npx, npy, psx, psy, wfrac)
        rng = np.random.default_rng(7)
        model = np.zeros((npx, npy))
        model[rng.integers(0, npx, 40), rng.integers(0, npy, 40)] = rng.uniform(1, 2, 40)
        model[0, 0], model[npx - 1, npy - 1] = 1.0, -1.0
        d = odft.ms2dirty(uvw, freq, ms, wgt, npx, npy, psx, psy, ws)
        v = odft.dirty2ms(uvw, freq, model, wgt, psx, psy, ws)
        _CACHE[key] = (uvw, freq, ms, wgt, model, d, v)
    return _CACHE[key]


@pytest.fixture(scope="module")
def wg(built):
    import torch
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    from ska_sdp_func_python_b200 import wgridder
    return wgridder


@pytest.mark.parametrize("geo", list(GEOS))
@pytest.mark.parametrize("ws", [False, True])
@pytest.mark.parametrize("prec,eps", [("double", 1e-12), ("double", 1e-7), ("single", 1e-5),
                                      ("single", 1e-3)])
@pytest.mark.parametrize("algo", ["simple", "warprun"])
def test_parity_against_dft(wg, geo, ws, prec, eps, algo):
    uvw, freq, ms, wgt, model, dref, vref = _oracle(geo, ws)
    nrow, nchan, npx, npy, psx, psy, _ = GEOS[geo]
    code = wg.ALGO_SIMPLE if algo == "simple" else wg.ALGO_WARPRUN
    try:
        plan = wg.Plan(uvw, freq, npx, npy, psx, psy, eps, ws, precision=prec, algo=code)
    except wg.WgridderError as exc:
        if "do not support" in str(exc):
            pytest.skip(str(exc))
        raise
    cdt, rdt = (np.complex128, np.float64) if prec == "double" else (np.complex64, np.float32)
    with plan:
        d = plan.invert(ms.astype(cdt), wgt.astype(rdt))
        v = plan.predict(model.astype(rdt), wgt.astype(rdt))
    assert rel_l2(d, dref) < eps and rel_l2(v, vref) < eps
    bound = 1e-10 if (prec == "double" and eps <= 1e-12) else 10 * eps
    assert rel_max(d, dref) < bound and rel_max(v, vref) < bound


@pytest.mark.parametrize("prec,eps,tol", [("double", 1e-9, 2e-11), ("single", 1e-4, 5e-5)])
@pytest.mark.parametrize("ws", [False, True])
def test_adjointness(wg, prec, eps, tol, ws):
    """<A x, y> == <x, A^H y>: gridding and degridding use identical kernel values."""
    uvw, freq, ms, _, _, _, _ = _oracle("narrow", ws)
    nrow, nchan, npx, npy, psx, psy, _ = GEOS["narrow"]
    rng = np.random.default_rng(3)
    cdt, rdt = (np.complex128, np.float64) if prec == "double" else (np.complex64, np.float32)
    # image energy away from the rim, where the gridding correction amplifies rounding noise
    x = np.zeros((npx, npy), dtype=rdt)
    x[npx // 8:-npx // 8, npy // 8:-npy // 8] = rng.normal(size=(npx - 2 * (npx // 8), npy - 2 * (npy // 8)))
    y = (rng.normal(size=ms.shape) + 1j * rng.normal(size=ms.shape)).astype(cdt)
    with wg.Plan(uvw, freq, npx, npy, psx, psy, eps, ws, precision=prec) as plan:
        Ax, Ahy = plan.predict(x), plan.invert(y)
    lhs = np.vdot(y.astype(np.complex128), Ax.astype(np.complex128)).real
    rhs = np.vdot(Ahy.astype(np.float64), x.astype(np.float64))
    assert abs(lhs - rhs) < tol * max(abs(lhs), abs(rhs))


def test_one_shot_c_abi_equals_plan(wg, built):
    """b2g_ms2dirty / b2g_dirty2ms (the two ducc0 calls) == plan + execute."""
    from ska_sdp_func_python_b200 import _lib
    uvw, freq, ms, wgt, model, dref, vref = _oracle("narrow", True)
    nrow, nchan, npx, npy, psx, psy, _ = GEOS["narrow"]
    dirty = np.empty((npx, npy))
    ms_c = np.ascontiguousarray(ms)
    vp = lambda a: ctypes.c_void_p(a.ctypes.data)  # noqa: E731
    _lib.check(built.b2g_ms2dirty(nrow, nchan, vp(uvw), vp(freq), vp(ms_c), vp(wgt), _lib.F64,
                                  _lib.HOST, npx, npy, psx, psy, 0, 0, 1e-9, 1, _lib.F64, vp(dirty),
                                  _lib.F64, None))
    assert rel_l2(dirty, dref) < 1e-9
    out = np.empty((nrow, nchan), dtype=np.complex128)
    _lib.check(built.b2g_dirty2ms(nrow, nchan, vp(uvw), vp(freq), vp(model), _lib.F64, vp(wgt),
                                  _lib.F64, _lib.HOST, npx, npy, psx, psy, 0, 0, 1e-9, 1, _lib.F64,
                                  vp(out), None))
    assert rel_l2(out, vref) < 1e-9
    assert rel_l2(wg.ms2dirty(uvw, freq, ms, wgt, npx, npy, psx, psy, 0, 0, 1e-9, True), dirty) < 1e-13


def test_device_pointers_equal_host_pointers(wg):
    import torch
    uvw, freq, ms, wgt, model, _, _ = _oracle("narrow", True)
    nrow, nchan, npx, npy, psx, psy, _ = GEOS["narrow"]
    dev = torch.device("cuda:0")
    with wg.Plan(uvw, freq, npx, npy, psx, psy, 1e-5, True, precision="single") as ph:
        dh = ph.invert(ms.astype(np.complex64), wgt.astype(np.float32))
        vh = ph.predict(model.astype(np.float32))
    with wg.Plan(torch.as_tensor(uvw, device=dev), freq, npx, npy, psx, psy, 1e-5, True,
                 precision="single") as pd:
        dd = pd.invert(torch.as_tensor(ms.astype(np.complex64), device=dev),
                       torch.as_tensor(wgt.astype(np.float32), device=dev))
        vd = pd.predict(torch.as_tensor(model.astype(np.float32), device=dev))
    assert dd.is_cuda and vd.is_cuda
    # same kernels, same sort: only the fp32 atomics' order may differ
    assert rel_max(dd.cpu().numpy(), dh) < 1e-5 and rel_max(vd.cpu().numpy(), vh) < 1e-6


def test_mixed_dtypes_through_the_abi(wg):
    """fp64 host data into an fp32 plan and back (the conversion kernels of the staging path)."""
    uvw, freq, ms, wgt, model, dref, vref = _oracle("narrow", True)
    nrow, nchan, npx, npy, psx, psy, _ = GEOS["narrow"]
    with wg.Plan(uvw, freq, npx, npy, psx, psy, 1e-5, True, precision="single") as p:
        d = p.invert(ms, wgt, out_dtype=np.float64)
        v = p.predict(model, wgt, out_dtype=np.complex128)
    assert d.dtype == np.float64 and v.dtype == np.complex128
    assert rel_l2(d, dref) < 1e-5 and rel_l2(v, vref) < 1e-5


@pytest.mark.parametrize("prec,eps", [("single", 1e-5), ("double", 1e-9)])
@pytest.mark.parametrize("ws", [False, True])
def test_host_streaming_chunks(wg, built, prec, eps, ws):
    """Plans built from host uvw stream the sample arrays in row chunks (sort key carries the
    chunk); the result must match the DFT oracle and the unchunked plan."""
    uvw, freq, ms, wgt, model, dref, vref = _oracle("narrow", ws)
    nrow, nchan, npx, npy, psx, psy, _ = GEOS["narrow"]
    cdt, rdt = (np.complex128, np.float64) if prec == "double" else (np.complex64, np.float32)
    msc, wc, mc = ms.astype(cdt), wgt.astype(rdt), model.astype(rdt)
    try:
        built.b2g_set_host_chunk_samples(1700)  # 12000 samples -> 7 chunks, ragged last one
        with wg.Plan(uvw, freq, npx, npy, psx, psy, eps, ws, precision=prec) as plan:
            assert plan.info()["key_bits"] > 0
            d1, v1 = plan.invert(msc, wc), plan.predict(mc)
            d1b = plan.invert(msc, None)
        built.b2g_set_host_chunk_samples(0)
        with wg.Plan(uvw, freq, npx, npy, psx, psy, eps, ws, precision=prec) as plan:
            d0, v0 = plan.invert(msc, wc), plan.predict(mc)
            d0b = plan.invert(msc, None)
    finally:
        built.b2g_set_host_chunk_samples(1 << 24)
    assert rel_l2(d1, dref) < eps
    # same samples, same kernel values; only the order of the floating-point sums differs
    # (rounding noise of the atomics, amplified towards the image rim by the gridding correction:
    # L2 for the images)
    tol = 1e-5 if prec == "single" else 1e-11
    errs = (rel_l2(d1, d0), rel_max(v1, v0), rel_l2(d1b, d0b))
    assert max(errs) < tol, errs
    # predict without weights covers every sample, also the zero-weight rows of `wgt`
    assert np.all(np.isfinite(v1)) and rel_l2(v1[wgt != 0], (vref / np.where(wgt == 0, 1, wgt))[wgt != 0]) < eps


@pytest.mark.parametrize("prec,eps", [("single", 1e-5), ("double", 1e-9)])
@pytest.mark.parametrize("space", ["host", "device"])
def test_wplane_shards_sum_to_full(wg, prec, eps, space):
    """w-plane sharding (b2g_plan_create_sel): plans over the per-rank sample selections cover
    fewer w-planes each, their dirty images add up to the full image (the all-reduce), and their
    predicted samples are the full prediction restricted to the selection."""
    import torch
    from ska_sdp_func_python_b200 import distributed as b2d
    uvw, freq, ms, wgt, model, dref, vref = _oracle("wide", True)
    nrow, nchan, npx, npy, psx, psy, _ = GEOS["wide"]
    cdt, rdt = (np.complex128, np.float64) if prec == "double" else (np.complex64, np.float32)
    world = 3
    th = b2d.wplane_thresholds(uvw, freq, world, npx, psx, prec)
    with wg.Plan(uvw, freq, npx, npy, psx, psy, eps, True, precision=prec) as full:
        nplanes_full = full.info()["nplanes"]
        v_full = full.predict(model.astype(rdt))
    dsum = np.zeros((npx, npy))
    vback = np.zeros((nrow, nchan), dtype=np.complex128)
    nsel_total = 0
    for r in range(world):
        sel = b2d.wplane_selection(uvw, freq, world, r, npx, psx, thresholds=th).numpy()
        nsel_total += len(sel)
        ms_r, wgt_r = ms.reshape(-1)[sel].astype(cdt), wgt.reshape(-1)[sel].astype(rdt)
        if space == "device":
            dev = torch.device("cuda:0")
            plan = wg.Plan(torch.as_tensor(uvw, device=dev), freq, npx, npy, psx, psy, eps, True,
                           precision=prec, sel=torch.as_tensor(sel, device=dev))
            d = plan.invert(torch.as_tensor(ms_r, device=dev), torch.as_tensor(wgt_r, device=dev)).cpu().numpy()
            v = plan.predict(torch.as_tensor(model.astype(rdt), device=dev)).cpu().numpy()
        else:
            plan = wg.Plan(uvw, freq, npx, npy, psx, psy, eps, True, precision=prec, sel=sel)
            d, v = plan.invert(ms_r, wgt_r), plan.predict(model.astype(rdt))
        assert plan.info()["nvis"] == len(sel) and v.shape == (len(sel),)
        assert plan.info()["nplanes"] < nplanes_full
        plan.close()
        dsum += d
        b2d.scatter_back(vback, torch.as_tensor(sel), v)
    assert nsel_total == nrow * nchan
    assert rel_l2(dsum, dref) < eps
    # predict has no weights here: compare with the full plan's prediction; the shard plans have
    # their own w-plane lattices, so the two differ by the kernel's approximation error (< eps)
    assert rel_l2(vback, v_full) < eps
