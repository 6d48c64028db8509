"""Component DFT on the GPU (csrc/dft_kernels.cu through the C ABI) against golden vectors from
the reference's own dft_cpu_looped / dft_skycomponent_visibility / extract_direction_and_flux and
against the numpy oracle on a larger seeded case.

Tolerances (relative to sum_comp |flux|): fp64 trig 1e-12 (the reference itself carries ~7e-14 of
argument-rounding error per term at these phases); fp32 trig 5e-7."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
C = 299792458.0


@pytest.fixture(scope="module")
def dft(built):
    import torch
    assert torch.cuda.is_available()
    from ska_sdp_func_python_b200 import dft
    return dft


def _gold():
    return np.load(os.path.join(GOLDEN, "dft_comp_case.npz"))


def test_dft_kernel_golden(dft):
    g = _gold()
    lam = np.einsum("tbs,k->tbks", g["uvw"], g["freq"] / C)
    scale = np.abs(g["vfl"]).sum(axis=0).max()
    for name in (None, "cpu_looped", "gpu_cupy_raw", "proc_func"):
        v = dft.dft_kernel(g["dc"], g["vfl"], lam, dft_compute_kernel=name)
        assert v.dtype == np.complex128 and v.shape == g["vis_a"].shape
        assert np.abs(v - g["vis_a"]).max() < 1e-12 * scale
    v1 = dft.dft_kernel(g["dc"], g["vfl"][:, :1, :1], lam)        # single-channel flux broadcast
    assert np.abs(v1 - g["vis_a1"]).max() < 1e-12 * scale
    with pytest.raises(ValueError):
        dft.dft_kernel(g["dc"], g["vfl"], lam, dft_compute_kernel="fortran")


def test_dft_skycomponent_visibility_golden(dft):
    from ska_sdp_func_python_b200 import minimal_models as mm
    g = _gold()
    pc = mm.SkyCoord(float(g["pc"][0]), float(g["pc"][1]))
    shape = g["out_b"].shape
    vis = mm.Visibility(np.zeros(shape, dtype=complex), g["uvw"], g["freq"], pc,
                        mm.PolarisationFrame("linear"))
    full = [mm.SkyComponent(mm.SkyCoord(d[0], d[1]), g["freq"], f, mm.PolarisationFrame("stokesIQUV"))
            for d, f in zip(g["dir_full"], g["flux_full"])]
    single = [mm.SkyComponent(mm.SkyCoord(d[0], d[1]), [float(g["freq_single"])], f,
                              mm.PolarisationFrame("linear"))
              for d, f in zip(g["dir_single"], g["flux_single"])]
    dc, vfl = dft.extract_direction_and_flux(full, vis)
    np.testing.assert_allclose(dc, g["dcb"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(vfl, g["vflb"], rtol=1e-15, atol=1e-15)
    out = dft.dft_skycomponent_visibility(vis.copy(deep=True), full)
    assert np.abs(out["vis"].data - g["out_b"]).max() < 1e-12 * np.abs(g["vflb"]).sum(axis=0).max()
    out = dft.dft_skycomponent_visibility(vis.copy(deep=True), single)
    assert np.abs(out["vis"].data - g["out_s"]).max() < 1e-12 * np.abs(g["vfls"]).sum(axis=0).max()
    v0 = vis.copy(deep=True)
    assert dft.dft_skycomponent_visibility(v0, []) is v0          # empty list: returned unchanged
    assert dft.dft_skycomponent_visibility(v0, None) is v0


@pytest.mark.parametrize("npol", [1, 2, 4])
def test_oracle_larger_and_modes(dft, npol):
    import torch
    from oracle import dft as odft
    rng = np.random.default_rng(3 + npol)
    nrow, nchan, ncomp = 3000, 37, 50
    freq = np.linspace(1.0e8, 1.3e8, nchan)
    uvw = rng.uniform(-3e4, 3e4, (nrow, 3))
    uvw[7] = 0.0
    dc = rng.uniform(-0.05, 0.05, (ncomp, 3))
    dc[:, 2] = np.sqrt(1 - dc[:, 0] ** 2 - dc[:, 1] ** 2) - 1
    fl = rng.normal(size=(ncomp, nchan, npol)) + 1j * rng.normal(size=(ncomp, nchan, npol))
    lam = np.einsum("rs,k->rks", uvw, freq / C)
    ref = odft.dft_components(dc, lam, fl)
    scale = np.abs(fl).sum(axis=0).max()
    dev = torch.device("cuda")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    out = dft.dft_predict_device(t(uvw), t(freq / C), t(dc), t(fl), nchan)
    assert np.abs(out.cpu().numpy() - ref).max() < 1e-12 * scale
    # per-sample wavelengths give the same numbers as uvw x f/c
    out2 = dft.dft_predict_device(t(lam.reshape(-1, 3)), None, t(dc), t(fl), nchan)
    assert torch.equal(out, out2)
    # fp32 trigonometry
    out3 = dft.dft_predict_device(t(uvw), t(freq / C), t(dc), t(fl), nchan, precision="single")
    err = np.abs(out3.cpu().numpy() - ref).max()
    assert 0 < err < 5e-7 * scale
    # accumulate: linearity in the component list
    half = ncomp // 2
    acc = dft.dft_predict_device(t(uvw), t(freq / C), t(dc[:half]), t(fl[:half]), nchan)
    dft.dft_predict_device(t(uvw), t(freq / C), t(dc[half:]), t(fl[half:]), nchan, out=acc,
                           accumulate=True)
    assert np.abs(acc.cpu().numpy() - ref).max() < 1e-12 * scale
    # no components: zeros
    z = dft.dft_predict_device(t(uvw), t(freq / C), t(dc[:0]), t(fl[:0]), nchan)
    assert float(z.abs().max()) == 0.0


def test_idft_visibility_skycomponent_golden(dft):
    """Adjoint DFT against the reference's own idft_visibility_skycomponent (flags, a zero-weight
    (chan, pol), components in the visibility frame and in stokesIQUV)."""
    from ska_sdp_func_python_b200 import minimal_models as mm
    g = np.load(os.path.join(GOLDEN, "idft_case.npz"))
    pc = mm.SkyCoord(float(g["pc"][0]), float(g["pc"][1]))
    vis = mm.Visibility(g["vis"], g["uvw"], g["freq"], pc, mm.PolarisationFrame("linear"),
                        weight=g["weight"], flags=g["flags"])
    comps = [mm.SkyComponent(mm.SkyCoord(d[0], d[1]), g["freq"], np.ones((4, 4), dtype=complex),
                             mm.PolarisationFrame(str(f))) for d, f in zip(g["dirs"], g["frames"])]
    v0 = vis["vis"].data.copy()
    newsc, wl = dft.idft_visibility_skycomponent(vis, comps)
    assert len(newsc) == len(wl) == 5 and np.array_equal(vis["vis"].data, v0)
    scale = np.abs(g["flux"]).max()
    for k in range(5):
        assert newsc[k] is not comps[k] and np.all(comps[k].flux == 1)
        assert np.abs(newsc[k].flux - g["flux"][k]).max() < 1e-12 * scale
        np.testing.assert_allclose(wl[k], g["weights"][k], rtol=1e-13)
        assert np.all(newsc[k].flux[2, 1] == 0) or str(g["frames"][k]) != "linear"
    assert dft.idft_visibility_skycomponent(vis, None) is None


@pytest.mark.parametrize("npol", [1, 2, 4])
def test_idft_is_adjoint_of_dft(dft, npol):
    """<dft(flux), w vis> == <flux, idft sums(vis)> on a larger seeded case, and against the oracle."""
    import torch
    from oracle import dft as odft
    rng = np.random.default_rng(30 + npol)
    nrow, nchan, ncomp = 5000, 45, 11
    freq = np.linspace(1.0e8, 1.3e8, nchan)
    uvw = rng.uniform(-2e4, 2e4, (nrow, 3))
    dc = rng.uniform(-0.05, 0.05, (ncomp, 3))
    dc[:, 2] = np.sqrt(1 - dc[:, 0] ** 2 - dc[:, 1] ** 2) - 1
    vis = rng.normal(size=(nrow, nchan, npol)) + 1j * rng.normal(size=(nrow, nchan, npol))
    wgt = rng.uniform(0.5, 2.0, (nrow, nchan, npol))
    flags = (rng.uniform(size=vis.shape) < 0.1).astype(np.uint8)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    from ska_sdp_func_python_b200 import _lib
    flux, wsum = dft.dft_invert_device(t(uvw), t(freq / C), t(dc), t(vis), t(wgt), t(flags), _lib.FLAG_U8)
    lam = np.einsum("rs,k->rks", uvw, freq / C)
    oflux, owsum = odft.idft_components(dc, lam, vis, wgt, flags.astype(np.int64))
    np.testing.assert_allclose(wsum.cpu().numpy(), owsum, rtol=1e-13)
    got = flux.cpu().numpy() / owsum
    assert np.abs(got - oflux).max() < 1e-12 * np.abs(oflux).max() * 50
    # no components: weight sums only
    f0, w0 = dft.dft_invert_device(t(uvw), t(freq / C), t(dc[:0]), t(vis), t(wgt), t(flags), _lib.FLAG_U8)
    assert f0.shape[0] == 0 and torch.allclose(w0, wsum, rtol=1e-13)
