"""Density weighting / tapering on the GPU (csrc/weight_kernels.cu through the C ABI) against
(1) golden vectors produced by the reference's own weighting code
(tests/golden/make_golden_weight.py) and (2) the numpy oracle on larger seeded inputs.

Tolerances: the cell assignment and the per-sample arithmetic are the reference's own fp64
operations, so imaging weights are BIT-EXACT whenever the gridded weights are; the gridded weights
themselves are sums of up to ~100 doubles per cell added in a different order (atomics), hence
rtol 1e-13 for non-integer weights and exact equality for 0/1 weights."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
RTOL = 1e-13


@pytest.fixture(scope="module")
def wt(built):
    import torch
    assert torch.cuda.is_available()
    from ska_sdp_func_python_b200 import weighting
    return weighting


def load_case(name):
    from ska_sdp_func_python_b200 import minimal_models as mm
    g = np.load(os.path.join(GOLDEN, name))
    pf = mm.PolarisationFrame(str(g["pf"]))
    pc = mm.SkyCoord.from_deg(float(g["img_crval"][0]), float(g["img_crval"][1]))
    nt, nbl, nchan, npol = g["weight"].shape
    vis = mm.Visibility(np.zeros(g["weight"].shape, dtype=complex), g["uvw"], g["freq"], pc, pf,
                        weight=g["weight"], imaging_weight=g["imaging_weight"], flags=g["flags"])
    wcs = mm.WCS(g["img_crval"], g["img_crpix"], g["img_cdelt"])
    npix = int(g["npix"])
    im = mm.Image(np.zeros((int(g["img_nchan"]), npol, npix, npix)), wcs, pf)
    return g, vis, im


@pytest.mark.parametrize("name", ["weight_case_a.npz", "weight_case_b.npz"])
def test_golden_grid_and_reweight(wt, name):
    from ska_sdp_func_python_b200 import minimal_models as mm
    g, vis, im = load_case(name)
    gd = mm.create_griddata_from_image(im, vis.visibility_acc.polarisation_frame)
    gd, sumwt = wt.grid_visibility_weight_to_griddata(vis, gd)
    assert gd["pixels"].data.dtype == np.complex128
    np.testing.assert_allclose(gd["pixels"].data.real, g["grid"], rtol=RTOL, atol=0)
    assert np.all(gd["pixels"].data.imag == 0)
    np.testing.assert_allclose(sumwt, g["sumwt"], rtol=RTOL)
    # re-weighting from the REFERENCE's grid: bit-exact
    gd["pixels"].data = g["grid"].astype(complex)
    for key, kw in (("imw_uniform", dict(weighting="uniform")),
                    ("imw_robust0", dict(weighting="robust", robustness=0.0, sumwt=g["sumwt"])),
                    ("imw_robustm1", dict(weighting="robust", robustness=-1.0, sumwt=g["sumwt"])),
                    ("imw_robust_nosumwt", dict(weighting="robust", robustness=0.5)),
                    ("imw_natural", dict(weighting="natural"))):
        v = wt.griddata_visibility_reweight(vis.copy(deep=True), gd, **kw)
        assert np.array_equal(v["imaging_weight"].data, g[key]), key


@pytest.mark.parametrize("name", ["weight_case_a.npz", "weight_case_b.npz"])
def test_golden_weight_visibility(wt, name):
    g, vis, im = load_case(name)
    w0 = vis["weight"].data.copy()
    for key, kw in (("wv_uniform", dict(weighting="uniform")),
                    ("wv_robust", dict(weighting="robust", robustness=-0.5)),
                    ("wv_natural", dict(weighting="natural"))):
        v = vis.copy(deep=True)
        out = wt.weight_visibility(v, im, **kw)
        assert out is v                                            # in place, like the reference
        np.testing.assert_allclose(v["imaging_weight"].data, g[key], rtol=4 * RTOL, atol=0)
        assert np.array_equal(v["weight"].data, w0)
        assert np.array_equal((v["imaging_weight"].data == 0), (g[key] == 0))
    with pytest.raises(AssertionError):
        wt.weight_visibility(vis.copy(deep=True), im, weighting="briggs")
    with pytest.raises(AssertionError):
        wt.weight_visibility(vis.copy(deep=True), None)


@pytest.mark.parametrize("name", ["weight_case_a.npz", "weight_case_b.npz"])
def test_golden_tapers(wt, name):
    g, vis, im = load_case(name)
    v = wt.taper_visibility_gaussian(vis.copy(deep=True), beam=float(g["beam"]))
    np.testing.assert_allclose(v["imaging_weight"].data, g["taper_gaussian"], rtol=4e-16, atol=0,
                               equal_nan=True)
    v = wt.taper_visibility_tukey(vis.copy(deep=True), tukey=float(g["tukey"]))
    np.testing.assert_allclose(v["imaging_weight"].data, g["taper_tukey"], rtol=1e-15, atol=1e-17,
                               equal_nan=True)
    with pytest.raises(ValueError):
        wt.taper_visibility_gaussian(vis.copy(deep=True))


def _random_job(seed, nrow, nchan, npol, npix, cell, unit_weights):
    rng = np.random.default_rng(seed)
    freq = np.linspace(1.0e8, 1.1e8, nchan)
    umax = 0.55 / cell / (freq[-1] / 299792458.0)        # a few per cent fall off the grid
    uvw = rng.uniform(-1, 1, (nrow, 3)) * umax
    uvw[: nrow // 2] *= 0.03                               # dense core: long same-cell runs
    uvw[::501] = 0.0                                       # autocorrelations: the centre cell
    shape = (nrow, nchan, npol)
    weight = np.ones(shape) if unit_weights else rng.uniform(0.1, 3.0, shape)
    weight[::13] = 0.0
    flags = (rng.uniform(size=shape) < 0.07).astype(np.int64)
    return uvw, freq, weight, flags


@pytest.mark.parametrize("unit_weights", [True, False])
def test_oracle_larger(wt, unit_weights):
    from oracle import weighting as ow
    nrow, nchan, npol, npix, cell = 60000, 5, 2, 256, 5e-4
    uvw, freq, weight, flags = _random_job(5, nrow, nchan, npol, npix, cell, unit_weights)
    vis_to_im = np.array([0, 0, 1, 1, 1])
    cd = (-1.0 / (npix * cell), 1.0 / (npix * cell), npix // 2 + 1, npix // 2 + 1)
    G = wt.WeightGrid(uvw, freq, vis_to_im, 2, npix, npix, *cd)
    grid, sumwt, sumsq = wt.grid_weights(G, weight, flags)
    ogrid, osumwt = ow.grid_weights(uvw, freq, weight, flags, vis_to_im, 2, npix, npix, *cd)
    gh = grid.cpu().numpy()
    if unit_weights:
        assert np.array_equal(gh, ogrid)                           # integers: exact
        assert np.array_equal(sumwt, osumwt)
    else:
        np.testing.assert_allclose(gh, ogrid, rtol=RTOL, atol=0)
        np.testing.assert_allclose(sumwt, osumwt, rtol=RTOL)
    assert ogrid.max() > 50 and (ogrid == 0).any()
    np.testing.assert_allclose(sumsq, np.sum(ogrid ** 2), rtol=RTOL)
    # re-weighting from identical grids is bit-exact
    import torch
    og = torch.from_numpy(ogrid).cuda()
    for weighting, rob in (("uniform", 0.0), ("robust", -0.7), ("robust", 1.5)):
        f2 = ow.robust_f2(ogrid, osumwt, rob) if weighting == "robust" else 0.0
        out = wt.reweight(G, weight, flags, og, weighting, f2).cpu().numpy()
        ref = ow.reweight(uvw, freq, weight, flags, weight, vis_to_im, ogrid, *cd,
                          weighting=weighting, robustness=rob, sumwt=osumwt)
        assert np.array_equal(out, ref), (weighting, rob)
    tg = wt.taper(G, weight, flags, "gaussian", 4 * cell).cpu().numpy()
    np.testing.assert_allclose(tg, ow.taper_gaussian(uvw, freq, flags, weight, 4 * cell),
                               rtol=1e-15, atol=0)  # exp: <= 1 ulp each side


def test_edges(wt):
    """Empty input, everything flagged, everything off the grid, uniform weights of a lone sample."""
    import torch
    npix, cell = 32, 1e-3
    cd = (-1.0 / (npix * cell), 1.0 / (npix * cell), npix // 2 + 1, npix // 2 + 1)
    freq = np.array([1e8])
    G = wt.WeightGrid(np.zeros((0, 3)), freq, [0], 1, npix, npix, *cd)
    grid, sumwt, sumsq = wt.grid_weights(G, np.zeros((0, 1, 1)), None)
    assert float(grid.abs().sum()) == 0 and sumwt[0, 0] == 0 and sumsq == 0
    uvw = np.array([[300.0, 600.0, 0.0], [1e9, 0.0, 0.0], [0.0, 0.0, 0.0]])
    G = wt.WeightGrid(uvw, freq, [0], 1, npix, npix, *cd)
    w = np.full((3, 1, 1), 2.5)
    grid, sumwt, _ = wt.grid_weights(G, w, np.ones((3, 1, 1), dtype=np.uint8))
    assert float(grid.abs().sum()) == 0 and sumwt[0, 0] == 0
    grid, sumwt, _ = wt.grid_weights(G, w, None)
    gh = grid.cpu().numpy()[0, 0]
    assert sumwt[0, 0] == 10.0                                    # two in-grid samples, twice each
    assert gh[npix // 2, npix // 2] == 5.0                        # the auto hits the centre twice
    assert np.count_nonzero(gh) == 3
    out = wt.reweight(G, w, None, grid, "uniform").cpu().numpy().ravel()
    assert out[0] == 1.0 and out[1] == 0.0 and out[2] == 0.5      # lone sample -> 1; off grid -> 0
    assert isinstance(grid, torch.Tensor) and grid.is_cuda
