"""The benchmark workload at its FULL size (BASELINE.json configs[1] "C2": SKA-LOW full array,
30 times x 131 328 baselines x 64 channels = 2.52e8 samples, npixel 2048, fp32, epsilon 1e-5)
through size-independent properties, everything resident on the device:

  * predict(point-source model) == the analytic visibilities of those sources, which the component
    DFT kernel (verified on its own against the oracle, test_gpu_dft.py) evaluates for all samples;
  * invert of those visibilities equals the exact DFT at sampled pixels (10 epsilon of the peak);
  * adjointness <invert(ms, wgt), image> == <wgt ms, predict(image)>;
  * uniform weighting (row N1) of the same samples followed by a dopsf-style invert gives a PSF with
    unit peak after normalisation by sumwt, and the gridded weights sum to 2 x the in-grid weight.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
C = 299792458.0


@pytest.fixture(scope="module")
def job(built):
    import torch
    assert torch.cuda.is_available()
    from ska_sdp_func_python_b200 import simulate, wgridder
    obs = simulate.observation("C2")
    uvw = simulate.flip_for_gridder(obs["uvw"])
    dev = torch.device("cuda")
    uvw_d = torch.from_numpy(uvw).to(dev)
    plan = wgridder.Plan(uvw_d, obs["freq"], obs["npix"], obs["npix"], obs["cellsize"],
                         obs["cellsize"], 1e-5, True, precision="single")
    yield dict(obs=obs, uvw=uvw, uvw_d=uvw_d, plan=plan, dev=dev, torch=torch)
    plan.close()


def test_point_sources_round_trip(job):
    from ska_sdp_func_python_b200 import dft
    torch, dev, obs, plan = job["torch"], job["dev"], job["obs"], job["plan"]
    npix, cell, freq = obs["npix"], obs["cellsize"], obs["freq"]
    rng = np.random.default_rng(77)
    nsrc = 24
    ix = rng.integers(npix // 8, npix - npix // 8, nsrc)
    iy = rng.integers(npix // 8, npix - npix // 8, nsrc)
    ix[0], iy[0] = npix // 2, npix // 2
    ix[1], iy[1] = 40, npix - 30                          # near a corner: largest w term
    flux = rng.uniform(0.5, 2.0, nsrc)
    model = np.zeros((npix, npix), dtype=np.float32)
    np.add.at(model, (ix, iy), flux.astype(np.float32))
    x, y = (ix - npix // 2) * cell, (iy - npix // 2) * cell
    r2 = x * x + y * y
    nm1 = -r2 / (np.sqrt(1.0 - r2) + 1.0)
    # gridder convention: exp(-2 pi i (u x + v y - w (n - 1))) * flux / n
    dc = np.stack([x, y, -nm1], axis=1)
    fl = np.ascontiguousarray(np.broadcast_to((flux / (nm1 + 1.0))[:, None, None], (nsrc, 1, 1))).astype(complex)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    exact = dft.dft_predict_device(job["uvw_d"], t(freq / C), t(dc), t(fl), len(freq))[..., 0]
    assert exact.shape == (job["uvw"].shape[0], len(freq))
    pred = plan.predict(t(model))
    num = torch.linalg.vector_norm((pred.to(torch.complex128) - exact).reshape(-1))
    den = torch.linalg.vector_norm(exact.reshape(-1))
    assert float(num / den) < 1e-5                          # relative L2 <= epsilon, all 2.5e8 samples
    wgt = torch.ones(exact.shape, device=dev, dtype=torch.float32)
    wgt[t(obs["auto"])] = 0
    dirty = plan.invert(exact.to(torch.complex64), wgt)
    sumwt = float(wgt.sum(dtype=torch.float64))
    d = (dirty.to(torch.float64) / sumwt).cpu().numpy()
    # the dirty image at sampled pixels (the two special sources, two corners, random ones)
    # against the exact DFT over all 2.5e8 samples: north_star's bound, 10 epsilon of the peak
    from oracle import dft as odft
    px = np.concatenate([ix[:2], [0, npix - 1], rng.integers(0, npix, 4)])
    py = np.concatenate([iy[:2], [0, npix - 1], rng.integers(0, npix, 4)])
    dref = odft.ms2dirty_points(job["uvw"], freq, exact.cpu().numpy(), wgt.cpu().numpy().astype(np.float64),
                                npix, npix, cell, cell, True, px, py) / sumwt
    peak = float(np.abs(d).max())
    assert np.abs(d[px, py] - dref).max() < 10 * 1e-5 * peak
    assert abs(d[ix[0], iy[0]] - dref[0]) < 10 * 1e-5 * peak
    # adjointness on the same data
    img = torch.randn((npix, npix), device=dev, dtype=torch.float32)
    ms = torch.view_as_complex(torch.randn(exact.shape + (2,), device=dev, dtype=torch.float32))
    lhs = float((plan.invert(ms, wgt).to(torch.float64) * img.to(torch.float64)).sum())
    pv = plan.predict(img)
    rhs = float(((ms.to(torch.complex128) * wgt) * pv.to(torch.complex128).conj()).real.sum())
    assert abs(lhs - rhs) < 2e-5 * max(abs(lhs), abs(rhs), float(torch.linalg.vector_norm(ms.reshape(-1))) *
                                       float(torch.linalg.vector_norm(pv.reshape(-1))))


def test_uniform_weighting_then_psf(job):
    from ska_sdp_func_python_b200 import weighting as wt
    torch, dev, obs, plan = job["torch"], job["dev"], job["obs"], job["plan"]
    npix, cell, freq = obs["npix"], obs["cellsize"], obs["freq"]
    nrow, nchan = job["uvw"].shape[0], len(freq)
    # weighting works on the UNFLIPPED uvw of the Visibility; the weight grid is conjugate-symmetric
    cd = (-1.0 / (npix * cell), 1.0 / (npix * cell), npix // 2 + 1, npix // 2 + 1)
    G = wt.WeightGrid(obs["uvw"], freq, np.zeros(nchan, dtype=np.int32), 1, npix, npix, *cd)
    w = torch.ones((nrow, nchan, 1), device=dev, dtype=torch.float64)
    w[torch.from_numpy(obs["auto"]).to(dev)] = 0
    grid, sumwt, sumsq = wt.grid_weights(G, w, None)
    nz = float(w.sum())
    assert sumwt[0, 0] == 2.0 * nz                          # integers: exact, nothing off the grid
    assert float(grid.sum()) == 2.0 * nz
    imw = wt.reweight(G, w, None, grid, "uniform")
    assert float(imw.max()) <= 1.0 and float(imw.min()) == 0.0
    # every occupied cell ends up with total uniform weight 1 from its direct hits + conjugate hits
    # of other cells: sum of uniform weights = sum over cells of (direct hits / gridded weight)
    assert 0 < float(imw.sum()) < nz
    wg32 = imw[..., 0].to(torch.float32).contiguous()
    ones = torch.ones((nrow, nchan), device=dev, dtype=torch.complex64)
    psf = plan.invert(ones, wg32)
    s = float(wg32.sum(dtype=torch.float64))
    assert abs(float(psf[npix // 2, npix // 2]) / s - 1.0) < 1e-5   # PSF peak == 1 after normalisation
    assert float(psf.max()) == float(psf[npix // 2, npix // 2])
