"""Major-cycle glue (row N2) on the GPU: Plan.residual / residual_ng / skymodel_* against
(1) golden vectors from the reference's own skymodel_predict_calibrate /
skymodel_calibrate_invert / subtract_visibility (tests/golden/make_golden_skymodel.py; ducc0
replaced by the exact DFT there) and (2) the identity residual == invert(ms - predict(model)) on
the same plan.  Tolerance: the gridder's epsilon (1e-12 requested -> < 1e-10 of the peak, the
north_star fp64 bound); fp32: 10 * epsilon."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, make_case, rel_max

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods(built):
    import torch
    assert torch.cuda.is_available()
    from ska_sdp_func_python_b200 import imaging, skymodel_imaging, wgridder
    return imaging, skymodel_imaging, wgridder


def _load():
    from ska_sdp_func_python_b200 import minimal_models as mm
    g = np.load(os.path.join(GOLDEN, "skymodel_case.npz"))
    pc = mm.SkyCoord(float(g["pc"][0]), float(g["pc"][1]))
    pf = mm.PolarisationFrame("stokesI")
    vis = mm.Visibility(g["vis"], g["uvw"], g["freq"], pc, pf, weight=g["weight"],
                        imaging_weight=g["imaging_weight"], flags=g["flags"])
    wcs = mm.WCS(g["crval"], g["crpix"], g["cdelt"])
    model = mm.Image(g["model"].copy(), wcs, pf)
    mask = mm.Image(g["mask"].copy(), wcs, pf)
    pb = mm.Image(g["pb"].copy(), wcs, pf)
    comps = [mm.SkyComponent(mm.SkyCoord(d[0], d[1]), g["freq"], f, pf)
             for d, f in zip(g["comp_dir"], g["comp_flux"])]
    return g, mm, vis, model, mask, pb, comps


def test_skymodel_predict_and_invert_golden(mods):
    _, sm, _ = mods
    g, mm, vis, model, mask, pb, comps = _load()
    sky = mm.SkyModel(image=model, components=comps)
    v0 = vis["vis"].data.copy()
    pred = sm.skymodel_predict_calibrate(vis, sky, context="ng", epsilon=1e-12)
    assert rel_max(pred["vis"].data, g["pred"]) < 1e-10
    assert np.array_equal(vis["vis"].data, v0)
    dirty, sumwt = sm.skymodel_calibrate_invert(vis, sky, context="ng", epsilon=1e-12)
    assert np.array_equal(sumwt, g["sumwt"])
    assert rel_max(dirty["pixels"].data, g["dirty"]) < 1e-10
    # masked image-only model
    skym = mm.SkyModel(image=model, components=[], mask=mask)
    pm = sm.skymodel_predict_calibrate(vis, skym, context="ng", epsilon=1e-12)
    assert rel_max(pm["vis"].data, g["pred_mask"]) < 1e-10
    with pytest.raises(ValueError):
        sm.skymodel_calibrate_invert(vis, mm.SkyModel(image=None))
    with pytest.raises(NotImplementedError):
        sm.skymodel_predict_calibrate(vis, mm.SkyModel(image=model, gaintable=object()), docal=True)


def test_skymodel_residual_golden(mods):
    imaging, sm, _ = mods
    g, mm, vis, model, mask, pb, comps = _load()
    peak = np.abs(g["dirty"]).max()
    res, sumwt = sm.skymodel_residual(vis, mm.SkyModel(image=model, components=comps), epsilon=1e-12)
    assert np.array_equal(sumwt, g["res_sumwt"])
    assert np.abs(res["pixels"].data - g["res_dirty"]).max() < 1e-10 * peak
    resm, _ = sm.skymodel_residual(vis, mm.SkyModel(image=model, components=[], mask=mask),
                                   epsilon=1e-12)
    assert np.abs(resm["pixels"].data - g["res_dirty_mask"]).max() < 1e-10 * peak
    # residual of a zero model is the dirty image; of zero data, minus the model's dirty image
    zero = model.copy(deep=True, zero=True)
    d0, _ = imaging.residual_ng(vis, zero, epsilon=1e-12)
    d1, _ = imaging.invert_ng(vis, zero, epsilon=1e-12)
    assert np.abs(d0["pixels"].data - d1["pixels"].data).max() < 1e-12 * peak


def test_skymodel_primary_beam_loop_golden(mods):
    _, sm, _ = mods
    g, mm, vis, model, mask, pb, comps = _load()
    sky = mm.SkyModel(image=model, components=[])
    calls = []

    def get_pb(v, image):
        calls.append(v["vis"].data.shape)
        return pb

    pred = sm.skymodel_predict_calibrate(vis, sky, context="ng", get_pb=get_pb, epsilon=1e-12)
    assert calls == [(1, 30, 3, 1)] * 3            # one call per time slice
    assert rel_max(pred["vis"].data, g["pred_pb"]) < 1e-10
    for flat in (False, True):
        d, f = sm.skymodel_calibrate_invert(vis, sky, context="ng", get_pb=get_pb, flat_sky=flat,
                                            epsilon=1e-12)
        assert rel_max(d["pixels"].data, g[f"dirty_pb_flat{int(flat)}"]) < 1e-10
        assert rel_max(f["pixels"].data, g[f"sens_pb_flat{int(flat)}"]) < 1e-12


# tol: the two paths grid the same samples and differ only by the order of the atomics, whose
# rounding noise the gridding correction may amplify up to epsilon / 4 at the image rim
@pytest.mark.parametrize("prec,eps,algo,tol", [("double", 1e-9, -1, 2.5e-10), ("double", 1e-12, -1, 1e-12),
                                               ("single", 1e-5, -1, 2.5e-6), ("double", 1e-9, 0, 2.5e-10)])
def test_plan_residual_identity(mods, prec, eps, algo, tol):
    """residual(model, ms, wgt) == invert(ms - predict(model), wgt) on the same plan, device
    and host arrays, with and without the residual visibilities returned / updated in place."""
    import torch
    _, _, wg = mods
    npix, cell = 96, 1.5e-3
    uvw, freq, ms, wgt = make_case(17, 4000, 6, npix, npix, cell, cell, 0.3)
    rng = np.random.default_rng(2)
    model = np.zeros((npix, npix))
    model[rng.integers(8, 88, 20), rng.integers(8, 88, 20)] = rng.uniform(0.5, 2, 20)
    cdt = np.complex64 if prec == "single" else np.complex128
    rdt = np.float32 if prec == "single" else np.float64
    ms, wgt, model = ms.astype(cdt), wgt.astype(rdt), model.astype(rdt)
    with wg.Plan(uvw, freq, npix, npix, cell, cell, eps, True, precision=prec, algo=algo) as plan:
        pred = plan.predict(model)
        ref = plan.invert(ms - pred, wgt)
        scale = np.abs(ref).max()
        r_host = plan.residual(model, ms, wgt)
        assert np.abs(r_host - ref).max() < tol * scale
        resid = np.empty_like(ms)
        r2 = plan.residual(model, ms, wgt, ms_resid=resid)
        assert np.abs(r2 - ref).max() < tol * scale
        assert np.abs(resid - (ms - pred)).max() < tol * np.abs(ms).max()
        dev = torch.device("cuda")
        ms_d, wgt_d, model_d = (torch.from_numpy(a).to(dev) for a in (ms, wgt, model))
        r_dev = plan.residual(model_d, ms_d, wgt_d, ms_resid=ms_d)      # in place
        assert np.abs(r_dev.cpu().numpy() - ref).max() < tol * scale
        assert np.abs(ms_d.cpu().numpy() - (ms - pred)).max() < tol * np.abs(ms).max()
        assert plan.info()["launches_last_call"] >= 4


def test_components_with_mask_use_builtin_apply_beam(mods):
    """Components combined with a mask (skymodel_imaging.py:22-45): fluxes scaled by the mask value at
    their pixel through sky_component.apply_beam_to_skycomponent, then the DFT; equal to doing the two
    steps by hand."""
    imaging, sm, _ = mods
    from ska_sdp_func_python_b200 import dft
    from ska_sdp_func_python_b200.sky_component import apply_beam_to_skycomponent
    g, mm, vis, model, mask, pb, comps = _load()
    mask["pixels"].data[:] = np.random.default_rng(1).uniform(0.3, 1.0, mask["pixels"].data.shape)
    sky = mm.SkyModel(image=model, components=comps, mask=mask)
    got = sm.skymodel_predict_calibrate(vis, sky, context="ng", epsilon=1e-12)
    zero = vis.copy(deep=True, zero=True)
    want = dft.dft_skycomponent_visibility(zero, apply_beam_to_skycomponent(comps, mask))
    masked = model.copy(deep=True)
    masked["pixels"].data *= mask["pixels"].data
    want["vis"].data += imaging.predict_ng(vis.copy(deep=True, zero=True), masked, epsilon=1e-12)["vis"].data
    assert rel_max(got["vis"].data, want["vis"].data) < 1e-10
    scaled = apply_beam_to_skycomponent(comps, mask)
    assert any(not np.array_equal(a.flux, b.flux) for a, b in zip(scaled, comps))
