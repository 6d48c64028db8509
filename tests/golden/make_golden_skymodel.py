#!/usr/bin/env python
"""Golden vectors for the major-cycle row from THE REFERENCE'S OWN
skymodel_predict_calibrate / skymodel_calibrate_invert / subtract_visibility
(/root/reference/src/ska_sdp_func_python/sky_model/skymodel_imaging.py:70-235,
visibility/operations.py:86-105), run in the build container:
    python tests/golden/make_golden_skymodel.py
Substitutions as in make_golden.py: ducc0.wgridder -> exact direct DFT (oracle/dft.py),
datamodels -> minimal_models (incl. SkyModel, groupby("time") slices, concatenate_visibility)."""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402
from ska_sdp_func_python_b200 import minimal_models as mm  # noqa: E402


class _GroupableVis(mm.Visibility):
    def groupby(self, key, squeeze=False):
        assert key == "time"
        return [(t, _as_groupable(v)) for t, v in mm.time_slices(self)]

    def copy(self, deep=True, zero=False):
        return _as_groupable(super().copy(deep=deep, zero=zero))


def _as_groupable(v):
    v.__class__ = _GroupableVis
    return v


def main():
    m = types.ModuleType("ska_sdp_datamodels.visibility.vis_utils")
    m.concatenate_visibility = lambda vl, dim="time": _as_groupable(mm.concatenate_visibility(vl, dim))
    sys.modules["ska_sdp_datamodels.visibility.vis_utils"] = m
    mg.load_reference()
    import ska_sdp_func_python.visibility as ref_visibility
    ref_visibility.concatenate_visibility = m.concatenate_visibility
    import ska_sdp_func_python.sky_model.skymodel_imaging as ref_sm
    ref_sm.concatenate_visibility = m.concatenate_visibility
    from ska_sdp_func_python.visibility.operations import subtract_visibility

    def no_components(comps, beam):
        # apply_beam_to_skycomponent needs astropy; the masked / primary-beam cases below carry
        # no components, for which it returns the empty list
        assert len(comps) == 0
        return comps

    ref_sm.apply_beam_to_skycomponent = no_components

    rng = np.random.default_rng(31)
    npix, cell = 32, 2.0e-3
    pc = mm.SkyCoord.from_deg(30.0, -50.0)
    vis = _as_groupable(mg.make_vis(rng, 3, 30, 3, "stokesI", npix, cell, pc, wfrac=0.2))
    freq = vis["frequency"].data
    im = mm.create_image(npix, cell, pc, frequency=[freq[0]], nchan=1,
                         channel_bandwidth=4.0 * (freq[-1] - freq[0] + 1e6),
                         polarisation_frame=mm.PolarisationFrame("stokesI"))
    model = im.copy(deep=True)
    model["pixels"].data[0, 0, rng.integers(4, 28, 8), rng.integers(4, 28, 8)] = rng.uniform(0.5, 2, 8)
    comps = [mm.SkyComponent(mm.SkyCoord.from_deg(30.0 + rng.uniform(-1, 1), -50.0 + rng.uniform(-1, 1)),
                             freq, rng.uniform(0.5, 2.0, (3, 1)), mm.PolarisationFrame("stokesI"))
             for _ in range(3)]
    mask = im.copy(deep=True)
    mask["pixels"].data[...] = 1.0
    mask["pixels"].data[..., :6] = 0.0
    out = {}
    sm = mm.SkyModel(image=model, components=comps)
    pred = ref_sm.skymodel_predict_calibrate(vis, sm, context="ng", epsilon=1e-12)
    out["pred"] = pred["vis"].data
    dirty, sumwt = ref_sm.skymodel_calibrate_invert(vis, sm, context="ng", epsilon=1e-12)
    out["dirty"], out["sumwt"] = dirty["pixels"].data, sumwt
    res_vis = subtract_visibility(vis, pred)
    rdirty, rsumwt = ref_sm.skymodel_calibrate_invert(res_vis, sm, context="ng", epsilon=1e-12)
    out["res_dirty"], out["res_sumwt"] = rdirty["pixels"].data, rsumwt
    # image-only model with a mask (components with a mask need astropy in the reference)
    smm = mm.SkyModel(image=model, components=[], mask=mask)
    predm = ref_sm.skymodel_predict_calibrate(vis, smm, context="ng", epsilon=1e-12)
    out["pred_mask"] = predm["vis"].data
    rm, _ = ref_sm.skymodel_calibrate_invert(subtract_visibility(vis, predm), smm, context="ng",
                                             epsilon=1e-12)
    out["res_dirty_mask"] = rm["pixels"].data
    # per-time-slice primary-beam loop
    yy, xx = np.mgrid[:npix, :npix]
    pbpix = np.exp(-((xx - npix // 2) ** 2 + (yy - npix // 2) ** 2) / (2 * 9.0 ** 2))
    pb = im.copy(deep=True)
    pb["pixels"].data[0, 0] = pbpix

    def get_pb(v, image):
        return pb

    smi = mm.SkyModel(image=model, components=[])
    out["pred_pb"] = ref_sm.skymodel_predict_calibrate(vis, smi, context="ng", get_pb=get_pb,
                                                      epsilon=1e-12)["vis"].data
    for flat in (False, True):
        d, f = ref_sm.skymodel_calibrate_invert(vis, smi, context="ng", get_pb=get_pb,
                                                flat_sky=flat, epsilon=1e-12)
        out[f"dirty_pb_flat{int(flat)}"] = d["pixels"].data
        out[f"sens_pb_flat{int(flat)}"] = f["pixels"].data
    w = im.image_acc.wcs.wcs
    np.savez_compressed(
        os.path.join(HERE, "skymodel_case.npz"), vis=vis["vis"].data, uvw=vis["uvw"].data, freq=freq,
        weight=vis["weight"].data, imaging_weight=vis["imaging_weight"].data, flags=vis["flags"].data,
        pc=np.array([pc.ra.rad, pc.dec.rad]), crval=w.crval, crpix=w.crpix, cdelt=w.cdelt,
        model=model["pixels"].data, mask=mask["pixels"].data, pb=pb["pixels"].data,
        comp_dir=np.array([[c.direction.ra.rad, c.direction.dec.rad] for c in comps]),
        comp_flux=np.array([c.flux for c in comps]), **out)
    print("wrote skymodel_case.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
