#!/usr/bin/env python
"""Golden vectors for the density-weighting row, produced by running
THE REFERENCE'S OWN CODE in the build container:

    python tests/golden/make_golden_weight.py

Imported unmodified from /root/reference/src:
  * ska_sdp_func_python.grid_data.gridding.grid_visibility_weight_to_griddata  (gridding.py:258-334)
  * ska_sdp_func_python.grid_data.gridding.griddata_visibility_reweight        (gridding.py:362-499)
  * ska_sdp_func_python.imaging.weighting.weight_visibility, taper_visibility_gaussian,
    taper_visibility_tukey                                                      (weighting.py:35-140)
Substituted (not installable here, see make_golden.py): ska_sdp_datamodels -> the containers of
ska_sdp_func_python_b200/minimal_models.py (Visibility, Image, GridData,
create_griddata_from_image restated from the published datamodels behaviour); astropy WCS -> the
linear WCS stand-in of the same module.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402
from ska_sdp_func_python_b200 import minimal_models as mm  # noqa: E402


def load_weighting():
    # seeded BEFORE the reference package is imported (its imaging/__init__ pulls weighting.py in)
    def seed(name, **attrs):
        m = types.ModuleType(name)
        m.__path__ = []
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m

    seed("ska_sdp_datamodels.gridded_visibility")
    seed("ska_sdp_datamodels.gridded_visibility.grid_vis_model", GridData=mm.GridData)
    seed("ska_sdp_datamodels.gridded_visibility.grid_vis_create",
         create_griddata_from_image=mm.create_griddata_from_image)
    mg.load_reference()
    import ska_sdp_func_python.grid_data.gridding as ref_gridding
    import ska_sdp_func_python.imaging.weighting as ref_weighting
    return ref_gridding, ref_weighting


def make_case(seed, pf, nchan, img_nchan, npix, cell, ntimes=3, nbl=70):
    rng = np.random.default_rng(seed)
    pc = mm.SkyCoord.from_deg(180.0, -60.0)
    vis = mg.make_vis(rng, ntimes, nbl, nchan, pf, npix, cell, pc, flag_fraction=0.1)
    # cluster a third of the baselines in the core (many samples per cell), push a few rows
    # beyond the grid edge, one NaN coordinate
    vis["uvw"].data[:, : nbl // 3] *= 0.05
    vis["uvw"].data[1, -3:] *= 1.6
    vis["uvw"].data[2, 5, 0] = np.nan
    freq = vis["frequency"].data
    if img_nchan == 1:
        im = mm.create_image(npix, cell, pc, frequency=[freq[0]], nchan=1,
                             channel_bandwidth=4.0 * (freq[-1] - freq[0] + 1e6),
                             polarisation_frame=mm.PolarisationFrame(pf))
    else:
        im = mm.create_image(npix, cell, pc, frequency=freq, nchan=nchan,
                             channel_bandwidth=freq[1] - freq[0],
                             polarisation_frame=mm.PolarisationFrame(pf))
    return vis, im


def run_case(path, ref_gridding, ref_weighting, seed, pf, nchan, img_nchan, npix, cell):
    vis, im = make_case(seed, pf, nchan, img_nchan, npix, cell)
    out = {}
    gd = mm.create_griddata_from_image(im, polarisation_frame=mm.PolarisationFrame(pf))
    w = gd.griddata_acc.griddata_wcs.wcs
    gd, sumwt = ref_gridding.grid_visibility_weight_to_griddata(vis.copy(deep=True), gd)
    out["grid"] = np.real(gd["pixels"].data)
    out["sumwt"] = sumwt
    for name, kw in (("uniform", dict(weighting="uniform")),
                     ("robust0", dict(weighting="robust", robustness=0.0, sumwt=sumwt)),
                     ("robustm1", dict(weighting="robust", robustness=-1.0, sumwt=sumwt)),
                     ("robust_nosumwt", dict(weighting="robust", robustness=0.5)),
                     ("natural", dict(weighting="natural"))):
        v = ref_gridding.griddata_visibility_reweight(vis.copy(deep=True), gd, **kw)
        out["imw_" + name] = v["imaging_weight"].data
    for name, kw in (("uniform", dict(weighting="uniform")),
                     ("robust", dict(weighting="robust", robustness=-0.5)),
                     ("natural", dict(weighting="natural"))):
        v = ref_weighting.weight_visibility(vis.copy(deep=True), im, **kw)
        out["wv_" + name] = v["imaging_weight"].data
    beam = 3.0 * cell
    out["beam"] = beam
    out["taper_gaussian"] = ref_weighting.taper_visibility_gaussian(
        vis.copy(deep=True), beam=beam)["imaging_weight"].data
    out["tukey"] = 0.25
    out["taper_tukey"] = ref_weighting.taper_visibility_tukey(
        vis.copy(deep=True), tukey=0.25)["imaging_weight"].data
    iw = im.image_acc.wcs.wcs
    np.savez_compressed(
        path, uvw=vis["uvw"].data, freq=vis["frequency"].data, weight=vis["weight"].data,
        imaging_weight=vis["imaging_weight"].data, flags=vis["flags"].data, pf=pf, npix=npix,
        cell=cell, img_nchan=img_nchan, img_crval=iw.crval, img_crpix=iw.crpix, img_cdelt=iw.cdelt,
        grid_cdelt=w.cdelt, grid_crpix=w.crpix, **out)
    print(path, "grid max", out["grid"].max(), "sumwt", sumwt.ravel()[:4],
          "uniform mean", out["imw_uniform"].mean())


def main():
    ref_gridding, ref_weighting = load_weighting()
    run_case(os.path.join(HERE, "weight_case_a.npz"), ref_gridding, ref_weighting, 11, "stokesI",
             nchan=3, img_nchan=1, npix=64, cell=1e-3)
    run_case(os.path.join(HERE, "weight_case_b.npz"), ref_gridding, ref_weighting, 12, "stokesIQ",
             nchan=4, img_nchan=4, npix=48, cell=2e-3)


if __name__ == "__main__":
    main()
