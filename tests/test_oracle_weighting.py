"""The weighting oracle (oracle/weighting.py) against golden vectors produced by the reference's
own grid_visibility_weight_to_griddata / griddata_visibility_reweight / weight_visibility /
taper_visibility_* (tests/golden/make_golden_weight.py).  CPU only."""
import os

import numpy as np
import pytest

from oracle import weighting as ow

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["weight_case_a.npz", "weight_case_b.npz"]


def load(name):
    g = np.load(os.path.join(GOLD, name))
    nt, nbl, nchan, npol = g["weight"].shape
    d = dict(g=g, uvw=g["uvw"].reshape(-1, 3), freq=g["freq"],
             weight=g["weight"].reshape(-1, nchan, npol), flags=g["flags"].reshape(-1, nchan, npol),
             imw=g["imaging_weight"].reshape(-1, nchan, npol), shape=(nt, nbl, nchan, npol))
    nimg = int(g["img_nchan"])
    d["vis_to_im"] = np.round((g["freq"] - g["img_crval"][3]) / g["img_cdelt"][3]
                              + g["img_crpix"][3] - 1).astype(int)
    d["geo"] = dict(cdelt_u=g["grid_cdelt"][0], cdelt_v=g["grid_cdelt"][1],
                    crpix_u=g["grid_crpix"][0], crpix_v=g["grid_crpix"][1])
    d["nimg"], d["npix"] = nimg, int(g["npix"])
    return d


@pytest.mark.parametrize("name", CASES)
def test_grid_weights_bit_exact(name):
    d = load(name)
    grid, sumwt = ow.grid_weights(d["uvw"], d["freq"], d["weight"], d["flags"], d["vis_to_im"],
                                  d["nimg"], d["npix"], d["npix"], **d["geo"])
    assert np.array_equal(grid, d["g"]["grid"])
    assert np.array_equal(sumwt, d["g"]["sumwt"])
    assert grid.max() > 10  # the clustered core really piles samples into cells


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("key,kw", [
    ("imw_uniform", dict(weighting="uniform")),
    ("imw_robust0", dict(weighting="robust", robustness=0.0, use_sumwt=True)),
    ("imw_robustm1", dict(weighting="robust", robustness=-1.0, use_sumwt=True)),
    ("imw_robust_nosumwt", dict(weighting="robust", robustness=0.5)),
    ("imw_natural", dict(weighting="natural")),
])
def test_reweight_bit_exact(name, key, kw):
    d = load(name)
    kw = dict(kw)
    sumwt = d["g"]["sumwt"] if kw.pop("use_sumwt", False) else None
    out = ow.reweight(d["uvw"], d["freq"], d["weight"], d["flags"], d["imw"], d["vis_to_im"],
                      d["g"]["grid"], sumwt=sumwt, **d["geo"], **kw)
    assert np.array_equal(out.reshape(d["shape"]), d["g"][key])


@pytest.mark.parametrize("name", CASES)
def test_weight_visibility_goldens_are_consistent(name):
    """weight_visibility = grid + reweight with sumwt (weighting.py:55-63)."""
    d = load(name)
    g = d["g"]
    assert np.array_equal(g["wv_uniform"], g["imw_uniform"])
    assert np.array_equal(g["wv_natural"], g["weight"])
    out = ow.reweight(d["uvw"], d["freq"], d["weight"], d["flags"], d["imw"], d["vis_to_im"],
                      g["grid"], weighting="robust", robustness=-0.5, sumwt=g["sumwt"], **d["geo"])
    assert np.array_equal(out.reshape(d["shape"]), g["wv_robust"])


@pytest.mark.parametrize("name", CASES)
def test_tapers_bit_exact(name):
    d = load(name)
    g = d["g"]
    uvw = np.nan_to_num(d["uvw"]) if False else d["uvw"]
    tg = ow.taper_gaussian(uvw, d["freq"], d["flags"], d["imw"], float(g["beam"]))
    assert np.array_equal(tg.reshape(d["shape"]), g["taper_gaussian"], equal_nan=True)
    tt = ow.taper_tukey(uvw, d["freq"], d["flags"], d["imw"], float(g["tukey"]))
    assert np.array_equal(tt.reshape(d["shape"]), g["taper_tukey"], equal_nan=True)
