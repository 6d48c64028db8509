"""The Hogbom and multi-scale CLEAN oracles (oracle/clean.py) against outputs of the reference's own
hogbom() / msclean() (tests/golden/make_golden_clean.py, make_golden_msclean.py).  CPU only,
bit-exact."""
import os

import numpy as np
import pytest

from oracle import clean as oc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "clean_case.npz")


@pytest.mark.parametrize("name", ["a", "b", "c", "d"])
def test_hogbom_bit_exact(name):
    g = np.load(GOLD)
    gain, thresh, niter, frac = g[f"{name}_par"]
    window = g[f"{name}_window"] if f"{name}_window" in g.files else None
    comps, res, it = oc.hogbom(g[f"{name}_dirty"], g[f"{name}_psf"], window, gain, thresh, int(niter), frac)
    assert np.array_equal(comps, g[f"{name}_comps"])
    assert np.array_equal(res, g[f"{name}_res"])
    assert 1 <= it <= int(niter)


MSGOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "msclean_case.npz")


def ms_case(g, name):
    gain, thresh, niter, frac = g[f"{name}_par"]
    window = g[f"{name}_window"] if f"{name}_window" in g.files else None
    sens = g[f"{name}_sens"] if f"{name}_sens" in g.files else None
    return (g[f"{name}_dirty"], g[f"{name}_psf"], window, sens, gain, thresh, int(niter),
            list(g[f"{name}_scales"]), frac)


@pytest.mark.parametrize("name", ["a", "b", "c", "d", "e"])
def test_msclean_bit_exact(name):
    g = np.load(MSGOLD)
    args = ms_case(g, name)
    comps, res, done = oc.msclean(*args)
    assert np.array_equal(comps, g[f"{name}_comps"])
    assert np.array_equal(res, g[f"{name}_res"])
    assert (done < args[6]) == (name == "e")            # only case e stops at the threshold


def test_scale_stack_properties():
    st = oc.scale_stack((3, 33, 40), [0, 4, 9])
    assert st[0, 17, 20] == 1.0 and st[0].sum() == 1.0   # delta at ceil(n / 2)
    assert np.allclose(st.sum(axis=(1, 2)), 1.0, rtol=1e-15) and (st >= 0).all()
    assert np.unravel_index(st[2].argmax(), st[2].shape) == (17, 20)
    assert oc.pswf(1.5) == 0.0 and oc.pswf(-0.1) == 0.0 and oc.pswf(0.0) > oc.pswf(0.9) > 0.0


MFGOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "msmfsclean_case.npz")


@pytest.mark.parametrize("name", ["a", "b", "c", "d", "e", "f"])
def test_msmfsclean_bit_exact(name):
    g = np.load(MFGOLD)
    gain, thresh, niter, frac = g[f"{name}_par"]
    window = g[f"{name}_window"] if f"{name}_window" in g.files else None
    sens = g[f"{name}_sens"] if f"{name}_sens" in g.files else None
    model, res, done = oc.msmfsclean(g[f"{name}_dirty"], g[f"{name}_psf"], window, sens, gain, thresh,
                                     int(niter), list(g[f"{name}_scales"]), frac,
                                     findpeak="CASA" if name in "ef" else "RASCIL")
    assert np.array_equal(model, g[f"{name}_model"])
    assert np.array_equal(res, g[f"{name}_res"])
    assert (done < int(niter)) == (name == "d")         # only case d stops at the threshold


CGOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "clean_complex_case.npz")


@pytest.mark.parametrize("name", ["a", "b", "c", "d"])
def test_hogbom_complex_bit_exact(name):
    g = np.load(CGOLD)
    gain, thresh, niter, frac = g[f"{name}_par"]
    window = g[f"{name}_window"] if f"{name}_window" in g.files else None
    cq, cu, rq, ru, it = oc.hogbom_complex(g[f"{name}_dq"], g[f"{name}_du"], g[f"{name}_psf"], g[f"{name}_psf"],
                                           window, gain, thresh, int(niter), frac)
    for got, key in ((cq, "cq"), (cu, "cu"), (rq, "rq"), (ru, "ru")):
        assert np.array_equal(got, g[f"{name}_{key}"])
    assert (it < int(niter)) == (name == "b")
