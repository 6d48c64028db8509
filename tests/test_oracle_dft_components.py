"""Component-DFT oracle against outputs of the reference's own dft_cpu_looped /
dft_skycomponent_visibility (tests/golden/make_golden_dft.py).  CPU only.
Tolerance: the reference evaluates exp(-2 pi i phase) at |phase| up to ~100 turns without range
reduction (argument error ~ 600 rad * 1.1e-16 = 7e-14 per term); 1e-12 of the flux scale."""
import os

import numpy as np

from oracle import dft as odft

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dft_comp_case.npz")
C = 299792458.0


def test_array_level():
    g = np.load(GOLD)
    lam = np.einsum("tbs,k->tbks", g["uvw"], g["freq"] / C)
    v = odft.dft_components(g["dc"], lam, g["vfl"])
    scale = np.abs(g["vfl"]).sum(axis=0).max()
    assert np.abs(v - g["vis_a"]).max() < 1e-12 * scale
    v1 = odft.dft_components(g["dc"], lam, g["vfl"][:, :1, :1])
    assert v1.shape == g["vis_a1"].shape == (3, 40, 5, 1)
    assert np.abs(v1 - g["vis_a1"]).max() < 1e-12 * scale


def test_extracted_components():
    g = np.load(GOLD)
    lam = np.einsum("tbs,k->tbks", g["uvw"], g["freq"] / C)
    for dc, vfl, out in ((g["dcb"], g["vflb"], g["out_b"]), (g["dcs"], g["vfls"], g["out_s"])):
        v = odft.dft_components(dc, lam, vfl)
        assert np.abs(v - out).max() < 1e-12 * np.abs(vfl).sum(axis=0).max()


def test_idft_against_reference_golden():
    """idft_components vs the reference's idft_visibility_skycomponent (linear-frame components:
    no polarisation conversion between the sums and the stored flux)."""
    g = np.load(os.path.join(os.path.dirname(GOLD), "idft_case.npz"))
    from ska_sdp_func_python_b200.imaging import skycoord_to_lmn
    nchan, npol = g["vis"].shape[2:]
    lam = np.einsum("tbs,k->tbks", g["uvw"], g["freq"] / C).reshape(-1, nchan, 3)
    dcs = []
    for ra, dec in g["dirs"]:
        l, m, _ = skycoord_to_lmn(ra, dec, g["pc"][0], g["pc"][1])
        dcs.append([l, m, np.sqrt(1 - l * l - m * m) - 1])
    flux, wsum = odft.idft_components(np.array(dcs), lam, g["vis"].reshape(-1, nchan, npol),
                                      g["weight"].reshape(-1, nchan, npol), g["flags"].reshape(-1, nchan, npol))
    assert np.array_equal(wsum == 0, g["weights"][0] == 0) and (wsum == 0).any()
    np.testing.assert_allclose(wsum, g["weights"][0], rtol=1e-13)
    lin = [k for k, f in enumerate(g["frames"]) if str(f) == "linear"]
    assert np.abs(flux[lin] - g["flux"][lin]).max() < 1e-12 * np.abs(g["flux"]).max()
