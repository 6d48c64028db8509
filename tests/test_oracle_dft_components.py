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
