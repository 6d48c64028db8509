"""Shared fixtures.  `-m "not gpu"` covers the oracle against the golden vectors, the host logic
and the C-ABI surface; `-m gpu` holds the parity tests proper (all through the C ABI)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
C_M_S = 299792458.0


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """The in-tree CUDA library and the CPU oracle, built once per session."""
    import __graft_entry__ as entry
    entry.build()
    from ska_sdp_func_python_b200 import _lib
    return _lib.lib()


def make_case(seed, nrow, nchan, npx, npy, psx, psy, wfrac, clump=True):
    """Random uvw / ms / wgt covering the grid, with a dense clump, exact repeats and
    zero-weight rows (the collisions / nulls of this domain)."""
    rng = np.random.default_rng(seed)
    freq = np.linspace(100e6, 112e6, nchan) if nchan > 1 else np.array([100e6])
    lam = freq[-1] / C_M_S
    uvw = rng.uniform(-1, 1, (nrow, 3))
    uvw[:, 0] *= 0.48 / psx / lam
    uvw[:, 1] *= 0.48 / psy / lam
    uvw[:, 2] *= wfrac * 0.48 / max(psx, psy) / lam
    if clump and nrow >= 40:
        k = nrow // 4
        uvw[:k, :2] *= 0.02
        uvw[k:k + 10] = uvw[k]
    ms = rng.normal(size=(nrow, nchan)) + 1j * rng.normal(size=(nrow, nchan))
    wgt = rng.uniform(0.5, 1.5, (nrow, nchan))
    wgt[::17] = 0.0
    return uvw, freq, ms, wgt


def rel_l2(a, b):
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / np.linalg.norm(np.asarray(b)))


def rel_max(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / np.abs(np.asarray(b)).max())


def load_ng_case(name):
    """Rebuild the Visibility / Image of a golden case with the minimal data models."""
    from ska_sdp_func_python_b200 import minimal_models as mm
    g = np.load(os.path.join(GOLDEN, name))
    vis = mm.Visibility(g["vis"], g["uvw"], g["frequency"],
                        mm.SkyCoord(float(g["vis_pc"][0]), float(g["vis_pc"][1])),
                        mm.PolarisationFrame(str(g["vis_pf"])), weight=g["weight"],
                        imaging_weight=g["imaging_weight"], flags=g["flags"])
    wcs = mm.WCS(g["crval"], g["crpix"], g["cdelt"])
    im = mm.Image(np.zeros(tuple(g["im_shape"])), wcs, mm.PolarisationFrame(str(g["im_pf"])))
    return g, vis, im
