#!/usr/bin/env python
"""Golden vectors for the adjoint component DFT from THE REFERENCE'S OWN
idft_visibility_skycomponent (/root/reference/src/ska_sdp_func_python/imaging/dft.py:334-384),
run in the build container:   python tests/golden/make_golden_idft.py
Substitutions as in make_golden.py (datamodels containers, SkyCoord, convert_pol_frame table)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402
from ska_sdp_func_python_b200 import minimal_models as mm  # noqa: E402


def main():
    _, ref_dft, _ = mg.load_reference()
    rng = np.random.default_rng(51)
    pc = mm.SkyCoord.from_deg(15.0, -45.0)
    vis = mg.make_vis(rng, 3, 35, 4, "linear", 256, 1e-3, pc, flag_fraction=0.1)
    vis["weight"].data[:, :, 2, 1] = 0.0       # a (chan, pol) with zero total weight
    freq = vis["frequency"].data
    comps = []
    for k in range(5):
        d = mm.SkyCoord.from_deg(15.0 + rng.uniform(-1, 1), -45.0 + rng.uniform(-1, 1))
        pf = "stokesIQUV" if k % 2 else "linear"
        comps.append(mm.SkyComponent(d, freq, np.ones((4, 4), dtype=complex), mm.PolarisationFrame(pf)))
    newsc, wl = ref_dft.idft_visibility_skycomponent(vis, comps)
    np.savez_compressed(
        os.path.join(HERE, "idft_case.npz"), vis=vis["vis"].data, uvw=vis["uvw"].data, freq=freq,
        weight=vis["weight"].data, flags=vis["flags"].data, pc=np.array([pc.ra.rad, pc.dec.rad]),
        dirs=np.array([[c.direction.ra.rad, c.direction.dec.rad] for c in comps]),
        frames=np.array([c.polarisation_frame.type for c in comps]),
        flux=np.array([c.flux for c in newsc]), weights=np.array(wl))
    print("wrote idft_case.npz", np.array([c.flux for c in newsc]).shape, np.array(wl).shape)


if __name__ == "__main__":
    main()
