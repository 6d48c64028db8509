#!/usr/bin/env python
"""Golden vectors for the component-DFT row from THE REFERENCE'S OWN
dft_cpu_looped / dft_skycomponent_visibility / extract_direction_and_flux
(/root/reference/src/ska_sdp_func_python/imaging/dft.py:32-118, :265-285), run in the build
container:   python tests/golden/make_golden_dft.py
Substitutions as in make_golden.py (datamodels containers, SkyCoord, convert_pol_frame table);
`ska_sdp_func` (imported by dft_kernel for its optional proc_func back end) is a MagicMock."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402
from ska_sdp_func_python_b200 import minimal_models as mm  # noqa: E402


def main():
    _, ref_dft, _ = mg.load_reference()
    rng = np.random.default_rng(21)
    pc = mm.SkyCoord.from_deg(15.0, -45.0)
    ntimes, nbl, nchan = 3, 40, 5
    vis = mg.make_vis(rng, ntimes, nbl, nchan, "linear", 256, 1e-3, pc)
    vis["uvw"].data[...] *= 6.0    # phases of hundreds of turns
    freq = vis["frequency"].data
    comps = []
    for k in range(7):
        d = mm.SkyCoord.from_deg(15.0 + rng.uniform(-2, 2), -45.0 + rng.uniform(-2, 2))
        if k % 2:   # stokesIQUV flux for every channel -> converted to linear
            flux = rng.uniform(-1, 1, (nchan, 4)) + np.array([3.0, 0, 0, 0])
            comps.append(mm.SkyComponent(d, freq, flux, mm.PolarisationFrame("stokesIQUV")))
        else:       # single-channel component in the visibility frame (broadcast over channels)
            flux = rng.uniform(0.5, 2.0, (1, 4))
            comps.append(mm.SkyComponent(d, [freq[2]], flux, mm.PolarisationFrame("linear")))
    # (a) array level: dft_cpu_looped on per-channel fluxes
    dc = rng.uniform(-0.03, 0.03, (9, 3))
    dc[:, 2] = np.sqrt(1 - dc[:, 0] ** 2 - dc[:, 1] ** 2) - 1
    vfl = rng.normal(size=(9, nchan, 4)) + 1j * rng.normal(size=(9, nchan, 4))
    lam = np.einsum("tbs,k->tbks", vis["uvw"].data, freq / mm.C_M_S)
    vis_a = ref_dft.dft_cpu_looped(dc, lam, vfl)
    vis_a1 = ref_dft.dft_cpu_looped(dc, lam, vfl[:, :1, :1])
    # (b) data-model level, all-channel components only (mixed channel counts do not stack)
    full = [c for c in comps if c.flux.shape[0] == nchan]
    single = [c for c in comps if c.flux.shape[0] == 1]
    dcb, vflb = ref_dft.extract_direction_and_flux(full, vis)
    out_b = ref_dft.dft_skycomponent_visibility(vis.copy(deep=True), full)["vis"].data
    dcs, vfls = ref_dft.extract_direction_and_flux(single, vis)
    out_s = ref_dft.dft_skycomponent_visibility(vis.copy(deep=True), single)["vis"].data

    def pack(cs):
        return (np.array([[c.direction.ra.rad, c.direction.dec.rad] for c in cs]),
                np.array([c.flux for c in cs]))

    dir_full, flux_full = pack(full)
    dir_single, flux_single = pack(single)
    np.savez_compressed(
        os.path.join(HERE, "dft_comp_case.npz"), uvw=vis["uvw"].data, freq=freq,
        pc=np.array([pc.ra.rad, pc.dec.rad]), dc=dc, vfl=vfl, vis_a=vis_a, vis_a1=vis_a1,
        dir_full=dir_full, flux_full=flux_full, dcb=dcb, vflb=vflb, out_b=out_b,
        dir_single=dir_single, flux_single=flux_single, freq_single=freq[2], dcs=dcs, vfls=vfls,
        out_s=out_s)
    print("wrote dft_comp_case.npz", vis_a.shape, out_b.shape, out_s.shape,
          "max phase (turns)", np.abs(lam).max() * 0.04)


if __name__ == "__main__":
    main()
