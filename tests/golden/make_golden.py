#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by running THE REFERENCE'S OWN CODE.

Run in the build container (needs /root/reference; the GPU box does not have it, which is why
the vectors are committed):

    python tests/golden/make_golden.py

What is imported from /root/reference/src (unmodified):
  * ska_sdp_func_python.imaging.ng.invert_ng / predict_ng           (ng.py:38-294)
  * ska_sdp_func_python.imaging.base.shift_vis_to_image / normalise_sumwt (base.py:48-155)
  * ska_sdp_func_python.visibility.base.phaserotate_visibility / calculate_visibility_phasor
  * ska_sdp_func_python.util.coordinate_support.skycoord_to_lmn     (coordinate_support.py:436)
  * ska_sdp_func_python.imaging.dft.dft_cpu_looped                  (dft.py:265-285)

What is NOT available here and is therefore substituted (each substitution is the only way to
run the reference wrapper in this container; they are listed in DESIGN.md "Oracle"):
  * ducc0.wgridder        -> the exact fp64 direct DFT of oracle/dft.py (so the goldens hold
                             the exact measurement equation, i.e. ducc0's answer up to its epsilon)
  * ska_sdp_datamodels    -> the duck-typed containers of ska_sdp_func_python_b200/minimal_models.py
                             and the polarisation table of ska_sdp_func_python_b200/polarisation.py
  * astropy               -> SkyCoord / pixel_to_skycoord stand-ins (same module)
Every other third-party import of the reference package is satisfied by a MagicMock module
(the technique the reference's own docs/src/conf.py:28-37 uses for ducc0).
"""
import importlib.abc
import importlib.machinery
import os
import sys
import types
from unittest import mock

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import dft as odft  # noqa: E402
from ska_sdp_func_python_b200 import minimal_models as mm  # noqa: E402
from ska_sdp_func_python_b200.polarisation import pol_matrix_table  # noqa: E402

MOCK_PREFIXES = ("astropy", "ska_sdp_datamodels", "ducc0", "photutils", "ska_sdp_func",
                 "xarray", "astroplan", "h5py", "radler", "dp3", "wagg", "cupy", "pyfftw")


class _MockLoader(importlib.abc.Loader):
    def create_module(self, spec):
        m = mock.MagicMock(name=spec.name)
        m.__name__ = spec.name
        m.__path__ = []
        m.__spec__ = spec
        m.__loader__ = self
        return m

    def exec_module(self, module):
        pass


class _MockFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path, target=None):
        if fullname.split(".")[0] in MOCK_PREFIXES and not fullname.startswith("ska_sdp_func_python"):
            return importlib.machinery.ModuleSpec(fullname, _MockLoader(), is_package=True)
        return None


def convert_pol_frame(polvec, ipf, opf, polaxis=1):
    """Stand-in for ska_sdp_datamodels' convert_pol_frame: constant matrix along polaxis."""
    M = pol_matrix_table(ipf, opf)
    v = np.moveaxis(np.asarray(polvec), polaxis, -1)
    return np.moveaxis(v @ M.T, -1, polaxis)


class _Backend:
    """ducc0.wgridder stand-in: the exact direct DFT."""

    @staticmethod
    def ms2dirty(uvw, freq, ms, wgt, npix_x, npix_y, pixsize_x, pixsize_y, nu, nv, epsilon,
                 do_wstacking, nthreads=1, verbosity=0, mask=None,
                 double_precision_accumulation=False):
        return odft.ms2dirty(uvw, freq, ms, wgt, npix_x, npix_y, pixsize_x, pixsize_y, do_wstacking)

    @staticmethod
    def dirty2ms(uvw, freq, dirty, wgt, pixsize_x, pixsize_y, nu, nv, epsilon, do_wstacking,
                 nthreads=1, verbosity=0, mask=None):
        return odft.dirty2ms(uvw, freq, dirty, wgt, pixsize_x, pixsize_y, do_wstacking)


def load_reference():
    sys.meta_path.insert(0, _MockFinder())
    stub = types.ModuleType

    def seed(name, **attrs):
        m = stub(name)
        m.__path__ = []
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m

    seed("ska_sdp_datamodels", physical_constants=types.SimpleNamespace(C_M_S=mm.C_M_S))
    seed("ska_sdp_datamodels.image")
    seed("ska_sdp_datamodels.image.image_model", Image=mm.Image)
    seed("ska_sdp_datamodels.visibility")
    seed("ska_sdp_datamodels.visibility.vis_model", Visibility=mm.Visibility)
    seed("ska_sdp_datamodels.science_data_model")
    seed("ska_sdp_datamodels.science_data_model.polarisation_functions",
         convert_pol_frame=convert_pol_frame)
    seed("ska_sdp_datamodels.science_data_model.polarisation_model",
         PolarisationFrame=mm.PolarisationFrame)
    seed("astropy")
    seed("astropy.wcs")
    seed("astropy.wcs.utils", pixel_to_skycoord=mm.pixel_to_skycoord)
    seed("astropy.coordinates", SkyCoord=mm.SkyCoord, CartesianRepresentation=object)
    seed("ducc0")
    seed("ducc0.wgridder", ms2dirty=_Backend.ms2dirty, dirty2ms=_Backend.dirty2ms)
    sys.path.insert(0, "/root/reference/src")
    import ska_sdp_func_python.imaging.ng as ref_ng
    import ska_sdp_func_python.imaging.dft as ref_dft
    import ska_sdp_func_python.imaging.imaging as ref_imaging
    return ref_ng, ref_dft, ref_imaging


def make_vis(rng, ntimes, nbl, nchan, pf, npix, cell, phasecentre, f0=1.0e8, df=2.0e6, wfrac=0.2,
             flag_fraction=0.05):
    freq = f0 + df * np.arange(nchan)
    umax = 0.4 / cell / (freq[-1] / mm.C_M_S)
    uvw = rng.uniform(-1, 1, (ntimes, nbl, 3)) * np.array([umax, umax, wfrac * umax])
    uvw[0, 0] = 0.0  # an "autocorrelation" row
    npol = mm.PolarisationFrame(pf).npol
    shape = (ntimes, nbl, nchan, npol)
    vis = rng.normal(size=shape) + 1j * rng.normal(size=shape)
    weight = rng.uniform(0.5, 2.0, shape)
    weight[0, 0] = 0.0
    flags = (rng.uniform(size=shape) < flag_fraction).astype(int)
    return mm.Visibility(vis, uvw, freq, phasecentre, mm.PolarisationFrame(pf), weight=weight,
                         imaging_weight=weight.copy(), flags=flags)


def dump_case(path, vis, im, **arrays):
    w = im.image_acc.wcs.wcs
    np.savez_compressed(
        path, vis=vis["vis"].data, uvw=vis["uvw"].data, frequency=vis["frequency"].data,
        weight=vis["weight"].data, imaging_weight=vis["imaging_weight"].data,
        flags=vis["flags"].data, vis_pc=np.array([vis.phasecentre.ra.rad, vis.phasecentre.dec.rad]),
        vis_pf=vis.visibility_acc.polarisation_frame.type, im_pf=im.image_acc.polarisation_frame.type,
        im_shape=np.array(im["pixels"].data.shape), crval=w.crval, crpix=w.crpix, cdelt=w.cdelt,
        **arrays)
    print("wrote", path, {k: np.asarray(v).shape for k, v in arrays.items()})


def main():
    ref_ng, ref_dft, ref_imaging = load_reference()
    rng = np.random.default_rng(1805550721)
    pc = mm.SkyCoord.from_deg(180.0, -35.0)

    # ---- A: stokesI, MFS, image centred on the visibility phase centre -------------------
    npix, cell = 64, 1.0e-3
    vis = make_vis(rng, 3, 40, 3, "stokesI", npix, cell, pc)
    im = mm.create_image(npix, cell, pc, vis["frequency"].data[:1], channel_bandwidth=6.0e6,
                         polarisation_frame=mm.PolarisationFrame("stokesI"), nchan=1)
    model = im.copy(deep=True)
    model["pixels"].data[0, 0] = rng.normal(size=(npix, npix)) * (rng.uniform(size=(npix, npix)) < 0.02)
    out = {}
    d, sw = ref_ng.invert_ng(vis, im, dopsf=False, normalise=True)
    out["dirty_norm"], out["sumwt"] = d["pixels"].data, sw
    d, sw = ref_ng.invert_ng(vis, im, dopsf=False, normalise=False)
    out["dirty_raw"] = d["pixels"].data
    d, sw = ref_ng.invert_ng(vis, im, dopsf=True, normalise=True)
    out["psf"] = d["pixels"].data
    d, sw = ref_imaging.invert_visibility(vis, im, context="2d")
    out["dirty_2d"] = d["pixels"].data
    p = ref_ng.predict_ng(vis, model)
    out["model"], out["predicted"] = model["pixels"].data, p["vis"].data
    p = ref_imaging.predict_visibility(vis, model, context="2d")
    out["predicted_2d"] = p["vis"].data
    dump_case(os.path.join(HERE, "ng_case_a.npz"), vis, im, **out)

    # ---- B: linear -> stokesIQUV cube, image centre offset from the phase centre ----------
    npix, cell = 48, 1.5e-3
    vis = make_vis(rng, 2, 30, 4, "linear", npix, cell, pc, wfrac=0.1)
    im_pc = mm.SkyCoord.from_deg(181.0, -34.5)
    im = mm.create_image(npix, cell, im_pc, vis["frequency"].data[:1] + 1.0e6, channel_bandwidth=4.0e6,
                         polarisation_frame=mm.PolarisationFrame("stokesIQUV"), nchan=2)
    model = im.copy(deep=True)
    model["pixels"].data[...] = rng.normal(size=model["pixels"].data.shape) * (
        rng.uniform(size=model["pixels"].data.shape) < 0.03)
    out = {}
    d, sw = ref_ng.invert_ng(vis, im, dopsf=False, normalise=True)
    out["dirty_norm"], out["sumwt"] = d["pixels"].data, sw
    d, sw = ref_ng.invert_ng(vis, im, dopsf=True, normalise=False)
    out["psf_raw"], out["sumwt_psf"] = d["pixels"].data, sw
    p = ref_ng.predict_ng(vis, model)
    out["model"], out["predicted"] = model["pixels"].data, p["vis"].data
    out["predicted_pc"] = np.array([p.phasecentre.ra.rad, p.phasecentre.dec.rad])
    dump_case(os.path.join(HERE, "ng_case_b.npz"), vis, im, **out)

    # ---- C: the reference's own direct DFT pins the oracle's sign / axis convention -------
    npix, cell = 32, 2.0e-3
    ntimes, nbl, nchan = 2, 25, 2
    freq = np.array([1.0e8, 1.1e8])
    umax = 0.4 / cell / (freq[-1] / mm.C_M_S)
    uvw = rng.uniform(-1, 1, (ntimes, nbl, 3)) * np.array([umax, umax, 0.3 * umax])
    ix = rng.integers(2, npix - 2, 6)
    iy = rng.integers(2, npix - 2, 6)
    flux = rng.uniform(0.5, 2.0, 6)
    l = -(ix - npix // 2) * cell       # RA (x) axis decreases with pixel index
    m = (iy - npix // 2) * cell
    n = np.sqrt(1 - l * l - m * m)
    dc = np.stack([l, m, n - 1.0], axis=1)
    vfl = np.zeros((6, nchan, 1), dtype=complex)
    vfl[:, :, 0] = (flux / n)[:, None]  # a pixel of a Jy/pixel image carries flux / n
    uvw_lambda = np.einsum("tbs,k->tbks", uvw, freq / mm.C_M_S)
    vis_ref = ref_dft.dft_cpu_looped(dc, uvw_lambda, vfl)
    image = np.zeros((npix, npix))      # [ny, nx]
    np.add.at(image, (iy, ix), flux)
    np.savez_compressed(os.path.join(HERE, "dft_case_c.npz"), uvw=uvw, frequency=freq, image=image,
                        cellsize=cell, vis=vis_ref)
    print("wrote dft_case_c.npz", vis_ref.shape)


if __name__ == "__main__":
    main()
