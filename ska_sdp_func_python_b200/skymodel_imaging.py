"""Major-cycle glue on a B200: drop-ins for
``skymodel_predict_calibrate`` and ``skymodel_calibrate_invert``
(/root/reference/src/ska_sdp_func_python/sky_model/skymodel_imaging.py:70-151, :154-235) built on
the CUDA component DFT (dft.py), degridder and gridder (imaging.py), plus ``skymodel_residual``:
the residual image of a SkyModel in one pass with the samples resident on the device
(imaging.residual_ng -> b2g_plan_residual), which is what a major cycle actually needs.

Same names, argument meaning and error behaviour as the reference.  Calibration (``docal`` with a
gain table) belongs to the reference's calibration package (calibration/operations.py) and is
outside this path: ``docal=True`` with a gain table present raises NotImplementedError rather than
silently skipping it.
"""
import numpy

from .dft import dft_skycomponent_visibility
from .imaging import invert_visibility, predict_visibility, residual_ng

__all__ = ["skymodel_predict_calibrate", "skymodel_calibrate_invert", "skymodel_residual", "subtract_visibility"]


def _time_slices(v):
    if hasattr(v, "groupby"):
        return list(v.groupby("time", squeeze=False))
    from .minimal_models import time_slices
    return time_slices(v)


def _concatenate(vis_slices):
    try:
        from ska_sdp_datamodels.visibility.vis_utils import concatenate_visibility
    except ImportError:
        from .minimal_models import concatenate_visibility
    return concatenate_visibility(vis_slices, "time")


def _no_calibration(docal, skymodel):
    if docal and skymodel.gaintable is not None:
        raise NotImplementedError(
            "apply_gaintable is part of the reference's calibration package, outside the B200 "
            "imaging path: apply the gain table with the reference before / after this call")


def _normalise_sumwt_image(im, sumwt, min_weight=0.1, flat_sky=False):
    """normalise_sumwt for a sum of weights that is an Image (primary-beam case), statement by
    statement as imaging/base.py:129-149 -- including its in-place updates of ``sumwt``."""
    maxwt = numpy.max(sumwt["pixels"].data)
    minwt = min_weight * maxwt
    nchan, npol, ny, nx = sumwt["pixels"].data.shape
    cx, cy = nx // 2, ny // 2
    for chan in range(nchan):
        for pol in range(npol):
            if flat_sky:
                norm = numpy.sqrt(sumwt["pixels"].data[chan, pol, cy, cx]
                                  * sumwt["pixels"].data[chan, pol, :, :])
                im["pixels"].data[chan, pol, :, :][norm > minwt] /= norm[norm > minwt]
                im["pixels"].data[chan, pol, :, :][norm <= minwt] /= maxwt
            else:
                im["pixels"].data[chan, pol, :, :] /= maxwt
                sumwt["pixels"].data[chan, pol, :, :] /= maxwt
                sumwt["pixels"].data = numpy.sqrt(sumwt["pixels"].data)
    return im


def _dft_sky_component(vis_slice, skymodel, pb=None, **kwargs):
    """Run DFT of sky components (skymodel_imaging.py:22-45).  With a mask or a primary beam the
    component fluxes are first scaled by ``apply_beam_to_skycomponent`` (``sky_component.py``: the
    reference's function with the SIN projection written out instead of astropy; another
    implementation can be passed as ``apply_beam_to_skycomponent=``)."""
    comps = skymodel.components
    if skymodel.mask is not None or pb is not None:
        comps = list(skymodel.components)
        if len(comps) > 0:
            apply_beam = kwargs.get("apply_beam_to_skycomponent")
            if apply_beam is None:
                from .sky_component import apply_beam_to_skycomponent as apply_beam
            if skymodel.mask is not None:
                comps = apply_beam(comps, skymodel.mask)
            if pb is not None:
                comps = apply_beam(comps, pb)
    return dft_skycomponent_visibility(vis_slice, comps,
                                       dft_compute_kernel=kwargs.get("dft_compute_kernel", None))


def _fft_image(vis_slice, context, skymodel, pb=None, **kwargs):
    """Add the degridded image (times mask and primary beam) to vis_slice
    (skymodel_imaging.py:48-67)."""
    imgv = vis_slice.copy(deep=True, zero=True)
    model = skymodel.image
    if skymodel.mask is not None or pb is not None:
        model = skymodel.image.copy(deep=True)
        if skymodel.mask is not None:
            model["pixels"].data *= skymodel.mask["pixels"].data
        if pb is not None:
            model["pixels"].data *= pb["pixels"].data
    imgv = predict_visibility(imgv, model, context=context, **kwargs)
    vis_slice["vis"].data += imgv["vis"].data


def skymodel_predict_calibrate(bvis, skymodel, context="ng", docal=False, inverse=True,
                               get_pb=None, **kwargs):
    """Predict visibility for a SkyModel: DFT of the components plus degridded image
    (skymodel_imaging.py:70-151)."""
    _no_calibration(docal, skymodel)
    kw = {k: v for k, v in kwargs.items() if k not in ("apply_beam_to_skycomponent",)}
    v = bvis.copy(deep=True, zero=True)
    has_image = (skymodel.image is not None
                 and numpy.max(numpy.abs(skymodel.image["pixels"].data)) > 0.0)
    if get_pb is not None:
        vis_slices = []
        for _, vis_slice in _time_slices(v):
            pb = get_pb(vis_slice, skymodel.image)
            if len(skymodel.components) > 0:
                vis_slice = _dft_sky_component(vis_slice, skymodel, pb=pb, **kwargs)
            if has_image:
                _fft_image(vis_slice, context, skymodel, pb=pb, **kw)
            vis_slices.append(vis_slice)
        return _concatenate(vis_slices)
    v = _dft_sky_component(v, skymodel, pb=None, **kwargs)
    if has_image:
        _fft_image(v, context, skymodel, pb=None, **kw)
    return v


def skymodel_calibrate_invert(bvis, skymodel, context="ng", docal=False, get_pb=None,
                              normalise=True, flat_sky=False, **kwargs):
    """Inverse Fourier sum of Visibility to Image (skymodel_imaging.py:154-235).  Returns
    (dirty Image, sumwt) -- with ``get_pb``, (weighted dirty Image, sensitivity Image)."""
    if skymodel.image is None:
        raise ValueError("skymodel image is None")
    _no_calibration(docal, skymodel)
    bvis_cal = bvis.copy(deep=True)
    if get_pb is not None:
        sum_flats = skymodel.image.copy(deep=True)
        sum_flats["pixels"].data[...] = 0.0
        sum_dirtys = skymodel.image.copy(deep=True)
        sum_dirtys["pixels"].data[...] = 0.0
        for _, vis_slice in _time_slices(bvis_cal):
            pb = get_pb(vis_slice, skymodel.image)
            result = invert_visibility(vis_slice, skymodel.image, context=context, normalise=False,
                                       **kwargs)
            flat = numpy.ones_like(result[0]["pixels"].data)
            if skymodel.mask is not None:
                flat *= skymodel.mask["pixels"].data
            if pb is not None:
                flat *= pb["pixels"].data
            sum_dirtys["pixels"].data += flat * result[0]["pixels"].data
            sum_flats["pixels"].data += flat * flat * result[1][:, :, numpy.newaxis, numpy.newaxis]
        if normalise:
            sum_dirtys = _normalise_sumwt_image(sum_dirtys, sum_flats, flat_sky=flat_sky)
            sum_flats["pixels"].data = numpy.sqrt(sum_flats["pixels"].data)
        return sum_dirtys, sum_flats
    result = invert_visibility(bvis_cal, skymodel.image, context=context, **kwargs)
    if skymodel.mask is not None:
        result[0]["pixels"].data *= skymodel.mask["pixels"].data
    return result


def skymodel_residual(bvis, skymodel, context="ng", normalise=True, **kwargs):
    """Residual image of a SkyModel against the observed visibilities:
    ``skymodel_calibrate_invert(bvis - skymodel_predict_calibrate(bvis, skymodel), skymodel)``
    of the reference's major cycle, with the image part fused on the device: the components are
    transformed by the CUDA DFT and subtracted, then ONE plan degrids the (masked) image, subtracts
    it sample by sample and grids the residual (imaging.residual_ng).  Returns (Image, sumwt)."""
    if skymodel.image is None:
        raise ValueError("skymodel image is None")
    if context not in ("ng", "2d"):
        raise ValueError(f"Unknown imaging context {context}")
    if context == "2d":
        kwargs = dict(kwargs, do_wstacking=False)
    work = bvis
    if len(skymodel.components) > 0:
        comp = _dft_sky_component(bvis.copy(deep=True, zero=True), skymodel, pb=None, **kwargs)
        work = bvis.copy(deep=True)
        work["vis"].data = work["vis"].data - comp["vis"].data
    model = skymodel.image
    if skymodel.mask is not None:
        model = skymodel.image.copy(deep=True)
        model["pixels"].data *= skymodel.mask["pixels"].data
    kw = {k: v for k, v in kwargs.items()
          if k not in ("dft_compute_kernel", "apply_beam_to_skycomponent")}
    result = residual_ng(work, model, normalise=normalise, **kw)
    if skymodel.mask is not None:
        result[0]["pixels"].data *= skymodel.mask["pixels"].data
    return result


def subtract_visibility(vis, model_vis, inplace=False):
    """vis - model_vis as a new Visibility, or in place (visibility/operations.py:86-105).  This is
    the stand-alone host form for callers that already hold both cubes in host memory; inside the
    major cycle the subtraction is fused into the degridder (``skymodel_residual`` /
    ``imaging.residual_ng``) and never leaves the device."""
    assert vis.vis.shape == model_vis.vis.shape, (
        f"Observed {vis.vis.shape} and model visibilities {model_vis.vis.shape} have different shapes.")
    if inplace:
        vis["vis"].data = vis["vis"].data - model_vis["vis"].data
        return vis
    residual_vis = vis.copy(deep=True)
    residual_vis["vis"].data = residual_vis["vis"].data - model_vis["vis"].data
    return residual_vis
