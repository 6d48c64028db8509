"""Major-cycle glue on a B200: drop-ins for
``skymodel_predict_calibrate`` and ``skymodel_calibrate_invert``
(/root/reference/src/ska_sdp_func_python/sky_model/skymodel_imaging.py:70-151, :154-235) built on
the CUDA component DFT (dft.py), degridder and gridder (imaging.py) -- the primary-beam-weighted
sums over time slices and their normalisation stay on the device -- plus ``skymodel_residual``:
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


def _normalise_weighted(sum_dirty, sum_flat, min_weight=0.1, flat_sky=False):
    """Primary-beam-weighted normalisation on the device: what ``normalise_sumwt`` does when the
    sum of weights is an Image (imaging/base.py:129-149) followed by the caller's final square root
    of the weights (skymodel_imaging.py:226-227), written as whole-cube tensor expressions.

    sum_dirty, sum_flat: CUDA tensors [nchan, npol, ny, nx].  Returns (image, sensitivity).
      flat_sky : image / sqrt(W_centre * W) where that exceeds min_weight * max(W), else image /
                 max(W); sensitivity = sqrt(W).
      otherwise: image / max(W).  The reference takes the square root of the WHOLE weight cube once
                 per (chan, pol) plane of its loop, so plane j (of K, in loop order) ends up as
                 (W_j^(2^-j) / max(W))^(2^-(K-j+1)) -- (W / max(W))^(1/4) for a single plane.  Kept,
                 because callers of the reference see exactly these numbers."""
    import torch
    maxwt = sum_flat.max()
    nchan, npol, ny, nx = sum_flat.shape
    if flat_sky:
        centre = sum_flat[:, :, ny // 2:ny // 2 + 1, nx // 2:nx // 2 + 1]
        norm = torch.sqrt(centre * sum_flat)
        image = torch.where(norm > min_weight * maxwt, sum_dirty / norm, sum_dirty / maxwt)
        return image, torch.sqrt(sum_flat)
    k = nchan * npol
    j = torch.arange(k, device=sum_flat.device, dtype=torch.float64).reshape(nchan, npol, 1, 1)
    sens = (sum_flat ** (0.5 ** j) / maxwt) ** (0.5 ** (k - j + 1))
    return sum_dirty / maxwt, sens


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
    if get_pb is not None:
        # one gridding per time slice (each has its own primary beam); the beam-weighted image
        # sum_t flat_t * dirty_t and the weight image sum_t flat_t^2 * sumwt_t stay on the device
        # until the single normalisation at the end
        import torch
        sum_dirty = sum_flat = None
        mask = None
        for _, vis_slice in _time_slices(bvis):
            pb = get_pb(vis_slice, skymodel.image)
            pix, sumwt = invert_visibility(vis_slice, skymodel.image, context=context, normalise=False,
                                           _device_result=True, **kwargs)
            if sum_dirty is None:
                sum_dirty, sum_flat = torch.zeros_like(pix), torch.zeros_like(pix)
                if skymodel.mask is not None:
                    mask = torch.as_tensor(numpy.asarray(skymodel.mask["pixels"].data, dtype=float),
                                           device=pix.device)
            flat = None if mask is None else mask
            if pb is not None:
                beam = torch.as_tensor(numpy.asarray(pb["pixels"].data, dtype=float), device=pix.device)
                flat = beam if flat is None else flat * beam
            wt = torch.as_tensor(sumwt, device=pix.device)[:, :, None, None]
            if flat is None:
                sum_dirty += pix
                sum_flat += wt
            else:
                sum_dirty.addcmul_(flat, pix)
                sum_flat.addcmul_(flat * flat, wt.expand_as(flat))
        if normalise:
            sum_dirty, sum_flat = _normalise_weighted(sum_dirty, sum_flat, flat_sky=flat_sky)
        dirty_im, flat_im = skymodel.image.copy(deep=True), skymodel.image.copy(deep=True)
        dirty_im["pixels"].data[...] = sum_dirty.cpu().numpy()
        flat_im["pixels"].data[...] = sum_flat.cpu().numpy()
        return dirty_im, flat_im
    result = invert_visibility(bvis, skymodel.image, context=context, **kwargs)
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
