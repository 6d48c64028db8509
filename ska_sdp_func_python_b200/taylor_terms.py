"""Frequency-moment (Taylor-term) images for the MSMFS CLEAN wrapper: the two list functions of
/root/reference/src/ska_sdp_func_python/image/taylor_terms.py that ``mmclean_kernel_list`` calls
(:160-237 calculate_image_list_frequency_moments, :240-288
calculate_image_list_from_frequency_taylor_terms).  Weighted sums over a handful of channel images:
data-model glue around the device kernel, done with numpy on the arrays the images hold."""
import copy as _copy

import numpy

from .cleaners import _new_image_like

__all__ = ["calculate_image_list_frequency_moments", "calculate_image_list_from_frequency_taylor_terms"]


def _image_frequency(im):
    """Frequency of channel 0 of a single-channel image, from its WCS (0-based pixel 0)."""
    w = im.image_acc.wcs.wcs
    return w.crval[3] + (1.0 - w.crpix[3]) * w.cdelt[3]


def calculate_image_list_frequency_moments(im_list, reference_frequency=None, nmoment=1):
    """moment k = sum over channels of ((nu - nu_ref) / nu_ref)^k x channel image; nu_ref defaults
    to the middle channel; the spectral axis becomes a MOMENT axis."""
    if nmoment <= 0:
        raise ValueError(f"Number of moments {nmoment} must be > 0")
    if im_list[0] is None:
        return None
    nchan = len(im_list)
    _, npol, ny, nx = im_list[0]["pixels"].data.shape
    freq = [_image_frequency(im) for im in im_list]
    if nmoment > nchan:
        raise ValueError(f"Number of moments {nmoment} cannot exceed the number of channels {nchan}")
    if reference_frequency is None:
        reference_frequency = freq[nchan // 2]
    moment_data = numpy.zeros([nmoment, npol, ny, nx])
    for moment in range(nmoment):
        for chan, im in enumerate(im_list):
            weight = numpy.power((freq[chan] - reference_frequency) / reference_frequency, moment)
            moment_data[moment, ...] += im["pixels"].data[0, ...] * weight
    out = _new_image_like(im_list[0], moment_data)
    wcs = _copy.deepcopy(im_list[0].image_acc.wcs)
    wcs.wcs.ctype[3] = "MOMENT"
    wcs.wcs.crval[3] = 0.0
    wcs.wcs.crpix[3] = 1.0
    wcs.wcs.cdelt[3] = 1.0
    if hasattr(out, "attrs") and "wcs" in out.attrs:
        out.attrs["wcs"] = wcs
    return out


def calculate_image_list_from_frequency_taylor_terms(im_list, moment_image, reference_frequency=None):
    """Channel images back from the moment cube: sum over k of moment k x ((nu - nu_ref) / nu_ref)^k."""
    nchan = len(im_list)
    nmoment = moment_image["pixels"].data.shape[0]
    frequency = numpy.array([_image_frequency(d) for d in im_list])
    if reference_frequency is None:
        reference_frequency = frequency[nchan // 2]
    newims = []
    for chan in range(nchan):
        newim_data = numpy.zeros_like(im_list[chan]["pixels"].data[...])
        for moment in range(nmoment):
            weight = numpy.power((frequency[chan] - reference_frequency) / reference_frequency, moment)
            newim_data[0, ...] += moment_image["pixels"].data[moment, ...] * weight
        newims.append(_new_image_like(im_list[chan], newim_data))
    return newims
