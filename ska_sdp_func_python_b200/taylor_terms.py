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


def _moment_weights(freq, reference_frequency, nmoment):
    """w[k, c] = ((nu_c - nu_ref) / nu_ref)^k; nu_ref defaults to the middle channel."""
    freq = numpy.asarray(freq, dtype=float)
    ref = freq[len(freq) // 2] if reference_frequency is None else reference_frequency
    return numpy.power(((freq - ref) / ref)[None, :], numpy.arange(nmoment)[:, None])


def calculate_image_list_frequency_moments(im_list, reference_frequency=None, nmoment=1):
    """moment k = sum over channels of ((nu - nu_ref) / nu_ref)^k x channel image; the spectral axis
    becomes a MOMENT axis.  Channels are accumulated one after the other, in list order."""
    if nmoment <= 0:
        raise ValueError(f"Number of moments {nmoment} must be > 0")
    if im_list[0] is None:
        return None
    if nmoment > len(im_list):
        raise ValueError(f"Number of moments {nmoment} cannot exceed the number of channels {len(im_list)}")
    w = _moment_weights([_image_frequency(im) for im in im_list], reference_frequency, nmoment)
    cube = numpy.zeros((nmoment,) + im_list[0]["pixels"].data.shape[1:])
    for chan, im in enumerate(im_list):
        plane = im["pixels"].data[0]
        for k in range(nmoment):
            cube[k] += plane * w[k, chan]
    out = _new_image_like(im_list[0], cube)
    wcs = _copy.deepcopy(im_list[0].image_acc.wcs)
    wcs.wcs.ctype[3], wcs.wcs.crval[3], wcs.wcs.crpix[3], wcs.wcs.cdelt[3] = "MOMENT", 0.0, 1.0, 1.0
    if hasattr(out, "attrs") and "wcs" in out.attrs:
        out.attrs["wcs"] = wcs
    return out


def calculate_image_list_from_frequency_taylor_terms(im_list, moment_image, reference_frequency=None):
    """Channel images back from the moment cube: image_c = sum over k of moment k x w[k, c]."""
    cube = moment_image["pixels"].data
    w = _moment_weights([_image_frequency(d) for d in im_list], reference_frequency, cube.shape[0])
    rebuilt = []
    for chan, template in enumerate(im_list):
        data = numpy.zeros_like(template["pixels"].data)
        for k in range(cube.shape[0]):
            data[0] += cube[k] * w[k, chan]
        rebuilt.append(_new_image_like(template, data))
    return rebuilt
