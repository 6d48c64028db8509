"""Minimal duck-typed ``Visibility`` / ``Image`` containers.

``invert_ng`` / ``predict_ng`` normally receive ``ska_sdp_datamodels`` objects (xarray Datasets
with accessors; external package, not installable in this build environment).  These classes
carry exactly the attributes the reference's hot path touches (SURVEY.md section 8 a9;
/root/reference/src/ska_sdp_func_python/imaging/ng.py:63-69, :134, :185-203,
imaging/base.py:60-70, :90, visibility/base.py:35-43) so that the drop-in functions -- and the
reference's own ng.py, for generating golden vectors -- can be exercised without datamodels,
xarray or astropy.  They are not a re-implementation of the data models.
"""
import copy as _copy

import numpy

C_M_S = 299792458.0


def _deep_copy(a):
    """numpy.array(a, copy=True); arrays of hundreds of MB are copied by torch's multi-threaded
    CPU copy when torch is importable (a Visibility of 2.5e8 samples holds 6 GB of weights and
    flags, which one numpy thread copies in 1.4 s)."""
    a = numpy.asarray(a)
    if a.nbytes >= (1 << 26) and a.flags.c_contiguous and a.dtype.kind in "fciub":
        try:
            import torch
            out = numpy.empty_like(a)
            torch.from_numpy(out).copy_(torch.from_numpy(a))
            return out
        except Exception:  # noqa: BLE001  (torch missing or dtype unsupported)
            pass
    return numpy.array(a, copy=True)


class Angle:
    def __init__(self, rad):
        self.rad = float(rad)

    @property
    def deg(self):
        return float(numpy.degrees(self.rad))


class _Scalar:
    def __init__(self, value):
        self.value = float(value)


class _Cartesian:
    def __init__(self, x, y, z):
        self.x, self.y, self.z = _Scalar(x), _Scalar(y), _Scalar(z)


class _OffsetFrame:
    def __init__(self, origin):
        self.origin = origin


class _OffsetCoord:
    """Position expressed in the sky-offset frame of ``origin`` (x towards the origin)."""

    def __init__(self, pos, origin):
        dra = pos.ra.rad - origin.ra.rad
        d, d0 = pos.dec.rad, origin.dec.rad
        self._l = numpy.cos(d) * numpy.sin(dra)
        self._m = numpy.sin(d) * numpy.cos(d0) - numpy.cos(d) * numpy.sin(d0) * numpy.cos(dra)
        self._n = numpy.sin(d) * numpy.sin(d0) + numpy.cos(d) * numpy.cos(d0) * numpy.cos(dra)

    def represent_as(self, _cls):
        return _Cartesian(self._n, self._l, self._m)


class SkyCoord:
    """ICRS direction with the few methods the path uses."""

    def __init__(self, ra_rad, dec_rad):
        self.ra, self.dec = Angle(ra_rad), Angle(dec_rad)

    @classmethod
    def from_deg(cls, ra_deg, dec_deg):
        return cls(numpy.radians(ra_deg), numpy.radians(dec_deg))

    def separation(self, other):
        dl = other.ra.rad - self.ra.rad
        d1, d2 = self.dec.rad, other.dec.rad
        num = numpy.hypot(numpy.cos(d2) * numpy.sin(dl),
                          numpy.cos(d1) * numpy.sin(d2) - numpy.sin(d1) * numpy.cos(d2) * numpy.cos(dl))
        den = numpy.sin(d1) * numpy.sin(d2) + numpy.cos(d1) * numpy.cos(d2) * numpy.cos(dl)
        return Angle(numpy.arctan2(num, den))

    def skyoffset_frame(self):
        return _OffsetFrame(self)

    def transform_to(self, frame):
        return _OffsetCoord(self, frame.origin)

    def __repr__(self):
        return f"SkyCoord(ra={self.ra.deg:.9f} deg, dec={self.dec.deg:.9f} deg)"


class PolarisationFrame:
    NPOL = {"stokesI": 1, "stokesIQ": 2, "stokesIV": 2, "linearnp": 2, "circularnp": 2,
            "stokesIQUV": 4, "linear": 4, "circular": 4}

    def __init__(self, name):
        if name not in self.NPOL:
            raise ValueError(f"Unknown polarisation frame {name}")
        self.type = name

    @property
    def npol(self):
        return self.NPOL[self.type]

    def __eq__(self, other):
        return isinstance(other, PolarisationFrame) and other.type == self.type

    def __hash__(self):
        return hash(self.type)


class _WcsParams:
    def __init__(self, crval, crpix, cdelt, ctype):
        self.crval = numpy.array(crval, dtype=float)
        self.crpix = numpy.array(crpix, dtype=float)
        self.cdelt = numpy.array(cdelt, dtype=float)
        self.ctype = list(ctype)


class WCS:
    """4-axis image WCS (RA---SIN, DEC--SIN, STOKES, FREQ), linear on the last two axes."""

    def __init__(self, crval, crpix, cdelt, ctype=("RA---SIN", "DEC--SIN", "STOKES", "FREQ")):
        self.wcs = _WcsParams(crval, crpix, cdelt, ctype)

    def sub(self, axes):
        idx = [a - 1 for a in axes]
        w = self.wcs
        return WCS(w.crval[idx], w.crpix[idx], w.cdelt[idx], [w.ctype[i] for i in idx])

    def wcs_world2pix(self, *args):
        *world, origin = args
        w = self.wcs
        return [(numpy.asarray(v, dtype=float) - w.crval[i]) / w.cdelt[i] + w.crpix[i] - (1 - origin)
                for i, v in enumerate(world)]

    def pixel_to_skycoord(self, xp, yp, origin=1):
        """SIN deprojection of pixel (xp, yp)."""
        w = self.wcs
        ra0, dec0 = numpy.radians(w.crval[0]), numpy.radians(w.crval[1])
        l = numpy.radians(w.cdelt[0] * (xp + (1 - origin) - w.crpix[0]))
        m = numpy.radians(w.cdelt[1] * (yp + (1 - origin) - w.crpix[1]))
        n = numpy.sqrt(max(0.0, 1.0 - l * l - m * m))
        dec = numpy.arcsin(m * numpy.cos(dec0) + n * numpy.sin(dec0))
        ra = ra0 + numpy.arctan2(l, n * numpy.cos(dec0) - m * numpy.sin(dec0))
        return SkyCoord(ra, dec)


def pixel_to_skycoord(xp, yp, wcs, origin=0):
    """Stand-in for astropy.wcs.utils.pixel_to_skycoord on the WCS class above."""
    return wcs.pixel_to_skycoord(xp, yp, origin=origin)


class _Var:
    """One data variable: ``.data`` (numpy array) and ``.shape``."""

    def __init__(self, data):
        self.data = data

    @property
    def shape(self):
        return self.data.shape

    def __array__(self, dtype=None, copy=None):
        return numpy.asarray(self.data, dtype=dtype)

    def __len__(self):
        return len(self.data)

    def __getitem__(self, key):
        return self.data[key]

    def __setitem__(self, key, value):
        self.data[key] = value


class _Dataset:
    def __init__(self, variables, attrs):
        self._vars = {k: _Var(v) for k, v in variables.items()}
        self.attrs = dict(attrs)

    def __getitem__(self, key):
        return self._vars[key]

    def __contains__(self, key):
        return key in self._vars

    def _copy_vars(self, zero_keys=()):
        out = {}
        for k, v in self._vars.items():
            # numpy.zeros maps untouched zero pages (calloc); zeros_like would write all of them
            out[k] = (numpy.zeros(numpy.shape(v.data), dtype=numpy.asarray(v.data).dtype) if k in zero_keys
                      else _deep_copy(v.data))
        return out


class _ImageAccessor:
    def __init__(self, im):
        self._im = im

    @property
    def wcs(self):
        return self._im.attrs["wcs"]

    @property
    def polarisation_frame(self):
        return self._im.attrs["polarisation_frame"]

    @property
    def phasecentre(self):
        w = self.wcs.wcs
        return SkyCoord.from_deg(w.crval[0], w.crval[1])

    def is_canonical(self):
        ct = self.wcs.wcs.ctype
        return (ct[0] == "RA---SIN" and ct[1] == "DEC--SIN" and ct[2] == "STOKES"
                and ct[3] == "FREQ")


class Image(_Dataset):
    """pixels [nchan, npol, ny, nx] float64 + WCS + polarisation frame."""

    def __init__(self, pixels, wcs, polarisation_frame):
        super().__init__({"pixels": numpy.asarray(pixels)},
                         {"wcs": wcs, "polarisation_frame": polarisation_frame})

    @property
    def image_acc(self):
        return _ImageAccessor(self)

    def copy(self, deep=True, zero=False):
        v = self._copy_vars(("pixels",) if zero else ())
        return Image(v["pixels"], _copy.deepcopy(self.attrs["wcs"]), self.attrs["polarisation_frame"])


def create_image(npixel, cellsize, phasecentre, frequency, channel_bandwidth=None,
                 polarisation_frame=None, nchan=None):
    """Canonical template image the way datamodels' create_image lays it out: cdelt = (-cell,
    +cell) degrees, crpix = npixel//2 + 1, crval = phase centre, channel axis linear."""
    pf = polarisation_frame or PolarisationFrame("stokesI")
    frequency = numpy.atleast_1d(numpy.asarray(frequency, dtype=float))
    nchan = nchan or len(frequency)
    if channel_bandwidth is None:
        channel_bandwidth = frequency[1] - frequency[0] if len(frequency) > 1 else 1e6
    cell = numpy.degrees(cellsize)
    wcs = WCS(crval=[phasecentre.ra.deg, phasecentre.dec.deg, 1.0, frequency[0]],
              crpix=[npixel // 2 + 1, npixel // 2 + 1, 1.0, 1.0],
              cdelt=[-cell, cell, 1.0, float(channel_bandwidth)])
    return Image(numpy.zeros((nchan, pf.npol, npixel, npixel)), wcs, pf)


class _VisAccessor:
    def __init__(self, vis):
        self._v = vis

    @property
    def polarisation_frame(self):
        return self._v.attrs["polarisation_frame"]

    @property
    def flagged_vis(self):
        return self._v["vis"].data * (1 - self._v["flags"].data)

    @property
    def flagged_imaging_weight(self):
        return self._v["imaging_weight"].data * (1 - self._v["flags"].data)

    @property
    def flagged_weight(self):
        return self._v["weight"].data * (1 - self._v["flags"].data)

    @property
    def u(self):
        return _Var(self._v["uvw"].data[..., 0])

    @property
    def v(self):
        return _Var(self._v["uvw"].data[..., 1])

    @property
    def uvw_lambda(self):
        k = self._v["frequency"].data / C_M_S
        # values of einsum("tbs,k->tbks") (visibility/base.py:54-55), stored so that the
        # [..., chan, axis] slices the reference takes (gridding.py:49-51) are contiguous:
        # numpy >= 2 hands out read-only copies for .flat of non-contiguous slices, which the
        # reference's nan_to_num call cannot write to
        lam = numpy.ascontiguousarray(numpy.einsum("tbs,k->kstb", self._v["uvw"].data, k))
        return _Var(lam.transpose(2, 3, 0, 1))


class Visibility(_Dataset):
    """vis [ntimes, nbaselines, nchan, npol] complex128, uvw [ntimes, nbaselines, 3] metres,
    weight / imaging_weight float64, flags int, frequency [nchan] Hz."""

    def __init__(self, vis, uvw, frequency, phasecentre, polarisation_frame, weight=None,
                 imaging_weight=None, flags=None):
        vis = numpy.asarray(vis, dtype=numpy.complex128)
        weight = numpy.ones(vis.shape) if weight is None else numpy.asarray(weight, dtype=float)
        imaging_weight = weight.copy() if imaging_weight is None else numpy.asarray(imaging_weight, dtype=float)
        flags = numpy.zeros(vis.shape, dtype=int) if flags is None else numpy.asarray(flags)
        super().__init__(
            {"vis": vis, "uvw": numpy.asarray(uvw, dtype=float), "weight": weight,
             "imaging_weight": imaging_weight, "flags": flags,
             "frequency": numpy.asarray(frequency, dtype=float)},
            {"phasecentre": phasecentre, "polarisation_frame": polarisation_frame})

    vis = property(lambda self: self["vis"])
    uvw = property(lambda self: self["uvw"])
    frequency = property(lambda self: self["frequency"])
    flags = property(lambda self: self["flags"])
    weight = property(lambda self: self["weight"])
    imaging_weight = property(lambda self: self["imaging_weight"])

    @property
    def phasecentre(self):
        return self.attrs["phasecentre"]

    @property
    def visibility_acc(self):
        return _VisAccessor(self)

    def copy(self, deep=True, zero=False):
        v = self._copy_vars(("vis",) if zero else ())
        out = Visibility(v["vis"], v["uvw"], v["frequency"], self.attrs["phasecentre"],
                         self.attrs["polarisation_frame"], v["weight"], v["imaging_weight"],
                         v["flags"])
        return out


class _GridAccessor:
    def __init__(self, gd):
        self._gd = gd

    @property
    def griddata_wcs(self):
        return self._gd.attrs["griddata_wcs"]

    @property
    def polarisation_frame(self):
        return self._gd.attrs["polarisation_frame"]

    @property
    def shape(self):
        return self._gd["pixels"].data.shape


class GridData(_Dataset):
    """pixels [nchan, npol, nv, nu] complex128 + UU/VV/STOKES/FREQ WCS (the attributes
    grid_data/gridding.py touches: griddata_acc.{griddata_wcs, polarisation_frame, shape})."""

    def __init__(self, pixels, griddata_wcs, polarisation_frame):
        super().__init__({"pixels": numpy.asarray(pixels)},
                         {"griddata_wcs": griddata_wcs, "polarisation_frame": polarisation_frame})

    @property
    def griddata_acc(self):
        return _GridAccessor(self)


def create_griddata_from_image(im, polarisation_frame=None, ft_types=None):
    """Stand-in for datamodels' create_griddata_from_image (external ska_sdp_datamodels 0.2.1,
    restated from its published behaviour): a grid of the image's shape whose UU / VV axes have
    cdelt = 1 / (n * cdelt_image[rad]) -- negative along u because RA decreases with x --
    crpix = n // 2 + 1, crval = 0."""
    ft_types = ft_types or ["UU", "VV"]
    nchan, npol, ny, nx = im["pixels"].data.shape
    w = im.image_acc.wcs.wcs
    d2r = numpy.pi / 180.0
    cdelt = [1.0 / (nx * w.cdelt[0] * d2r), 1.0 / (ny * w.cdelt[1] * d2r), w.cdelt[2], w.cdelt[3]]
    gwcs = WCS(crval=[0.0, 0.0, w.crval[2], w.crval[3]],
               crpix=[nx // 2 + 1, ny // 2 + 1, w.crpix[2], w.crpix[3]],
               cdelt=cdelt, ctype=[ft_types[0], ft_types[1], w.ctype[2], w.ctype[3]])
    pf = polarisation_frame or im.image_acc.polarisation_frame
    return GridData(numpy.zeros((nchan, pf.npol, ny, nx), dtype=complex), gwcs, pf)


class SkyComponent:
    """direction (SkyCoord), frequency [nchan], flux [nchan, npol], polarisation_frame: the
    attributes imaging/dft.py:59-118 reads."""

    def __init__(self, direction, frequency, flux, polarisation_frame, name="", shape="Point"):
        self.direction = direction
        self.frequency = numpy.atleast_1d(numpy.asarray(frequency, dtype=float))
        self.flux = numpy.asarray(flux)
        self.polarisation_frame = polarisation_frame
        self.name, self.shape = name, shape

    def copy(self):
        return SkyComponent(self.direction, self.frequency.copy(), self.flux.copy(),
                            self.polarisation_frame, self.name, self.shape)


class SkyModel:
    """image, components, mask, gaintable: the attributes sky_model/skymodel_imaging.py reads."""

    def __init__(self, image=None, components=None, gaintable=None, mask=None, fixed=False):
        self.image = image
        self.components = [] if components is None else list(components)
        self.gaintable = gaintable
        self.mask = mask
        self.fixed = fixed


def time_slices(vis):
    """[(index, Visibility of one time sample)]: what ``vis.groupby("time", squeeze=False)``
    yields for the xarray data model."""
    out = []
    for t in range(vis["vis"].data.shape[0]):
        sl = slice(t, t + 1)
        out.append((t, Visibility(vis["vis"].data[sl].copy(), vis["uvw"].data[sl].copy(),
                                  vis["frequency"].data.copy(), vis.attrs["phasecentre"],
                                  vis.attrs["polarisation_frame"], vis["weight"].data[sl].copy(),
                                  vis["imaging_weight"].data[sl].copy(), vis["flags"].data[sl].copy())))
    return out


def concatenate_visibility(vis_list, dim="time"):
    """Stand-in for datamodels' concatenate_visibility along time."""
    assert dim == "time"
    first = vis_list[0]
    cat = lambda k: numpy.concatenate([v[k].data for v in vis_list], axis=0)  # noqa: E731
    return Visibility(cat("vis"), cat("uvw"), first["frequency"].data.copy(),
                      first.attrs["phasecentre"], first.attrs["polarisation_frame"], cat("weight"),
                      cat("imaging_weight"), cat("flags"))
