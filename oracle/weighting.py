"""CPU oracle (TEST INFRASTRUCTURE ONLY) for the density-weighting row N1 of SURVEY.md section 8f.

numpy restatement, at the array level, of
  * spatial_mapping                     /root/reference/src/ska_sdp_func_python/grid_data/gridding.py:146-157
  * grid_visibility_weight_to_griddata  gridding.py:258-334
  * griddata_visibility_reweight        gridding.py:362-499
  * taper_visibility_gaussian / _tukey  imaging/weighting.py:68-140 (tukey_filter:
                                        util/array_functions.py:85-99)
Pinned against outputs of those reference functions themselves, run in the build container by
tests/golden/make_golden_weight.py (tests/golden/weight_case_*.npz; tests/test_oracle_weighting.py).
Only tests/ may import this module; the product path is csrc/weight_kernels.cu.

Array conventions: uvw [nrow, 3] metres; freq [nchan] Hz; weight / flags / imaging_weight
[nrow, nchan, npol]; grid [nchan_img, npol, ny, nx] float64; the UU / VV axes are linear with
(cdelt, crpix 1-based, crval 0): pixel = round(x / cdelt + crpix - 1).
"""
import numpy

C_M_S = 299792458.0


def cells(uvw, freq, chan, cdelt_u, cdelt_v, crpix_u, crpix_v):
    """(pu, pv, pu_conj, pv_conj) of every row at one channel (gridding.py:49-57, :146-157)."""
    k = freq[chan] / C_M_S
    u = numpy.nan_to_num(uvw[:, 0] * k)
    v = numpy.nan_to_num(uvw[:, 1] * k)

    def pix(x, cdelt, crpix):
        return numpy.round(x / cdelt + crpix - 1).astype("int")

    return (pix(u, cdelt_u, crpix_u), pix(v, cdelt_v, crpix_v),
            pix(-u, cdelt_u, crpix_u), pix(-v, cdelt_v, crpix_v))


def _ingrid(pu, pv, puc, pvc, nx, ny):
    return ((pv >= 0) & (pv < ny) & (pu >= 0) & (pu < nx) &
            (pvc >= 0) & (pvc < ny) & (puc >= 0) & (puc < nx))


def grid_weights(uvw, freq, weight, flags, vis_to_im, nchan_img, nx, ny, cdelt_u, cdelt_v,
                 crpix_u, crpix_v):
    """Returns (grid [nchan_img, npol, ny, nx], sumwt [nchan_img, npol]); sample order and
    addition order of the reference loops (row-serial, cell then conjugate cell)."""
    nrow, nchan, npol = weight.shape
    fw = weight * (1 - flags)
    grid = numpy.zeros((nchan_img, npol, ny, nx))
    sumwt = numpy.zeros((nchan_img, npol))
    for chan in range(nchan):
        imchan = vis_to_im[chan]
        pu, pv, puc, pvc = cells(uvw, freq, chan, cdelt_u, cdelt_v, crpix_u, crpix_v)
        ok = _ingrid(pu, pv, puc, pvc, nx, ny)
        idx = numpy.empty(2 * int(ok.sum()), dtype=numpy.int64)
        idx[0::2] = pv[ok] * nx + pu[ok]
        idx[1::2] = pvc[ok] * nx + puc[ok]
        for pol in range(npol):
            w = fw[ok, chan, pol]
            numpy.add.at(grid[imchan, pol].reshape(-1), idx, numpy.repeat(w, 2))
            # sequential accumulation, as the reference's "sumwt += w * 2" per row
            sumwt[imchan, pol] = numpy.add.accumulate(
                numpy.concatenate(([sumwt[imchan, pol]], w * 2)))[-1]
    return grid, sumwt


def reweight(uvw, freq, weight, flags, imaging_weight, vis_to_im, grid, cdelt_u, cdelt_v, crpix_u,
             crpix_v, weighting="uniform", robustness=0.0, sumwt=None):
    """New imaging_weight [nrow, nchan, npol] (gridding.py:362-499)."""
    assert weighting in ("natural", "uniform", "robust")
    if weighting == "natural":
        return numpy.array(weight, copy=True)
    nrow, nchan, npol = weight.shape
    ny, nx = grid.shape[2], grid.shape[3]
    fw = weight * (1 - flags)
    out = imaging_weight * (1 - flags)
    f2 = None
    if weighting == "robust":
        sumlocwt = numpy.sum(grid ** 2)
        total = numpy.sum(fw) * 2 if sumwt is None else numpy.sum(sumwt)
        f2 = (5.0 * numpy.power(10.0, -robustness)) ** 2 * total / sumlocwt
    for chan in range(nchan):
        imchan = vis_to_im[chan]
        pu, pv, puc, pvc = cells(uvw, freq, chan, cdelt_u, cdelt_v, crpix_u, crpix_v)
        ok = _ingrid(pu, pv, puc, pvc, nx, ny)
        for pol in range(npol):
            g = numpy.zeros(nrow)
            g[ok] = grid[imchan, pol, pv[ok], pu[ok]]
            col = numpy.zeros(nrow)
            pos = ok & (g > 0.0)
            if weighting == "uniform":
                col[pos] = fw[pos, chan, pol] / g[pos]
            else:
                col[pos] = fw[pos, chan, pol] / (1 + f2 * g[pos])
            out[:, chan, pol] = col
    return out


def robust_f2(grid, sumwt, robustness):
    return (5.0 * numpy.power(10.0, -robustness)) ** 2 * numpy.sum(sumwt) / numpy.sum(grid ** 2)


def taper_gaussian(uvw, freq, flags, imaging_weight, beam):
    scale_factor = numpy.pi ** 2 * beam ** 2 / (4.0 * numpy.log(2.0))
    out = numpy.array(imaging_weight, copy=True)
    fimw = imaging_weight * (1 - flags)
    for chan, f in enumerate(freq):
        wave = C_M_S / f
        uvdistsq = (uvw[:, 0] ** 2 + uvw[:, 1] ** 2) / wave ** 2
        out[:, chan, :] = fimw[:, chan, :] * numpy.exp(-scale_factor * uvdistsq)[:, None]
    return out


def tukey_filter(x, r):
    if 0.0 <= x < r / 2.0:
        return 0.5 * (1.0 + numpy.cos(2.0 * numpy.pi * (x - r / 2.0) / r))
    if 1 - r / 2.0 <= x <= 1.0:
        return 0.5 * (1.0 + numpy.cos(2.0 * numpy.pi * (x - 1 + r / 2.0) / r))
    return 1.0


def taper_tukey(uvw, freq, flags, imaging_weight, tukey):
    out = numpy.array(imaging_weight, copy=True)
    fimw = imaging_weight * (1 - flags)
    for chan, f in enumerate(freq):
        wave = C_M_S / f
        uvdist = numpy.sqrt(uvw[:, 0] ** 2 + uvw[:, 1] ** 2) / wave
        uvdist = uvdist / numpy.max(uvdist)
        wt = numpy.array([tukey_filter(x, tukey) for x in uvdist])
        out[:, chan, :] = fimw[:, chan, :] * wt[:, None]
    return out
