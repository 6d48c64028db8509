"""Polarisation-frame conversion matrices for the gridder wrapper.

The reference converts visibilities between polarisation frames with
``ska_sdp_datamodels.science_data_model.polarisation_functions.convert_pol_frame``
(external package ska-sdp-datamodels 0.2.1, /root/reference/poetry.lock:1152-1153; call sites
/root/reference/src/ska_sdp_func_python/imaging/ng.py:131-136, :193-198).  That function is a
constant complex matrix applied along the polarisation axis; the GPU wrapper fuses the matrix
into its load / store kernels, so all it needs is the matrix.

``pol_matrix(ipf, opf)`` asks the real package for it when it is importable (by converting the
unit vectors), and otherwise uses the table below, which restates the package's published
conventions (external package, restated from memory of its public source; Thompson, Moran &
Swenson eq. 4.58/4.59):  XX = I + Q, XY = U + iV, YX = U - iV, YY = I - Q;
RR = I + V, RL = -iQ + U, LR = -iQ - U, LL = I - V.
"""
import numpy

_I = 1j

# opf <- ipf : matrix M with  out = M @ in
_STOKES_TO_LINEAR = numpy.array([[1, 1, 0, 0], [0, 0, 1, _I], [0, 0, 1, -_I], [1, -1, 0, 0]],
                                dtype=complex)
_LINEAR_TO_STOKES = numpy.array([[0.5, 0, 0, 0.5], [0.5, 0, 0, -0.5], [0, 0.5, 0.5, 0],
                                 [0, -0.5 * _I, 0.5 * _I, 0]], dtype=complex)
_STOKES_TO_CIRCULAR = numpy.array([[1, 0, 0, 1], [0, -_I, 1, 0], [0, -_I, -1, 0], [1, 0, 0, -1]],
                                  dtype=complex)
_CIRCULAR_TO_STOKES = numpy.linalg.inv(_STOKES_TO_CIRCULAR)
_NP_TO_STOKES2 = numpy.array([[0.5, 0.5], [0.5, -0.5]], dtype=complex)
_STOKES2_TO_NP = numpy.array([[1, 1], [1, -1]], dtype=complex)

_TABLE = {
    ("linear", "stokesIQUV"): _LINEAR_TO_STOKES,
    ("stokesIQUV", "linear"): _STOKES_TO_LINEAR,
    ("circular", "stokesIQUV"): _CIRCULAR_TO_STOKES,
    ("stokesIQUV", "circular"): _STOKES_TO_CIRCULAR,
    ("linearnp", "stokesIQ"): _NP_TO_STOKES2,
    ("stokesIQ", "linearnp"): _STOKES2_TO_NP,
    ("circularnp", "stokesIV"): _NP_TO_STOKES2,
    ("stokesIV", "circularnp"): _STOKES2_TO_NP,
}


def _frame_type(pf):
    return pf if isinstance(pf, str) else pf.type


def _frame_npol(pf):
    if not isinstance(pf, str) and hasattr(pf, "npol"):
        return int(pf.npol)
    return {"stokesI": 1, "stokesIQ": 2, "stokesIV": 2, "linearnp": 2, "circularnp": 2}.get(
        _frame_type(pf), 4)


def pol_matrix_table(ipf, opf):
    """Matrix from the restated table (no dependency on ska_sdp_datamodels)."""
    it, ot = _frame_type(ipf), _frame_type(opf)
    if it == ot:
        return numpy.eye(_frame_npol(ipf), dtype=complex)
    try:
        return _TABLE[(it, ot)].copy()
    except KeyError:
        raise ValueError(f"Unknown polarisation conversion: {it} to {ot}") from None


def pol_matrix(ipf, opf):
    """Complex matrix M [npol_out, npol_in] with  convert_pol_frame(v, ipf, opf) == M @ v."""
    try:
        from ska_sdp_datamodels.science_data_model.polarisation_functions import (
            convert_pol_frame,
        )
    except ImportError:
        return pol_matrix_table(ipf, opf)
    if isinstance(ipf, str) or isinstance(opf, str):
        return pol_matrix_table(ipf, opf)
    n = _frame_npol(ipf)
    basis = numpy.eye(n, dtype=complex).reshape(1, n, n)  # [1, vector index, polaxis]
    out = numpy.asarray(convert_pol_frame(basis, ipf, opf, polaxis=2))
    return numpy.ascontiguousarray(out[0].T)  # column i = image of unit vector i
