"""Host-side kernel mathematics of the library (b2g_kernel_select / _eval / _corr): the
piecewise polynomial reproduces the ES kernel, its tabulated Fourier transform matches a direct
quadrature, and the (W, sigma) selection honours the error budget."""
import ctypes
import json
import os

import numpy as np
import pytest

from conftest import ROOT


def _pd(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def _rows():
    with open(os.path.join(ROOT, "ska_sdp_func_python_b200", "kernel_db.json")) as fh:
        return json.load(fh)


def es_phi(t, W, beta, e0):
    return np.exp(beta * W * (np.clip(1 - t * t, 0, None) ** e0 - 1.0))


@pytest.mark.parametrize("W", [4, 7, 8, 12, 16])
def test_polynomial_reproduces_es_kernel(built, W):
    row = [r for r in _rows() if r["W"] == W and abs(r["sigma"] - 1.5) < 1e-9][0]
    t = np.linspace(-1, 1, 4001)
    out = np.empty_like(t)
    assert built.b2g_kernel_eval(W, row["beta"], row["e0"], _pd(t), len(t), _pd(out)) == 0
    err = np.abs(out - es_phi(t, W, row["beta"], row["e0"])).max()
    # the polynomial IS the kernel; its fit error only needs to sit below the aliasing error
    assert err < max(0.5 * row["eps"], 2e-15), (err, row["eps"])


@pytest.mark.parametrize("W", [5, 8, 14])
def test_kernel_fourier_transform(built, W):
    row = [r for r in _rows() if r["W"] == W and abs(r["sigma"] - 1.5) < 1e-9][0]
    x = np.linspace(0, 0.5, 33)
    out = np.empty_like(x)
    assert built.b2g_kernel_corr(W, row["beta"], row["e0"], _pd(x), len(x), _pd(out)) == 0
    g, w = np.polynomial.legendre.leggauss(400)
    cells = 0.5 * W * g  # integration variable in cells, kernel support [-W/2, W/2]
    ref = (0.5 * W * w * es_phi(g, W, row["beta"], row["e0"])) @ np.cos(2 * np.pi * np.outer(cells, x))
    # psi-hat is the transform of the POLYNOMIAL kernel: it may differ from the exact ES
    # transform by the polynomial fit error, which sits below the aliasing error
    assert np.abs(out - ref).max() < 0.1 * row["eps"] * ref.max()


@pytest.mark.parametrize("prec,eps", [(0, 1e-3), (0, 1e-5), (1, 1e-7), (1, 1e-12)])
@pytest.mark.parametrize("ws", [0, 1])
def test_selection_meets_budget(built, prec, eps, ws):
    out = (ctypes.c_double * 5)()
    assert built.b2g_kernel_select(eps, ws, prec, 2048, 2048, 1e8, out) == 0
    W, sigma, eps_k = int(out[0]), out[1], out[4]
    assert 4 <= W <= (8 if prec == 0 else 16)
    assert 1.15 <= sigma <= 2.6
    assert eps_k * 1.75 * np.sqrt(3 if ws else 2) <= eps


def test_selection_rejects_unreachable_epsilon(built):
    out = (ctypes.c_double * 5)()
    assert built.b2g_kernel_select(1e-9, 1, 0, 512, 512, 1e6, out) != 0
    assert b"single precision" in built.b2g_last_error()
    assert built.b2g_kernel_select(1e-15, 1, 1, 512, 512, 1e6, out) != 0
    assert built.b2g_kernel_select(-1.0, 1, 1, 512, 512, 1e6, out) != 0


def test_table_is_monotonic():
    """More support or more oversampling never makes the tabulated error worse."""
    rows = _rows()
    by = {(r["W"], r["sigma"]): r["eps"] for r in rows}
    for (W, s), e in by.items():
        if e < 1e-13:  # fp64 quadrature floor of the table generator
            continue
        if (W + 1, s) in by:
            assert by[(W + 1, s)] < e * 1.05
        s2 = round(s + 0.05, 2)
        if (W, s2) in by:
            assert by[(W, s2)] < e * 1.05
