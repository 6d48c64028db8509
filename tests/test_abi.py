"""The C-ABI boundary (include/b200_wgridder.h): the library loads, exports every declared
symbol, validates arguments, and fails loudly -- never falls back -- without a CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200_wgridder.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2g_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(built):
    syms = _declared_symbols()
    assert len(syms) >= 16
    for s in syms:
        assert hasattr(built, s), f"{s} declared in include/b200_wgridder.h but not exported"


def test_python_binding_lists_every_symbol(built):
    from ska_sdp_func_python_b200 import _lib
    assert sorted(_lib.EXPORTS) == _declared_symbols()


def test_no_reference_or_oracle_in_product():
    """The product never routes through the oracle or a CPU path."""
    pkg = os.path.join(ROOT, "ska_sdp_func_python_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "liboracle" not in src and "wgridder_cpu" not in src, f


def test_version_and_error_strings(built):
    assert b"b200-wgridder" in built.b2g_version()
    assert isinstance(built.b2g_last_error(), bytes)


def test_argument_validation_precedes_device_use(built):
    """Bad arguments are rejected with a message (no GPU needed for these paths)."""
    from ska_sdp_func_python_b200 import _lib
    h = ctypes.c_void_p()
    uvw = np.zeros((4, 3))
    freq = np.array([1e8])
    args = lambda npx, eps, prec: (ctypes.byref(h), 4, 1, uvw.ctypes.data, freq.ctypes.data,  # noqa: E731
                                   _lib.HOST, npx, 64, 1e-3, 1e-3, 0, 0, eps, 1, prec, -1, 0, None)
    assert built.b2g_plan_create(*args(63, 1e-5, _lib.F32)) != 0
    assert b"even" in built.b2g_last_error()
    assert built.b2g_plan_create(*args(64, 1e-5, 7)) != 0
    assert b"precision" in built.b2g_last_error()
    assert built.b2g_plan_invert(None, None, None, 0, 0, None, 0, 0) != 0
    assert b"NULL" in built.b2g_last_error()


def test_fails_loudly_without_gpu(built):
    """No CPU fallback: with no CUDA device every compute entry point raises."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from ska_sdp_func_python_b200 import wgridder
    uvw = np.random.default_rng(0).normal(size=(10, 3))
    with pytest.raises(wgridder.WgridderError):
        wgridder.ms2dirty(uvw, np.array([1e8]), np.ones((10, 1), complex), None, 64, 64, 1e-3, 1e-3,
                          0, 0, 1e-5, True)
    with pytest.raises(wgridder.WgridderError):
        wgridder.dirty2ms(uvw, np.array([1e8]), np.ones((64, 64)), None, 1e-3, 1e-3, 0, 0, 1e-5, True)
