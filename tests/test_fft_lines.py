"""Host check of the line-FFT core of the hand-written plane transform (csrc/fft_inplace.cuh: static
plans, in-place Stockham stages staged through registers, padded layouts, generated butterflies,
three-table twiddle scheme): the header is __host__ __device__, tests/native/test_fft_inplace_host.cpp
runs the same templates on one host thread -- the read phase of every "thread" before any write phase,
as the barriers order them on the device -- against a direct O(N^2) DFT in double precision.  No GPU
needed."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
PLANS = {1600: 40, 3200: 80, 6400: 160, 12800: 320}       # line length -> threads


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("fft") / "test_fft_inplace_host")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe,
                           os.path.join(HERE, "native", "test_fft_inplace_host.cpp")])
    return exe


def test_line_fft_against_direct_dft(harness):
    out = subprocess.run([harness], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    rows = [line.split() for line in out.stdout.strip().splitlines()]
    assert {int(r[0]): int(r[1]) for r in rows} == PLANS
    for row in rows:
        efwd, einv = float(row[2]), float(row[3])
        assert efwd < 3e-7 and einv < 3e-7, row       # fp32 rounding level, both signs


def test_codelets_are_current():
    """fft_codelets.inc is the output of tools/gen_fft_codelets.py."""
    import importlib.util
    root = os.path.dirname(HERE)
    spec = importlib.util.spec_from_file_location("gen", os.path.join(root, "tools", "gen_fft_codelets.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    text = open(os.path.join(root, "ska_sdp_func_python_b200", "csrc", "fft_codelets.inc")).read()
    for R in gen.SIZES:
        assert gen.codelet(R) in text
