"""Host check of the line-FFT core of the hand-written plane transform (csrc/fft_lines.cuh: radix
schedule, Stockham index arithmetic, generated butterflies, two-level twiddle tables): the header is
__host__ __device__, tests/native/test_fft_host.cpp runs it on one host thread against a direct
O(N^2) DFT in double precision.  No GPU needed."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SIZES = [8, 12, 16, 24, 30, 32, 48, 64, 96, 128, 192, 200, 256, 288, 320, 384, 400, 512, 640, 768, 800,
         1024, 3200, 4096, 6400, 9600, 12800, 12288, 8192]


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("fft") / "test_fft_host")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe,
                           os.path.join(HERE, "native", "test_fft_host.cpp")])
    return exe


def test_line_fft_against_direct_dft(harness):
    out = subprocess.run([harness] + [str(n) for n in SIZES], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    rows = [line.split() for line in out.stdout.strip().splitlines()]
    assert len(rows) == len(SIZES)
    for n, row in zip(SIZES, rows):
        assert int(row[0]) == n and row[1] != "unsupported", row
        nstage = int(row[1])
        radices = [int(x) for x in row[2:2 + nstage]]
        prod = 1
        for r in radices:
            prod *= r
        assert prod == n and nstage <= 4
        efwd, einv = float(row[-2]), float(row[-1])
        assert efwd < 5e-7 and einv < 5e-7, row       # fp32 rounding level, both signs
    # the benchmark grids: three stages, odd first radix (conflict-free shared-memory writes)
    by_n = {int(r[0]): [int(x) for x in r[2:2 + int(r[1])]] for r in rows}
    assert by_n[12800] == [25, 32, 16] and by_n[3200][0] == 25 and by_n[6400][0] == 25


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
