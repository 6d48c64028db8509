"""b2g_sumwt_static: the allocation-free variant of b2g_sumwt (csrc/sum_static.cu; scratch in a
file-scope __device__ array, no cudaMalloc / cudaFree in the call) against b2g_sumwt through the
C ABI.  No counterpart in the reference; b2g_sumwt
itself mirrors numpy.sum(wgt) at /root/reference/src/ska_sdp_func_python/imaging/ng.py:258, :289.

Both entry points use the same fixed reduction tree, so they must agree bit for bit (the bound
of this file; the other tests hold b2g_sumwt to 1e-12 relative against numpy).  Two calls in a
row on different inputs: the shared scratch keeps the first call's partials, the second call must
not see them.  This file sorts last on purpose."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

F32, F64, HOST, DEVICE = 0, 1, 0, 1
MAX_N = 1 << 32      # B2G_SUMWT_STATIC_MAX_N


@pytest.fixture(scope="module")
def torch_dev(built):
    import torch
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    return torch, torch.device("cuda", 0)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _sumwt(lib, t, code):
    out = ctypes.c_double(0.0)
    rc = lib.b2g_sumwt(_ptr(t), code, DEVICE, t.numel(), ctypes.byref(out), None)
    assert rc == 0, lib.b2g_last_error().decode()
    return out.value


@pytest.mark.parametrize("dtype,code", [("float32", F32), ("float64", F64)])
def test_static_sum_equals_allocating_sum_twice_in_a_row(built, torch_dev, dtype, code):
    torch, dev = torch_dev
    rng = np.random.default_rng(7)
    # ragged sizes: fewer elements than partials, a non-multiple of 1024 * 256, and an empty input
    sizes = (1_000_003, 777, 3_500_000, 0)
    inputs = [torch.as_tensor(rng.uniform(0.0, 2.0, n).astype(dtype), device=dev) for n in sizes]
    outs = torch.full((len(sizes),), float("nan"), dtype=torch.float64, device=dev)
    want = [_sumwt(built, t, code) for t in inputs]
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    # all calls back to back on one stream, nothing read in between: every call after the first
    # starts from the previous call's partials in the shared scratch
    for i, t in enumerate(inputs):
        rc = built.b2g_sumwt_static(_ptr(t), code, t.numel(), ctypes.c_void_p(outs.data_ptr() + 8 * i),
                                    ctypes.c_void_p(stream))
        assert rc == 0, built.b2g_last_error().decode()
    got = outs.cpu().numpy()
    for i, n in enumerate(sizes):
        assert got[i] == want[i], (n, got[i], want[i])
        ref = float(inputs[i].double().sum().item())
        assert abs(got[i] - ref) <= 1e-12 * max(ref, 1.0), (n, got[i], ref)
    # and once more over the first input: same bits as the first time
    again = torch.zeros(1, dtype=torch.float64, device=dev)
    assert built.b2g_sumwt_static(_ptr(inputs[0]), code, sizes[0], _ptr(again), ctypes.c_void_p(stream)) == 0
    assert float(again.item()) == want[0]


def test_static_sum_allocates_nothing(built, torch_dev):
    """Free device memory as the driver reports it does not move over many calls (the allocating
    entry point is allowed to: its stream-ordered pool grows on first use)."""
    torch, dev = torch_dev
    w = torch.ones(1 << 20, dtype=torch.float32, device=dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    assert built.b2g_sumwt_static(_ptr(w), F32, w.numel(), _ptr(out), None) == 0   # set-up call
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    for _ in range(50):
        assert built.b2g_sumwt_static(_ptr(w), F32, w.numel(), _ptr(out), None) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0
    assert float(out.item()) == float(1 << 20)


def test_static_sum_rejects_bad_arguments(built, torch_dev):
    torch, dev = torch_dev
    w = torch.ones(16, dtype=torch.float32, device=dev)
    out = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
    # beyond the stated maximum: refused before anything is launched (the pointer is never read)
    assert built.b2g_sumwt_static(_ptr(w), F32, MAX_N + 1, _ptr(out), None) != 0
    assert b"B2G_SUMWT_STATIC_MAX_N" in built.b2g_last_error()
    assert built.b2g_sumwt_static(_ptr(w), F32, -1, _ptr(out), None) != 0
    assert built.b2g_sumwt_static(_ptr(w), 7, 16, _ptr(out), None) != 0
    assert built.b2g_sumwt_static(None, F32, 16, _ptr(out), None) != 0
    assert built.b2g_sumwt_static(_ptr(w), F32, 16, None, None) != 0
    torch.cuda.synchronize()
    assert float(out.item()) == -1.0
    assert built.b2g_sumwt_static(_ptr(w), F32, 16, _ptr(out), None) == 0
    assert float(out.item()) == 16.0
