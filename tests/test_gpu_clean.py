"""Hogbom CLEAN on the GPU (csrc/clean_kernels.cu through the C ABI): bit-exact against golden
vectors from the reference's own hogbom() and against the numpy oracle on a larger seeded image
(ties, windows, PSF smaller / larger than the image, early stop, niter exhaustion)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cl(built):
    import torch
    assert torch.cuda.is_available()
    from ska_sdp_func_python_b200 import cleaners
    return cleaners


@pytest.mark.parametrize("name", ["a", "b", "c", "d"])
def test_golden_bit_exact(cl, name):
    g = np.load(os.path.join(GOLDEN, "clean_case.npz"))
    gain, thresh, niter, frac = g[f"{name}_par"]
    window = g[f"{name}_window"] if f"{name}_window" in g.files else None
    dirty = g[f"{name}_dirty"].copy()
    comps, res = cl.hogbom(dirty, g[f"{name}_psf"], window, gain, thresh, int(niter), frac)
    assert isinstance(comps, np.ndarray) and comps.dtype == np.float64
    assert np.array_equal(comps, g[f"{name}_comps"])
    assert np.array_equal(res, g[f"{name}_res"])
    assert np.array_equal(dirty, g[f"{name}_dirty"])          # input not modified


def test_larger_image_against_oracle(cl):
    import torch
    from oracle import clean as oc
    rng = np.random.default_rng(8)
    n = 512
    yy, xx = np.meshgrid(np.arange(n) - n // 2, np.arange(n) - n // 2)
    r = np.hypot(xx, yy)
    psf = np.exp(-(r / 2.0) ** 2) + 0.05 * np.cos(r / 3.1) * np.exp(-r / 40.0) * (r > 0)
    dirty = 0.02 * rng.normal(size=(n, n))
    for _ in range(40):
        x, y = rng.integers(20, n - 20, 2)
        dirty[x - 10:x + 10, y - 10:y + 10] += rng.uniform(0.5, 4.0) * psf[n // 2 - 10:n // 2 + 10, n // 2 - 10:n // 2 + 10]
    dirty[100, 100] = dirty[300, 50] = 9.0                      # exact tie: the first one wins
    comps, res, it = oc.hogbom(dirty, psf, None, 0.2, 0.0, 400, 0.0)
    gc, gr = cl.hogbom(torch.from_numpy(dirty).cuda(), torch.from_numpy(psf).cuda(), None, 0.2, 0.0, 400, 0.0)
    assert gc.is_cuda and np.array_equal(gc.cpu().numpy(), comps) and np.array_equal(gr.cpu().numpy(), res)
    assert comps[100, 100] != 0 and it == 400
    # early stop through the fractional threshold, boolean window
    window = np.zeros((n, n), dtype=bool)
    window[64:448, 64:448] = True
    dirty[100, 100] = dirty[300, 50] = 0.0                      # back to PSF-shaped sources only
    comps, res, it = oc.hogbom(dirty, psf, window, 0.3, 0.0, 5000, 0.3)
    gc, gr = cl.hogbom(dirty, psf, window, 0.3, 0.0, 5000, 0.3)
    assert it < 5000
    assert np.array_equal(gc, comps) and np.array_equal(gr, res)
    dirty[100, 100] = dirty[300, 50] = 9.0
    v, (x, y) = cl.image_peak(dirty)
    assert v == 9.0 and (x, y) == (100, 100)


def test_argument_checks(cl):
    d = np.zeros((8, 8))
    p = np.zeros((8, 8))
    p[4, 4] = 1.0
    with pytest.raises(AssertionError):
        cl.hogbom(d, p, None, 2.5, 0.0, 10, 0.0)
    with pytest.raises(AssertionError):
        cl.hogbom(d, p, None, 0.1, 0.0, 0, 0.0)
    with pytest.raises(AssertionError):
        cl.hogbom(d, 0.5 * p, None, 0.1, 0.0, 10, 0.0)          # PSF peak not 1
    comps, res = cl.hogbom(d, p, None, 0.1, 0.0, 10, 0.0)       # all-zero image: niter no-op iterations
    assert not comps.any() and not res.any()


def test_hogbom_kernel_list(cl):
    """Image-level loop over channels and polarisations (image/deconvolution.py:543-643)."""
    from oracle import clean as oc
    from ska_sdp_func_python_b200 import minimal_models as mm
    g = np.load(os.path.join(GOLDEN, "clean_case.npz"))
    pc = mm.SkyCoord.from_deg(0.0, -30.0)
    pf = mm.PolarisationFrame("stokesIQ")
    dirty = mm.create_image(64, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    dirty["pixels"].data[0, 0] = g["a_dirty"]
    dirty["pixels"].data[0, 1] = 0.5 * g["a_dirty"][::-1]
    psf = mm.create_image(64, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    psf["pixels"].data[0, 0] = g["a_psf"]                      # pol 1 PSF all zero: skipped
    comps, resid = cl.hogbom_kernel_list([dirty], "test", [psf], None, gain=0.1, niter=300,
                                         fractional_threshold=0.01)
    assert len(comps) == len(resid) == 1 and isinstance(comps[0], mm.Image)
    assert np.array_equal(comps[0]["pixels"].data[0, 0], g["a_comps"])
    assert np.array_equal(resid[0]["pixels"].data[0, 0], g["a_res"])
    assert not comps[0]["pixels"].data[0, 1].any() and not resid[0]["pixels"].data[0, 1].any()
    psf["pixels"].data[0, 1] = g["a_psf"]
    win = mm.create_image(64, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    win["pixels"].data[0, :, 16:48, 16:48] = 1.0
    comps, resid = cl.hogbom_kernel_list([dirty], "test", [psf], [win], gain=0.2, niter=50)
    oc_c, oc_r, _ = oc.hogbom(dirty["pixels"].data[0, 1], g["a_psf"], win["pixels"].data[0, 1], 0.2,
                              0.0, 50, 0.01)
    assert np.array_equal(comps[0]["pixels"].data[0, 1], oc_c)
    assert np.array_equal(resid[0]["pixels"].data[0, 1], oc_r)
    with pytest.raises(ValueError):
        cl.hogbom_kernel_list([dirty], "test", [psf], None, gain=2.0)


# ---- multi-scale CLEAN -------------------------------------------------------------------------

def _ms_case(g, name):
    gain, thresh, niter, frac = g[f"{name}_par"]
    window = g[f"{name}_window"] if f"{name}_window" in g.files else None
    sens = g[f"{name}_sens"] if f"{name}_sens" in g.files else None
    return (g[f"{name}_dirty"], g[f"{name}_psf"], window, sens, gain, thresh, int(niter),
            list(g[f"{name}_scales"]), frac)


@pytest.mark.parametrize("name", ["a", "b", "c", "d", "e"])
def test_msclean_golden(cl, name):
    """Against the reference's own msclean(): the stacks come from cuFFT instead of numpy's FFT
    (rounding-level differences), the loop arithmetic is identical -> 1e-11 absolute on O(1) data."""
    g = np.load(os.path.join(GOLDEN, "msclean_case.npz"))
    args = _ms_case(g, name)
    dirty = args[0].copy()
    comps, res = cl.msclean(dirty, *args[1:])
    assert isinstance(comps, np.ndarray) and comps.shape == dirty.shape
    assert np.abs(comps - g[f"{name}_comps"]).max() < 1e-11
    assert np.abs(res - g[f"{name}_res"]).max() < 1e-11
    assert np.array_equal(dirty, args[0])


@pytest.mark.parametrize("name", ["a", "b", "c", "e"])
def test_msclean_loop_bit_exact(cl, name):
    """The minor cycle alone, through the C ABI, on the ORACLE's stacks: bit-exact components and
    residual stack (sensitivity applied twice in the argmax, strict '>' across scales, windows)."""
    import ctypes
    import torch
    from oracle import clean as oc
    from ska_sdp_func_python_b200 import _lib
    g = np.load(os.path.join(GOLDEN, "msclean_case.npz"))
    dirty, psf, window, sens, gain, thresh, niter, scales, frac = _ms_case(g, name)
    pmax, pstack, res, pss, coupling, wstack = oc.msclean_setup(dirty, psf, window, scales)
    absthresh = max(thresh, frac * np.abs(res[0]).max())
    dev = torch.device("cuda", 0)
    up = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
    ptr = lambda t: ctypes.c_void_p(0 if t is None else t.data_ptr())
    d_res, d_pss, d_ps, d_w, d_s = up(res), up(pss), up(pstack), up(wstack), up(sens)
    d_comps = torch.empty(dirty.shape, dtype=torch.float64, device=dev)
    cdiag = np.ascontiguousarray(np.diagonal(coupling))
    iters = ctypes.c_int64(0)
    _lib.check(_lib.lib().b2g_msclean(ptr(d_res), ptr(d_pss), ptr(d_ps), ptr(d_w), ptr(d_s),
                                      cdiag.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), len(scales),
                                      dirty.shape[0], dirty.shape[1], psf.shape[0], psf.shape[1], gain,
                                      absthresh, niter, ptr(d_comps), ctypes.byref(iters),
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    o_comps, done = oc.msclean_loop(res, pss, pstack, coupling, wstack, sens, psf.shape, gain, absthresh, niter)
    assert iters.value == done
    assert np.array_equal(d_comps.cpu().numpy(), o_comps)
    assert np.array_equal(d_res.cpu().numpy(), res)


def test_scalestack_matches_oracle(cl):
    from oracle import clean as oc
    for shape, scales in (((3, 33, 40), [0, 4, 9]), ((4, 64, 64), [0, 3, 10, 30]), ((1, 8, 8), [0])):
        assert np.array_equal(cl.create_scalestack(list(shape), scales), oc.scale_stack(shape, scales))
    with pytest.raises(ValueError):
        cl.create_scalestack([1, 16, 16], [30])
    for v in (0.0, 0.3, 0.75, 0.9, 1.0, 1.2, -0.5):
        assert cl.spheroidal_function(v) == oc.pswf(v)


def test_msclean_larger_image_and_device_tensors(cl):
    import torch
    from oracle import clean as oc
    rng = np.random.default_rng(21)
    n = 256
    yy, xx = np.meshgrid(np.arange(n) - n // 2, np.arange(n) - n // 2)
    r = np.hypot(xx, yy)
    psf = np.exp(-(r / 2.0) ** 2) + 0.05 * np.cos(r / 3.1) * np.exp(-r / 25.0) * (r > 0)
    dirty = 0.02 * rng.normal(size=(n, n))
    for _ in range(6):
        cx, cy, wd = rng.uniform(60, 190), rng.uniform(60, 190), rng.uniform(1.5, 8.0)
        dirty += rng.uniform(0.5, 2.0) * np.exp(-((xx.T - cx + n // 2) ** 2 + (yy.T - cy + n // 2) ** 2) / (2 * wd ** 2))
    scales = [0, 3, 10, 30]
    o_comps, o_res, done = oc.msclean(dirty, psf, None, None, 0.7, 0.0, 200, scales, 0.01)
    comps, res = cl.msclean(torch.from_numpy(dirty).cuda(), torch.from_numpy(psf).cuda(), None, None,
                            0.7, 0.0, 200, scales, 0.01)
    assert comps.is_cuda and res.is_cuda
    assert np.abs(comps.cpu().numpy() - o_comps).max() < 1e-11
    assert np.abs(res.cpu().numpy() - o_res).max() < 1e-11
    assert done > 10


def test_msclean_kernel_list(cl):
    from ska_sdp_func_python_b200 import minimal_models as mm
    g = np.load(os.path.join(GOLDEN, "msclean_case.npz"))
    pc = mm.SkyCoord.from_deg(0.0, -30.0)
    pf = mm.PolarisationFrame("stokesIQ")
    dirty = mm.create_image(64, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    dirty["pixels"].data[0, 0] = g["a_dirty"]
    dirty["pixels"].data[0, 1] = g["c_dirty"]
    psf = mm.create_image(64, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    psf["pixels"].data[0, 0] = g["a_psf"]
    comps, resid = cl.msclean_kernel_list([dirty], "test", [psf], None, gain=0.7, niter=60,
                                          fractional_threshold=0.01, scales=[0, 3, 10])
    assert np.abs(comps[0]["pixels"].data[0, 0] - g["a_comps"]).max() < 1e-11
    assert np.abs(resid[0]["pixels"].data[0, 0] - g["a_res"]).max() < 1e-11
    assert not comps[0]["pixels"].data[0, 1].any()            # zero PSF plane: skipped
    psf["pixels"].data[0, 1] = g["c_psf"]
    win = mm.create_image(64, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    win["pixels"].data[0, 1] = g["c_window"]
    win["pixels"].data[0, 0] = 1.0
    sens = mm.create_image(64, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    sens["pixels"].data[0, :] = g["c_sens"]
    comps, resid = cl.msclean_kernel_list([dirty], "test", [psf], [win], [sens], gain=0.5, niter=40,
                                          fractional_threshold=0.05, scales=[0, 3, 8, 14])
    assert np.abs(comps[0]["pixels"].data[0, 1] - g["c_comps"]).max() < 1e-11
    assert np.abs(resid[0]["pixels"].data[0, 1] - g["c_res"]).max() < 1e-11


# ---- multi-scale multi-frequency CLEAN ---------------------------------------------------------

def _mf_case(g, name):
    gain, thresh, niter, frac = g[f"{name}_par"]
    window = g[f"{name}_window"] if f"{name}_window" in g.files else None
    sens = g[f"{name}_sens"] if f"{name}_sens" in g.files else None
    return (g[f"{name}_dirty"], g[f"{name}_psf"], window, sens, gain, thresh, int(niter),
            list(g[f"{name}_scales"]), frac)


@pytest.mark.parametrize("name", ["a", "b", "c", "d", "e", "f"])
def test_msmfsclean_golden(cl, name):
    """Against the reference's own msmfsclean().  The stacks come from cuFFT and pass through the
    inverse moment Hessians (condition number up to ~1e4 here), so the tolerance is 1e-9 on O(1)
    data rather than the 1e-11 of msclean."""
    g = np.load(os.path.join(GOLDEN, "msmfsclean_case.npz"))
    args = _mf_case(g, name)
    model, res = cl.msmfsclean(*args, findpeak="CASA" if name in "ef" else "RASCIL")
    assert model.shape == args[0].shape and res.shape == args[0].shape
    assert np.abs(model - g[f"{name}_model"]).max() < 1e-9
    assert np.abs(res - g[f"{name}_res"]).max() < 1e-9


@pytest.mark.parametrize("name", ["a", "b", "c", "d", "e", "f"])
def test_msmfsclean_loop_on_oracle_stacks(cl, name):
    """The minor cycle alone, through the C ABI, on the ORACLE's stacks: same arithmetic as the
    reference up to the summation order inside numpy.einsum -> 1e-13, same iteration count."""
    import ctypes
    import torch
    from oracle import clean as oc
    from ska_sdp_func_python_b200 import _lib
    g = np.load(os.path.join(GOLDEN, "msmfsclean_case.npz"))
    dirty, psf, window, sens, gain, thresh, niter, scales, frac = _mf_case(g, name)
    pmax, pstack, smres, ssmm, hess, ihess, wstack = oc.msmfs_setup(dirty, psf, window, scales)
    ns, nm = smres.shape[:2]
    psfm = np.zeros((ns, ns, 2 * nm - 1) + psf.shape[1:])
    for t in range(nm):
        for q in range(nm):
            psfm[:, :, t + q] = ssmm[:, :, t, q]
    absthresh = max(thresh, frac * np.abs(smres[0, 0]).max())
    dev = torch.device("cuda", 0)
    up = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
    ptr = lambda t: ctypes.c_void_p(0 if t is None else t.data_ptr())
    d_res, d_psfm, d_ps, d_w, d_s = up(smres), up(psfm), up(pstack), up(wstack), up(sens)
    d_model = torch.empty((nm,) + dirty.shape[1:], dtype=torch.float64, device=dev)
    ih, hh = np.ascontiguousarray(ihess), np.ascontiguousarray(hess)
    casa = name in "ef"
    iters = ctypes.c_int64(0)
    _lib.check(_lib.lib().b2g_msmfsclean(ptr(d_res), ptr(d_psfm), ptr(d_ps), ptr(d_w), ptr(d_s),
                                         ih.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                                         hh.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), int(casa), ns, nm,
                                         dirty.shape[1], dirty.shape[2], psf.shape[1], psf.shape[2], gain,
                                         absthresh, niter, ptr(d_model), ctypes.byref(iters),
                                         ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    o_model, done = oc.msmfs_loop(smres, ssmm, pstack, ihess, wstack, sens, gain, absthresh, niter,
                                  hess, "CASA" if casa else "RASCIL")
    assert iters.value == done
    assert np.abs(d_model.cpu().numpy() - o_model).max() < 1e-13
    assert np.abs(d_res.cpu().numpy() - smres).max() < 1e-13


def test_msmfsclean_arguments(cl):
    g = np.load(os.path.join(GOLDEN, "msmfsclean_case.npz"))
    args = list(_mf_case(g, "b"))
    with pytest.raises(AssertionError):
        cl.msmfsclean(args[0], args[1][:3], *args[2:])           # psf must hold 2 x nmoment planes
    with pytest.raises(ValueError):
        cl.msmfsclean(args[0][0], args[1][0], *args[2:])         # 2-D input


# ---- complex Hogbom (Q + iU) -------------------------------------------------------------------

@pytest.mark.parametrize("name", ["a", "b", "c", "d"])
def test_hogbom_complex_golden_bit_exact(cl, name):
    """Against the reference's own hogbom_complex(): windows (bool / float), PSF smaller / larger than
    the image, a PSF peak != 1 (numpy divides a complex by a real through 1 / pmax), threshold stop."""
    g = np.load(os.path.join(GOLDEN, "clean_complex_case.npz"))
    gain, thresh, niter, frac = g[f"{name}_par"]
    window = g[f"{name}_window"] if f"{name}_window" in g.files else None
    outs = cl.hogbom_complex(g[f"{name}_dq"], g[f"{name}_du"], g[f"{name}_psf"], g[f"{name}_psf"].copy(),
                             window, gain, thresh, int(niter), frac)
    for got, key in zip(outs, ("cq", "cu", "rq", "ru")):
        assert isinstance(got, np.ndarray) and np.array_equal(got, g[f"{name}_{key}"]), key


def test_hogbom_complex_larger_and_list(cl):
    import torch
    from oracle import clean as oc
    from ska_sdp_func_python_b200 import deconvolution as dc, minimal_models as mm
    rng = np.random.default_rng(12)
    n = 256
    yy, xx = np.meshgrid(np.arange(n) - n // 2, np.arange(n) - n // 2)
    r = np.hypot(xx, yy)
    psf = np.exp(-(r / 2.0) ** 2) + 0.05 * np.cos(r / 3.1) * np.exp(-r / 30.0) * (r > 0)
    dq, du = 0.01 * rng.normal(size=(n, n)), 0.01 * rng.normal(size=(n, n))
    pos = rng.integers(20, n - 20, (30, 2))
    dq[pos[:, 0], pos[:, 1]] += rng.normal(size=30) * 2
    du[pos[:, 0], pos[:, 1]] += rng.normal(size=30) * 2
    want = oc.hogbom_complex(dq, du, psf, psf, None, 0.2, 0.0, 400, 0.001)
    t = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    got = cl.hogbom_complex(t(dq), t(du), t(psf), t(psf), None, 0.2, 0.0, 400, 0.001)
    for a, b in zip(got, want[:4]):
        assert a.is_cuda and np.array_equal(a.cpu().numpy(), b)
    with pytest.raises(AssertionError):
        cl.hogbom_complex(dq, du, psf, 2 * psf, None, 0.2, 0.0, 10, 0.0)
    # image-list level: I and V by hogbom, Q + iU by hogbom_complex (image/deconvolution.py:398-511)
    pf = mm.PolarisationFrame("stokesIQUV")
    pc = mm.SkyCoord.from_deg(0.0, -30.0)
    dirty = mm.create_image(n, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    psf_im = mm.create_image(n, 1e-3, pc, frequency=[1e8], polarisation_frame=pf)
    dirty["pixels"].data[0] = np.stack([dq + du, dq, du, dq - du])
    psf_im["pixels"].data[0, :3] = psf                        # V PSF all zero: skipped
    comp, resid = dc.deconvolve_cube(dirty, psf_im, algorithm="hogbom-complex", gain=0.2, niter=400,
                                     fractional_threshold=0.001, psf_support=n // 2)
    assert np.array_equal(comp["pixels"].data[0, 1], want[0]) and np.array_equal(comp["pixels"].data[0, 2], want[1])
    assert np.array_equal(resid["pixels"].data[0, 1], want[2]) and np.array_equal(resid["pixels"].data[0, 2], want[3])
    ci, ri, _ = oc.hogbom(dq + du, psf, None, 0.2, 0.0, 400, 0.001)
    assert np.array_equal(comp["pixels"].data[0, 0], ci) and np.array_equal(resid["pixels"].data[0, 0], ri)
    assert not comp["pixels"].data[0, 3].any()


def test_public_helpers_match_oracle(cl):
    import torch
    from oracle import clean as oc
    rng = np.random.default_rng(4)
    img = rng.normal(size=(33, 40))
    stack = oc.scale_stack((3, 33, 40), [0, 4, 9])
    assert np.abs(cl.convolve_scalestack(stack, img) - oc.convolve_stack(stack, img)).max() < 1e-13
    assert np.abs(cl.convolve_convolve_scalestack(stack, img) - oc.convolve_convolve_stack(stack, img)).max() < 1e-13
    assert cl.convolve_scalestack(torch.from_numpy(stack).cuda(), torch.from_numpy(img).cuda()).is_cuda
    for shape, pshape, px, py in (((64, 64), (64, 64), 3, 60), ((48, 80), (33, 41), 47, 0), ((40, 40), (80, 80), 20, 20)):
        assert cl.overlapIndices(np.zeros(shape), np.zeros(pshape), px, py) == oc.overlap(shape, pshape, px, py)
    a = rng.normal(size=(17, 23))
    assert tuple(cl.argmax(a)) == tuple(np.unravel_index(a.argmax(), a.shape)) == cl.argmax(torch.from_numpy(a).cuda())
