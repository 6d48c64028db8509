"""ctypes binding of the C ABI in include/b200_wgridder.h (libb200wgridder.so).

The shared library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a).  There
is deliberately no fallback: if the library is missing, or no CUDA device is present when a
compute entry point is called, an exception is raised.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2G_LIB: development override (kernel variants built side by side)
LIB_PATH = os.environ.get("B2G_LIB") or os.path.join(_HERE, "libb200wgridder.so")

F32, F64 = 0, 1
HOST, DEVICE = 0, 1
ALGO_AUTO, ALGO_SIMPLE, ALGO_WARPRUN = -1, 0, 1

EXPORTS = (
    "b2g_last_error", "b2g_version", "b2g_plan_create", "b2g_plan_destroy",
    "b2g_plan_invert", "b2g_plan_predict", "b2g_plan_info", "b2g_plan_timings",
    "b2g_ms2dirty", "b2g_dirty2ms", "b2g_sumwt", "b2g_kernel_select", "b2g_kernel_eval",
    "b2g_kernel_corr", "b2g_vis_prepare", "b2g_vis_finish", "b2g_set_host_chunk_samples", "b2g_set_max_slab_planes", "b2g_set_tc_mode", "b2g_plan_create_sel",
    "b2g_weight_grid", "b2g_weight_apply", "b2g_weight_taper", "b2g_dft_predict",
    "b2g_plan_residual", "b2g_hogbom", "b2g_image_peak", "b2g_dft_invert", "b2g_msclean", "b2g_msmfsclean", "b2g_hogbom_complex",
    "b2g_sumwt_static",
)
FLAG_NONE, FLAG_U8, FLAG_I32, FLAG_I64, FLAG_F64 = 0, 1, 2, 3, 4

_lib = None


class WgridderError(RuntimeError):
    """Raised when a C-ABI call returns a non-zero status."""


def lib():
    """Load (once) and return the ctypes handle of libb200wgridder.so."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` (needs nvcc); there is no CPU fallback")
    L = ctypes.CDLL(LIB_PATH)
    i64, dbl, vp, ci = ctypes.c_int64, ctypes.c_double, ctypes.c_void_p, ctypes.c_int
    pd = ctypes.POINTER(ctypes.c_double)
    L.b2g_last_error.restype = ctypes.c_char_p
    L.b2g_last_error.argtypes = []
    L.b2g_version.restype = ctypes.c_char_p
    L.b2g_version.argtypes = []
    L.b2g_plan_create.restype = ci
    L.b2g_plan_create.argtypes = [ctypes.POINTER(vp), i64, i64, vp, vp, ci, i64, i64, dbl, dbl,
                                  i64, i64, dbl, ci, ci, ci, ci, vp]
    L.b2g_plan_create_sel.restype = ci
    L.b2g_plan_create_sel.argtypes = [ctypes.POINTER(vp), i64, i64, vp, vp, ci, vp, i64, i64, i64,
                                      dbl, dbl, i64, i64, dbl, ci, ci, ci, ci, vp]
    L.b2g_plan_destroy.restype = None
    L.b2g_plan_destroy.argtypes = [vp]
    L.b2g_plan_invert.restype = ci
    L.b2g_plan_invert.argtypes = [vp, vp, vp, ci, ci, vp, ci, ci]
    L.b2g_plan_predict.restype = ci
    L.b2g_plan_predict.argtypes = [vp, vp, ci, ci, vp, vp, ci, ci]
    L.b2g_plan_info.restype = ci
    L.b2g_plan_info.argtypes = [vp, ctypes.POINTER(i64), ci, pd, ci]
    L.b2g_plan_timings.restype = ci
    L.b2g_plan_timings.argtypes = [vp, pd, ci]
    L.b2g_ms2dirty.restype = ci
    L.b2g_ms2dirty.argtypes = [i64, i64, vp, vp, vp, vp, ci, ci, i64, i64, dbl, dbl, i64, i64,
                               dbl, ci, ci, vp, ci, vp]
    L.b2g_dirty2ms.restype = ci
    L.b2g_dirty2ms.argtypes = [i64, i64, vp, vp, vp, ci, vp, ci, ci, i64, i64, dbl, dbl, i64,
                               i64, dbl, ci, ci, vp, vp]
    L.b2g_sumwt.restype = ci
    L.b2g_sumwt.argtypes = [vp, ci, ci, i64, pd, vp]
    L.b2g_sumwt_static.restype = ci
    L.b2g_sumwt_static.argtypes = [vp, ci, i64, vp, vp]
    L.b2g_set_host_chunk_samples.restype = ci
    L.b2g_set_host_chunk_samples.argtypes = [i64]
    L.b2g_set_max_slab_planes.restype = ci
    L.b2g_set_max_slab_planes.argtypes = [ci]
    L.b2g_set_tc_mode.restype = ci
    L.b2g_set_tc_mode.argtypes = [ci]
    L.b2g_kernel_select.restype = ci
    L.b2g_kernel_select.argtypes = [dbl, ci, ci, i64, i64, dbl, pd]
    L.b2g_kernel_eval.restype = ci
    L.b2g_kernel_eval.argtypes = [ci, dbl, dbl, pd, ci, pd]
    L.b2g_kernel_corr.restype = ci
    L.b2g_kernel_corr.argtypes = [ci, dbl, dbl, pd, ci, pd]
    L.b2g_vis_prepare.restype = ci
    L.b2g_vis_prepare.argtypes = [i64, i64, ci, ci, vp, ci, vp, ci, vp, ci, pd, pd, vp, vp, i64,
                                  i64, ci, vp, vp, ci, vp]
    L.b2g_vis_finish.restype = ci
    L.b2g_vis_finish.argtypes = [i64, i64, ci, ci, vp, ci, pd, pd, vp, vp, i64, i64, vp, ci, vp]
    L.b2g_weight_grid.restype = ci
    L.b2g_weight_grid.argtypes = [i64, i64, ci, vp, vp, vp, vp, ci, vp, ci, i64, i64, dbl, dbl,
                                  dbl, dbl, vp, pd, pd, vp]
    L.b2g_weight_apply.restype = ci
    L.b2g_weight_apply.argtypes = [i64, i64, ci, vp, vp, vp, vp, ci, vp, ci, i64, i64, dbl, dbl,
                                   dbl, dbl, vp, ci, dbl, vp, vp]
    L.b2g_weight_taper.restype = ci
    L.b2g_weight_taper.argtypes = [i64, i64, ci, vp, vp, vp, ci, ci, dbl, vp, vp]
    L.b2g_dft_predict.restype = ci
    L.b2g_dft_predict.argtypes = [i64, i64, ci, i64, vp, vp, vp, vp, i64, vp, ci, ci, vp]
    L.b2g_plan_residual.restype = ci
    L.b2g_plan_residual.argtypes = [vp, vp, ci, ci, vp, vp, ci, ci, vp, vp]
    L.b2g_hogbom.restype = ci
    L.b2g_hogbom.argtypes = [vp, vp, vp, i64, i64, i64, i64, dbl, dbl, dbl, dbl, i64, vp, vp,
                             ctypes.POINTER(i64), vp]
    L.b2g_image_peak.restype = ci
    L.b2g_image_peak.argtypes = [vp, vp, i64, i64, ci, pd, ctypes.POINTER(i64), vp]
    L.b2g_dft_invert.restype = ci
    L.b2g_dft_invert.argtypes = [i64, i64, ci, i64, vp, vp, vp, vp, vp, vp, ci, vp, vp, vp]
    L.b2g_msclean.restype = ci
    L.b2g_msclean.argtypes = [vp, vp, vp, vp, vp, pd, i64, i64, i64, i64, i64, dbl, dbl, i64, vp,
                              ctypes.POINTER(i64), vp]
    L.b2g_msmfsclean.restype = ci
    L.b2g_msmfsclean.argtypes = [vp, vp, vp, vp, vp, pd, pd, ci, i64, i64, i64, i64, i64, i64, dbl, dbl, i64, vp,
                                 ctypes.POINTER(i64), vp]
    L.b2g_hogbom_complex.restype = ci
    L.b2g_hogbom_complex.argtypes = [vp, vp, vp, vp, i64, i64, i64, i64, dbl, dbl, dbl, dbl, i64, vp, vp, vp, vp,
                                     ctypes.POINTER(i64), vp]
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise WgridderError(lib().b2g_last_error().decode("utf-8", "replace"))
