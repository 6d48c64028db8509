"""B200-native visibility gridder / degridder: a drop-in for the hot path of
``ska_sdp_func_python.imaging`` (``invert_ng`` / ``predict_ng`` and the ``ducc0.wgridder``
calls underneath them).  The arithmetic lives in hand-written sm_100a CUDA kernels behind
the C ABI in include/b200_wgridder.h; importing this package does not need a GPU, calling
any compute entry point does (there is no CPU fallback)."""
from . import _lib  # noqa: F401

__version__ = "0.1.0"
