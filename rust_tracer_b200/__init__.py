"""rust_tracer_b200 -- B200 (sm_100a) render path for erichgess/rust_tracer.

Only the hot path lives here: csrc/ (CUDA kernels + the C ABI of include/rt_b200.h, built
into librt_b200.so), the C++ host mirror of the reference's scene API (csrc/host/), and
this thin ctypes binding.  There is no CPU implementation in this package.
"""
from . import abi  # noqa: F401
from .api import Forest, MultiScene, Scene, SceneDesc, camera, device_count, measure_fp32_peak, render_multi, tile_rows, trim, tuned  # noqa: F401
from ._lib import RtError, lib  # noqa: F401
