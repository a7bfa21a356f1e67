"""plutracer_b200 — B200-native render path for andrew-pa/plutracer behind a C ABI (libplugpu.so).

Host-side mirror of the reference's plugin surface for the render path; the compute is hand-written
sm_100a CUDA (csrc/).  No CPU fallback, no PyTorch fallback: importing the FFI without the built
library raises.
"""
from . import _ffi  # noqa: F401
from ._build import build_lib  # noqa: F401
from .renderer import DEFAULT_SEED, GpuContext, postprocesser, renderer, write_bmp  # noqa: F401
from .scene import Scene, stratified_sampler, texture2d  # noqa: F401
from .shard import reduce_framebuffer, shard_plan, sharded_renderer  # noqa: F401
