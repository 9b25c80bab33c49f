"""raysnail_b200 — B200 (sm_100a) path-tracing backend for the Raysnail renderer.

The package holds only what the hot path needs: `csrc/` (CUDA kernels, the C ABI and the C++
host mirror of the reference's scene interface) and this thin ctypes binding.
"""
from ._abi import RsCamera, RsRenderParams, RsStats, RS_FLAG_PIXEL_CENTRE, RS_FLAG_KERNEL_TIMING, RS_FLAG_MIS, RS_REDO_REFERENCE, RS_REDO_NOISE, RS_REDO_NOISE_FIXED, LIB_PATH  # noqa: F401
from .api import (CameraBuilder, MultiRenderer, RaysnailError, Renderer, Scene, SdlParser, TakePhotoSettings, make_params,  # noqa: F401
                  painter_samples, take_photo, DEFAULT_SEED, philox4x32_10, philox_uniform, pass_seed)

__version__ = "0.1.0"
