"""tinypathtracer_b200 — B200-native (sm_100a) path-tracing core behind the C ABI of include/tpt.h.

The package is a thin ctypes binding plus mirrors of the reference's classes; all compute lives in
tinypathtracer_b200/lib/libtpt_b200.so (hand-written CUDA, built by build.py).  Importing the package
without the built library raises ImportError: there is no fallback path.
"""
from .api import (Accum, BVH, Camera, Context, DeviceScene, EnvLight, PathTracer, HIT_DTYPE, RAY_DTYPE, REF_NODE_DTYPE,
                  MATERIAL_DTYPE, MTL_INTERVAL_DTYPE, LIGHT_DTYPE, ACCUM_FRAC_BITS, render_config, comm_init_all, accum_allreduce, FrameExchange)
from ._capi import TptError, LIB_PATH

__all__ = ["Accum", "BVH", "Camera", "Context", "DeviceScene", "EnvLight", "PathTracer", "TptError", "LIB_PATH", "HIT_DTYPE",
           "RAY_DTYPE", "REF_NODE_DTYPE", "MATERIAL_DTYPE", "MTL_INTERVAL_DTYPE", "LIGHT_DTYPE", "ACCUM_FRAC_BITS", "render_config", "comm_init_all", "accum_allreduce", "FrameExchange"]
