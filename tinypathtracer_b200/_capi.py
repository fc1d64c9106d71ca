"""ctypes declarations for libtpt_b200.so (include/tpt.h).  No compute happens in Python.

Importing this module loads the shared library; if it has not been built the import fails loudly —
there is no Python, PyTorch or CPU stand-in for any entry point.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
# TPT_B200_LIB selects a tuning variant of the same library (tinypathtracer_b200/build.py -D... -o...)
LIB_PATH = Path(os.environ.get("TPT_B200_LIB") or (_PKG / "lib" / "libtpt_b200.so"))

if not LIB_PATH.exists():
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(nvcc, sm_100a). tinypathtracer_b200 has no fallback implementation.")

lib = C.CDLL(str(LIB_PATH))

TPT_OK = 0
TPT_RENDER_COUNTERS = 1
TPT_RENDER_STAGE_TIMES = 2
TPT_RENDER_ENV_NEE = 4
TPT_RENDER_NO_PIXEL_CANDIDATES = 8
TPT_RENDER_REFIT = 16
TPT_RENDER_ROUGH_METALS = 32
TPT_RENDER_NO_SPECULAR_CONTINUATION = 64
TPT_TRACE_STATIC_MAX_RAYS = 65536


class Ray(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("dir", C.c_float * 3)]


class Hit(C.Structure):
    _fields_ = [("prim", C.c_int32), ("t", C.c_float), ("u", C.c_float), ("v", C.c_float)]


class SceneDesc(C.Structure):
    _fields_ = [("indices", C.c_void_p), ("n_indices", C.c_size_t),
                ("vertices", C.c_void_p), ("normals", C.c_void_p), ("n_vertices", C.c_size_t),
                ("materials", C.c_void_p), ("n_materials", C.c_size_t),
                ("objects", C.c_void_p), ("n_objects", C.c_size_t),
                ("vert_trans", C.c_void_p), ("normal_trans", C.c_void_p),
                ("lights", C.c_void_p), ("n_lights", C.c_size_t)]


class Camera(C.Structure):
    _fields_ = [("vfov", C.c_float), ("aspect", C.c_float), ("near_plane", C.c_float), ("c2w", C.c_float * 16)]


class RenderConfig(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("max_depth", C.c_int32),
                ("seed", C.c_uint64), ("sample_offset", C.c_int32), ("sample_stride", C.c_int32),
                ("flags", C.c_uint32), ("paths_in_flight", C.c_int32)]


class BvhStats(C.Structure):
    _fields_ = [("build_ms", C.c_float), ("n_faces", C.c_uint32), ("n_nodes", C.c_uint32), ("max_depth", C.c_uint32),
                ("sah_cost", C.c_float), ("bounds_min", C.c_float * 3), ("bounds_max", C.c_float * 3)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays_primary", "rays_extension", "rays_shadow", "rays_direct", "node_visits",
                                          "tri_tests", "samples", "dropped_nonfinite", "kernel_launches", "rays_culled", "rays_candidates", "queue_overflow", "rays_queued")]


class StageTimes(C.Structure):
    _fields_ = [("ms", C.c_double * 4), ("launches", C.c_uint64 * 4)]


# every symbol include/tpt.h declares, with its signature
TPT_EXCHANGE_HANDLE_BYTES = 168

SIGNATURES = {
    "tpt_last_error": (C.c_char_p, []),
    "tpt_abi_version": (C.c_int, []),
    "tpt_context_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "tpt_context_destroy": (C.c_int, [C.c_void_p]),
    "tpt_context_synchronize": (C.c_int, [C.c_void_p]),
    "tpt_context_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tpt_get_counters": (C.c_int, [C.c_void_p, C.POINTER(Counters)]),
    "tpt_reset_counters": (C.c_int, [C.c_void_p]),
    "tpt_get_stage_times": (C.c_int, [C.c_void_p, C.POINTER(StageTimes)]),
    "tpt_scene_create": (C.c_int, [C.c_void_p, C.POINTER(SceneDesc), C.POINTER(C.c_void_p)]),
    "tpt_scene_destroy": (C.c_int, [C.c_void_p]),
    "tpt_scene_update": (C.c_int, [C.c_void_p, C.POINTER(SceneDesc)]),
    "tpt_scene_set_transforms": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "tpt_bvh_build": (C.c_int, [C.c_void_p, C.POINTER(BvhStats)]),
    "tpt_bvh_refit": (C.c_int, [C.c_void_p, C.POINTER(BvhStats)]),
    "tpt_bvh_build_from_world": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p), C.POINTER(BvhStats)]),
    "tpt_do_trace_device": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.POINTER(RenderConfig), C.c_void_p, C.c_void_p]),
    "tpt_pointer_is_device": (C.c_int, [C.c_void_p]),
    "tpt_bvh_export_reference_nodes": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "tpt_bvh_export_order": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "tpt_scene_export_world": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "tpt_comm_init_all": (C.c_int, [C.c_int, C.c_void_p]),
    "tpt_accum_allreduce": (C.c_int, [C.c_void_p, C.c_int]),
    "tpt_exchange_create": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_ubyte, C.POINTER(C.c_void_p), C.c_void_p]),
    "tpt_exchange_connect": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tpt_exchange_resolve": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "tpt_exchange_framebuffer": (C.c_void_p, [C.c_void_p]),
    "tpt_exchange_read_framebuffer": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tpt_exchange_status": (C.c_int, [C.c_void_p]),
    "tpt_exchange_destroy": (C.c_int, [C.c_void_p]),
    "tpt_env_set_rgba8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "tpt_env_set_rgb32f": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "tpt_env_clear": (C.c_int, [C.c_void_p]),
    "tpt_env_sample": (C.c_int, [C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p]),
    "tpt_env_sampler_tables": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "tpt_trace_rays": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int]),
    "tpt_trace_rays_counted": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int]),
    "tpt_trace_rays_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int]),
    "tpt_generate_primary_rays": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_int]),
    "tpt_accum_create": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "tpt_accum_wrap": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "tpt_accum_destroy": (C.c_int, [C.c_void_p]),
    "tpt_accum_clear": (C.c_int, [C.c_void_p]),
    "tpt_accum_add": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tpt_accum_device_ptr": (C.c_void_p, [C.c_void_p]),
    "tpt_accum_export_f32": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tpt_accum_export_i64": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tpt_accum_import_i64": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tpt_render": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.POINTER(RenderConfig), C.c_void_p]),
    "tpt_resolve_bgra8": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int]),
    "tpt_do_trace": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.POINTER(RenderConfig), C.c_void_p, C.c_void_p]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)   # AttributeError here = the library does not export what tpt.h declares
    _fn.restype = _res
    _fn.argtypes = _args


class TptError(RuntimeError):
    """Non-zero status from the C ABI (the reference throws std::runtime_error from CUDA_CHECK)."""


def check(rc: int) -> None:
    if rc != TPT_OK:
        raise TptError(f"tpt status {rc}: {lib.tpt_last_error().decode(errors='replace')}")
