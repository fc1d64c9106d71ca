"""ctypes front-end to the CPU oracle (oracle/oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package (tinypathtracer_b200/) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB_PATH = _HERE / "_build" / "liboracle.so"

MATERIAL_DTYPE = np.dtype([("base", "<f4", 3), ("emission", "<f4"), ("eta", "<f4"), ("metallic", "<f4"),
                           ("unused", "<f4", 9)])          # include/material.h:86-120 (60 B)
MTL_INTERVAL_DTYPE = np.dtype([("begin", "<i4"), ("mtl", "<i4")])   # include/mesh.cuh:74-78
LIGHT_DTYPE = np.dtype([("type", "<i4"), ("color", "<f4", 3), ("intensity", "<f4"), ("a", "<f4", 3),
                        ("b", "<f4", 3), ("cos_outer", "<f4"), ("inv_cone_diff", "<f4")])  # delta_light.h (52 B)
REF_NODE_DTYPE = np.dtype([("parent", "<u4"), ("a", "<i4"), ("b", "<i4"), ("mn", "<f4", 3), ("mx", "<f4", 3)])  # bvh.cuh:34-58
HIT_DTYPE = np.dtype([("prim", "<i4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4")])  # path_tracer.cu:19-26
assert MATERIAL_DTYPE.itemsize == 60 and LIGHT_DTYPE.itemsize == 52 and REF_NODE_DTYPE.itemsize == 36


def build(force: bool = False) -> Path:
    src = _HERE / "oracle.cpp"
    if force or not _LIB_PATH.exists() or _LIB_PATH.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(_HERE), "_build/liboracle.so"], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(str(_LIB_PATH))
        L.orc_frsqrt.restype = C.c_float
        L.orc_frsqrt.argtypes = [C.c_float]
        L.orc_cosf.restype = C.c_float
        L.orc_cosf.argtypes = [C.c_float]
        L.orc_morton.restype = C.c_int64
        L.orc_morton.argtypes = [C.c_float] * 3
        L.orc_schlick.restype = C.c_float
        L.orc_schlick.argtypes = [C.c_float, C.c_float]
        L.orc_refract.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        L.orc_hemisphere_cosine.argtypes = [C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_rng_pair.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_mat4_mul_point.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        L.orc_transform.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_build_reference_bvh.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_primary_rays.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_void_p]
        L.orc_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int,
                                     C.c_void_p, C.c_void_p]
        L.orc_retest_triangle.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_float, C.c_void_p]
        L.orc_render.argtypes = ([C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                  C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
                                 + [C.c_int] * 10 + [C.c_uint64, C.c_void_p, C.c_void_p])
        L.orc_render_env.argtypes = ([C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                      C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
                                     + [C.c_int] * 10 + [C.c_uint64, C.c_void_p, C.c_void_p])
        L.orc_resolve_bgra8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orc_sample_env.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


# ---- small known-answer wrappers -------------------------------------------------------------
def frsqrt(x: float) -> float:
    return float(lib().orc_frsqrt(x))


def morton(x, y, z) -> int:
    return int(lib().orc_morton(x, y, z))


def ray_hit_bbox(ray6, mn, mx) -> bool:
    return bool(lib().orc_ray_hit_bbox(_p(f32(ray6)), _p(f32(mn)), _p(f32(mx))))


def ray_hit_triangle(ray6, tri9):
    out = np.zeros(3, np.float32)
    ok = lib().orc_ray_hit_triangle(_p(f32(ray6)), _p(f32(tri9)), _p(out))
    return bool(ok), out


def reflect(d, n):
    out = np.zeros(3, np.float32)
    lib().orc_reflect(_p(f32(d)), _p(f32(n)), _p(out))
    return out


def refract(d, n, ior):
    out = np.zeros(6, np.float32)
    lib().orc_refract(_p(f32(d)), _p(f32(n)), ior, _p(out))
    return out[:3].copy(), float(out[3]), float(out[4]), bool(out[5])


def schlick(cos_i, eta) -> float:
    return float(lib().orc_schlick(cos_i, eta))


def hemisphere_cosine(xi1, xi2, n):
    out = np.zeros(3, np.float32)
    lib().orc_hemisphere_cosine(xi1, xi2, _p(f32(n)), _p(out))
    return out


def light_sample(light: np.ndarray, p):
    out = np.zeros(7, np.float32)
    rec = np.ascontiguousarray(light.reshape(1), dtype=LIGHT_DTYPE)
    lib().orc_light_sample(_p(rec), _p(f32(p)), _p(out))
    return out[:3].copy(), out[3:6].copy(), float(out[6])


def vec2uv(d):
    out = np.zeros(2, np.float32)
    lib().orc_vec2uv(_p(f32(d)), _p(out))
    return out


def sample_env(rgba: np.ndarray, d):
    h, w = rgba.shape[:2]
    out = np.zeros(3, np.float32)
    lib().orc_sample_env(_p(np.ascontiguousarray(rgba, np.uint8)), w, h, _p(f32(d)), _p(out))
    return out


def rng_pair(seed, pixel, sample, pair_idx):
    out = np.zeros(2, np.float32)
    lib().orc_rng_pair(seed, pixel, sample, pair_idx, _p(out))
    return out


def trs_to_mat4(t, q, s):
    out = np.zeros(16, np.float32)
    lib().orc_trs_to_mat4(_p(f32(t)), _p(f32(q)), _p(f32(s)), _p(out))
    return out


def normal_matrix(l2w):
    out = np.zeros(16, np.float32)
    lib().orc_normal_matrix(_p(f32(l2w)), _p(out))
    return out


def mat4_mul_point(m, p, w):
    out = np.zeros(3, np.float32)
    lib().orc_mat4_mul_point(_p(f32(m)), _p(f32(p)), float(w), _p(out))
    return out


def cosf(x) -> float:
    return float(lib().orc_cosf(float(np.float32(x))))


# ---- scene-level -----------------------------------------------------------------------------
def camera19(scene: dict) -> np.ndarray:
    return np.concatenate([[scene["vfov"], scene["aspect"], scene["near"]], scene["c2w"]]).astype(np.float32)


def transform(scene: dict):
    """src/path_tracer.cu:239-263 — returns (world vertices, world normals)."""
    V = scene["vertices"].shape[0]
    wv = np.zeros((V, 3), np.float32)
    wn = np.zeros((V, 3), np.float32)
    F = scene["indices"].size // 3
    lib().orc_transform(_p(scene["indices"]), F, _p(scene["vertices"]), _p(scene["normals"]), V, _p(scene["objects"]),
                        scene["objects"].shape[0], _p(scene["vert_trans"]), _p(scene["normal_trans"]), _p(wv), _p(wn))
    return wv, wn


def build_reference_bvh(wverts, indices):
    F = indices.size // 3
    nodes = np.zeros(2 * F - 1, REF_NODE_DTYPE)
    keys = np.zeros(F, np.int64)
    rc = lib().orc_build_reference_bvh(_p(wverts), _p(indices), F, _p(nodes), _p(keys))
    return nodes, keys, rc


def primary_rays(scene: dict, W: int, H: int, jitter: bool = False, seed: int = 0) -> np.ndarray:
    rays = np.zeros((H * W, 6), np.float32)
    lib().orc_primary_rays(_p(camera19(scene)), W, H, int(jitter), seed, _p(rays))
    return rays


def trace_rays(wverts, indices, rays, nodes=None, mode: int = 0, counters: bool = False):
    F = indices.size // 3
    rays = f32(rays)
    hits = np.zeros(rays.shape[0], HIT_DTYPE)
    cnt = np.zeros(3, np.uint64)
    bad = lib().orc_trace_rays(_p(wverts), _p(indices), F, _p(nodes), _p(rays), rays.shape[0], mode, _p(hits), _p(cnt))
    if mode == 1 and bad:
        raise RuntimeError(f"reference traversal stack overflow on {bad} rays")
    return (hits, cnt) if counters else hits


def retest_triangle(wverts, indices, fid, ray6, eps=1e-5):
    t = C.c_float(0)
    ok = lib().orc_retest_triangle(_p(wverts), _p(indices), int(fid), _p(f32(ray6)), eps, C.byref(t))
    return bool(ok), float(t.value)


def render(scene: dict, W: int, H: int, spp: int, max_depth: int = 8, seed: int = 1, env: np.ndarray | None = None,
           region=None, sample_offset: int = 0, sample_stride: int = 1, prepared=None, counters: bool = False,
           rough_metals: bool = False):
    """Restates src/path_tracer.cu:491-554 doTrace up to (not including) copyToFB.

    Returns the float32 radiance SUM buffer [H, W, 3] (row 0 = bottom of the image, as in the
    reference's `color` buffer).  `region` = (x0, y0, x1, y1) bounds the work.
    """
    if prepared is None:
        wv, wn = transform(scene)
        nodes, _, rc = build_reference_bvh(wv, scene["indices"])
        if rc:
            raise RuntimeError("reference hierarchy is malformed for this input (SURVEY fact 9)")
    else:
        wv, wn, nodes = prepared
    out = np.zeros((H, W, 3), np.float32)
    x0, y0, x1, y1 = region if region is not None else (0, 0, W, H)
    ew = eh = 0
    envp = None
    envfp = None
    if env is not None and np.asarray(env).dtype.kind == "f":      # HDR: linear float RGB [h, w, 3]
        env = np.ascontiguousarray(env, np.float32)
        assert env.ndim == 3 and env.shape[2] == 3
        eh, ew = env.shape[:2]
        envfp = _p(env)
    elif env is not None:
        env = np.ascontiguousarray(env, np.uint8)
        eh, ew = env.shape[:2]
        envp = _p(env)
    cnt = np.zeros(3, np.uint64)
    F = scene["indices"].size // 3
    lib().orc_set_extensions(1 if rough_metals else 0)
    try:
        _render_env(scene, F, wv, wn, nodes, envp, envfp, ew, eh, W, H, x0, y0, x1, y1, spp, sample_offset, sample_stride, max_depth, seed,
                    out, cnt if counters else None)
    finally:
        lib().orc_set_extensions(0)
    return (out, cnt) if counters else out


def _render_env(scene, F, wv, wn, nodes, envp, envfp, ew, eh, W, H, x0, y0, x1, y1, spp, sample_offset, sample_stride, max_depth, seed, out, cnt):
    lib().orc_render_env(_p(scene["indices"]), F, _p(wv), _p(wn), _p(scene["materials"]), scene["materials"].shape[0],
                     _p(scene["objects"]), scene["objects"].shape[0], _p(scene["lights"]), scene["lights"].shape[0],
                     _p(nodes), envp, envfp, ew, eh, _p(camera19(scene)), W, H, x0, y0, x1, y1, spp, sample_offset, sample_stride,
                     max_depth, seed, _p(out), _p(cnt))


def resolve_bgra8(radiance: np.ndarray, spp: int, fb: np.ndarray | None = None) -> np.ndarray:
    H, W = radiance.shape[:2]
    if fb is None:
        fb = np.zeros((H, W, 4), np.uint8)
    lib().orc_resolve_bgra8(_p(f32(radiance)), W, H, spp, _p(fb))
    return fb


def num_threads() -> int:
    return int(lib().orc_num_threads())
