"""ctypes front-end to oracle/_ref/libref_host.so: the reference's OWN host-callable functions, compiled from
/root/reference by oracle/ref_build.py.  TEST INFRASTRUCTURE ONLY — used to pin oracle/oracle.cpp and to generate
tests/golden/ref_host_vectors.npz."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from .pyoracle import HIT_DTYPE, LIGHT_DTYPE, MATERIAL_DTYPE, REF_NODE_DTYPE

LIB_PATH = Path(__file__).resolve().parent / "_ref" / "libref_host.so"
_lib = None


def available() -> bool:
    return LIB_PATH.exists()


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(str(LIB_PATH))
        L.ref_frsqrt.restype = C.c_float
        L.ref_frsqrt.argtypes = [C.c_float]
        L.ref_schlick.restype = C.c_float
        L.ref_schlick.argtypes = [C.c_float, C.c_float]
        L.ref_morton.restype = C.c_longlong
        L.ref_morton.argtypes = [C.c_float] * 3
        L.ref_refract.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        L.ref_to_uchar.argtypes = [C.c_void_p, C.c_float, C.c_void_p]
        L.ref_traverse.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_scene_load.restype = C.c_void_p
        L.ref_scene_load.argtypes = [C.c_char_p]
        for n in ("ref_scene_free", "ref_scene_counts", "ref_scene_camera"):
            getattr(L, n).argtypes = [C.c_void_p] + ([C.c_void_p] if n != "ref_scene_free" else [])
        L.ref_mesh_sizes.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.ref_mesh_data.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_char_p, C.c_int]
        L.ref_material.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_char_p, C.c_int]
        L.ref_light.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_char_p, C.c_int]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def f32(a):
    return np.ascontiguousarray(a, np.float32)


def frsqrt(x):
    return float(lib().ref_frsqrt(x))


def morton(x, y, z):
    return int(lib().ref_morton(x, y, z))


def normalize(v):
    out = np.zeros(3, np.float32)
    lib().ref_normalize(_p(f32(v)), _p(out))
    return out


def ray_hit_bbox(ray6, mn, mx):
    return bool(lib().ref_ray_hit_bbox(_p(f32(ray6)), _p(f32(mn)), _p(f32(mx))))


def ray_hit_triangle(ray6, tri9):
    out = np.zeros(3, np.float32)
    ok = lib().ref_ray_hit_triangle(_p(f32(ray6)), _p(f32(tri9)), _p(out))
    return bool(ok), out


def traverse(nodes, wverts, indices, rays):
    rays = f32(rays)
    hits = np.zeros(rays.shape[0], HIT_DTYPE)
    nodes = np.ascontiguousarray(nodes)
    lib().ref_traverse(_p(nodes), _p(f32(wverts)), _p(np.ascontiguousarray(indices, np.uint32)), indices.size // 3, _p(rays),
                       rays.shape[0], _p(hits))
    return hits


def reflect(d, n):
    out = np.zeros(3, np.float32)
    lib().ref_reflect(_p(f32(d)), _p(f32(n)), _p(out))
    return out


def refract(d, n, ior):
    out = np.zeros(6, np.float32)
    lib().ref_refract(_p(f32(d)), _p(f32(n)), ior, _p(out))
    return out


def schlick(c, eta):
    return float(lib().ref_schlick(c, eta))


def light_sample(light, p):
    out = np.zeros(7, np.float32)
    rec = np.ascontiguousarray(np.asarray(light).reshape(1), LIGHT_DTYPE)
    lib().ref_light_sample(_p(rec), _p(f32(p)), _p(out))
    return out


def vec2uv(d):
    out = np.zeros(2, np.float32)
    lib().ref_vec2uv(_p(f32(d)), _p(out))
    return out


def to_uchar(rgb, spp):
    out = np.zeros(3, np.uint8)
    lib().ref_to_uchar(_p(f32(rgb)), float(spp), _p(out))
    return out


def trs_to_mat4(t, q, s):
    out = np.zeros(16, np.float32)
    lib().ref_trs_to_mat4(_p(f32(t)), _p(f32(q)), _p(f32(s)), _p(out))
    return out


def normal_matrix(m):
    out = np.zeros(16, np.float32)
    lib().ref_normal_matrix(_p(f32(m)), _p(out))
    return out


def mat4_mul_vec4(m, v):
    out = np.zeros(4, np.float32)
    lib().ref_mat4_mul_vec4(_p(f32(m)), _p(f32(v)), _p(out))
    return out


def load_scene(path) -> dict:
    """What Scene::readFromGLTF parsed, flattened as Scene::copySceneToDevice would (src/mesh.cu:309-397)."""
    L = lib()
    h = L.ref_scene_load(str(path).encode())
    if not h:
        raise RuntimeError(f"reference loader failed on {path}")
    try:
        cnt = np.zeros(3, np.int32)
        L.ref_scene_counts(h, _p(cnt))
        cam = np.zeros(19, np.float32)
        L.ref_scene_camera(h, _p(cam))
        names, mats = [], np.zeros(max(int(cnt[1]), 0), MATERIAL_DTYPE)
        buf = C.create_string_buffer(256)
        for i in range(int(cnt[1])):
            rec = np.zeros(1, MATERIAL_DTYPE)
            L.ref_material(h, i, _p(rec), buf, 256)
            mats[i] = rec[0]
            names.append(buf.value.decode())
        lights = np.zeros(int(cnt[2]), LIGHT_DTYPE)
        lnames = []
        for i in range(int(cnt[2])):
            rec = np.zeros(1, LIGHT_DTYPE)
            L.ref_light(h, i, _p(rec), buf, 256)
            lights[i] = rec[0]
            lnames.append(buf.value.decode())
        meshes = []
        for i in range(int(cnt[0])):
            sz = np.zeros(2, np.int32)
            L.ref_mesh_sizes(h, i, _p(sz))
            idx = np.zeros(int(sz[0]), np.uint32)
            v = np.zeros((int(sz[1]), 3), np.float32)
            n = np.zeros((int(sz[1]), 3), np.float32)
            m = np.zeros(16, np.float32)
            L.ref_mesh_data(h, i, _p(idx), _p(v), _p(n), _p(m), buf, 256)
            meshes.append(dict(indices=idx, vertices=v, normals=n, l2w=m, material=buf.value.decode()))
        return dict(camera=cam, material_names=names, materials=mats, light_names=lnames, lights=lights, meshes=meshes)
    finally:
        L.ref_scene_free(h)
