"""Object wrappers over the C ABI, named after the reference's classes.

  reference (C++)                                  here
  ------------------------------------------------ ----------------------------------------------
  DeviceScene / Scene::copySceneToDevice            DeviceScene(ctx, arrays)      (include/mesh.cuh:80-96)
  BVH(nFaces).construct(wVerts, indices)            BVH(scene).construct()        (include/bvh.cuh:60-68)
  Camera.getVFov/getAspRatio/localToWorld           Camera(vfov, aspect, near, c2w)
  EnvLight(file)                                    EnvLight(ctx, rgba8)          (include/env_light.cuh:9-19)
  PathTracer(env).doTrace(scene, camera, fb, spp)   PathTracer(ctx, w, h).do_trace(...)  (src/path_tracer.cu:491-554)

Everything here is argument marshalling; all compute is in libtpt_b200.so.
"""
from __future__ import annotations

import ctypes as C
import weakref

import numpy as np

from . import _capi as capi
from ._capi import check, lib

MATERIAL_DTYPE = np.dtype([("base", "<f4", 3), ("emission", "<f4"), ("eta", "<f4"), ("metallic", "<f4"), ("unused", "<f4", 9)])
MTL_INTERVAL_DTYPE = np.dtype([("begin", "<i4"), ("mtl", "<i4")])
LIGHT_DTYPE = np.dtype([("type", "<i4"), ("color", "<f4", 3), ("intensity", "<f4"), ("a", "<f4", 3), ("b", "<f4", 3),
                        ("cos_outer", "<f4"), ("inv_cone_diff", "<f4")])
REF_NODE_DTYPE = np.dtype([("parent", "<u4"), ("a", "<i4"), ("b", "<i4"), ("mn", "<f4", 3), ("mx", "<f4", 3)])
HIT_DTYPE = np.dtype([("prim", "<i4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4")])
RAY_DTYPE = np.dtype([("origin", "<f4", 3), ("dir", "<f4", 3)])
ACCUM_FRAC_BITS = 36


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Context:
    def __init__(self, device: int = 0, _handle=None):
        h = C.c_void_p()
        if _handle is None:
            check(lib.tpt_context_create(device, C.byref(h)))
        else:
            h = _handle
        self._h = h
        self.device = device
        self._children = weakref.WeakSet()   # scenes / accumulators must be destroyed before their context

    def close(self):
        if self._h:
            kids = list(self._children)
            for child in sorted(kids, key=lambda c: not isinstance(c, FrameExchange)):   # exchanges first: they use accumulators
                child.close()
            lib.tpt_context_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def synchronize(self):
        check(lib.tpt_context_synchronize(self._h))

    def set_stream(self, cuda_stream: int):
        check(lib.tpt_context_set_stream(self._h, C.c_void_p(cuda_stream)))

    def counters(self) -> dict:
        c = capi.Counters()
        check(lib.tpt_get_counters(self._h, C.byref(c)))
        return {n: int(getattr(c, n)) for n, _ in capi.Counters._fields_}

    def stage_times(self) -> dict:
        st = capi.StageTimes()
        check(lib.tpt_get_stage_times(self._h, C.byref(st)))
        names = ("primary", "wave")
        return {n: {"ms": float(st.ms[i]), "launches": int(st.launches[i])} for i, n in enumerate(names)}

    def reset_counters(self):
        check(lib.tpt_reset_counters(self._h))

    def primary_rays(self, cam: "Camera", width: int, height: int, jitter: bool = False, seed: int = 0) -> np.ndarray:
        rays = np.zeros(width * height, RAY_DTYPE)
        check(lib.tpt_generate_primary_rays(self._h, C.byref(cam.c), width, height, int(jitter), seed, _ptr(rays), 0))
        return rays


def primary_rays_device(ctx: Context, cam: "Camera", width: int, height: int, rays_ptr: int, jitter: bool = False, seed: int = 0):
    check(lib.tpt_generate_primary_rays(ctx._h, C.byref(cam.c), width, height, int(jitter), seed, C.c_void_p(rays_ptr), 1))


class Camera:
    """include/camera.h: vfov in radians, aspect from the glTF camera, c2w = Transform::localToWorld()."""

    def __init__(self, vfov: float, aspect: float, near: float, c2w):
        self.c = capi.Camera()
        self.c.vfov = float(vfov)
        self.c.aspect = float(aspect)
        self.c.near_plane = float(near)
        m = np.ascontiguousarray(c2w, np.float32).reshape(16)
        for i in range(16):
            self.c.c2w[i] = float(m[i])

    @classmethod
    def from_scene(cls, scene: dict) -> "Camera":
        return cls(scene["vfov"], scene["aspect"], scene["near"], scene["c2w"])

    def getVFov(self):
        return self.c.vfov

    def getAspRatio(self):
        return self.c.aspect

    def getNearPlane(self):
        return self.c.near_plane


def comm_init_all(ngpu: int) -> list:
    """tpt_comm_init_all: one Context per GPU 0..ngpu-1 of THIS process, peer access enabled between all of them."""
    arr = (C.c_void_p * ngpu)()
    check(lib.tpt_comm_init_all(ngpu, arr))
    return [Context(d, _handle=C.c_void_p(arr[d])) for d in range(ngpu)]


def accum_allreduce(accums: list):
    """tpt_accum_allreduce: every accumulator (one per context) ends up holding the exact sum of all of them."""
    arr = (C.c_void_p * len(accums))(*[a._h for a in accums])
    check(lib.tpt_accum_allreduce(arr, len(accums)))


class FrameExchange:
    """tpt_exchange_*: the end of a frame with one process PER GPU.  Every rank sums its 1/world of the pixels straight out of
    all ranks' accumulators (peer memory, CUDA IPC), resolves them like tpt_resolve_bgra8 and stores them into the root's
    framebuffer — one fused pass instead of all-reduce + resolve.  `handle` is what the ranks exchange (any transport);
    `connect` takes all of them, in rank order."""

    def __init__(self, ctx: Context, accum: "Accum", rank: int, world: int, alpha: int = 0):
        h = C.c_void_p()
        buf = (C.c_ubyte * capi.TPT_EXCHANGE_HANDLE_BYTES)()
        check(lib.tpt_exchange_create(ctx._h, accum._h, rank, world, alpha, C.byref(h), buf))
        self._h, self.ctx, self.accum, self.rank, self.world = h, ctx, accum, rank, world
        self.handle = bytes(buf)
        ctx._children.add(self)

    def connect(self, handles):
        """handles: the `handle` of rank 0, 1, ... world-1 (a list of bytes, or their concatenation)."""
        raw = handles if isinstance(handles, (bytes, bytearray)) else b"".join(handles)
        if len(raw) != self.world * capi.TPT_EXCHANGE_HANDLE_BYTES:
            raise capi.TptError(f"FrameExchange.connect: {len(raw)} bytes for {self.world} ranks")
        check(lib.tpt_exchange_connect(self._h, C.c_char_p(bytes(raw))))

    def connect_over(self, dist, group=None):
        """Gather the handles with torch.distributed (all_gather_object on `group`: a gloo group keeps it off the GPU) and connect."""
        out = [None] * self.world
        dist.all_gather_object(out, self.handle, group=group)
        self.connect(out)

    def resolve(self, spp_total: int, root: int = 0):
        check(lib.tpt_exchange_resolve(self._h, spp_total, root))

    def framebuffer_ptr(self) -> int:
        return int(lib.tpt_exchange_framebuffer(self._h) or 0)

    def framebuffer(self) -> np.ndarray:
        """Synchronous copy of this rank's framebuffer: uint8 [height][width][4] (B, G, R, alpha), row 0 = top."""
        fb = np.zeros((self.accum.height, self.accum.width, 4), np.uint8)
        check(lib.tpt_exchange_read_framebuffer(self._h, _ptr(fb)))
        return fb

    def status(self):
        """Synchronises; raises if a wait for another rank timed out."""
        check(lib.tpt_exchange_status(self._h))

    def close(self):
        if self._h:
            lib.tpt_exchange_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DeviceScene:
    """Device copy of the reference's DeviceScene (include/mesh.cuh:80-96); arrays as produced by
    Scene::copySceneToDevice."""

    @staticmethod
    def _desc(scene: dict):
        keep = dict(
            indices=np.ascontiguousarray(scene["indices"], np.uint32).reshape(-1),
            vertices=np.ascontiguousarray(scene["vertices"], np.float32).reshape(-1, 3),
            normals=np.ascontiguousarray(scene["normals"], np.float32).reshape(-1, 3),
            materials=np.ascontiguousarray(scene["materials"]).view(MATERIAL_DTYPE).reshape(-1),
            objects=np.ascontiguousarray(scene["objects"]).view(MTL_INTERVAL_DTYPE).reshape(-1),
            vert_trans=np.ascontiguousarray(scene["vert_trans"], np.float32).reshape(-1, 16),
            normal_trans=np.ascontiguousarray(scene["normal_trans"], np.float32).reshape(-1, 16),
            lights=np.ascontiguousarray(scene["lights"]).view(LIGHT_DTYPE).reshape(-1),
        )
        d = capi.SceneDesc()
        d.indices, d.n_indices = _ptr(keep["indices"]), keep["indices"].size
        d.vertices, d.normals, d.n_vertices = _ptr(keep["vertices"]), _ptr(keep["normals"]), keep["vertices"].shape[0]
        d.materials, d.n_materials = _ptr(keep["materials"]), keep["materials"].shape[0]
        d.objects, d.n_objects = _ptr(keep["objects"]), keep["objects"].shape[0]
        d.vert_trans, d.normal_trans = _ptr(keep["vert_trans"]), _ptr(keep["normal_trans"])
        d.lights, d.n_lights = (_ptr(keep["lights"]) if keep["lights"].size else None), keep["lights"].shape[0]
        return d, keep

    def __init__(self, ctx: Context, scene: dict):
        self.ctx = ctx
        d, keep = self._desc(scene)
        h = C.c_void_p()
        check(lib.tpt_scene_create(ctx._h, C.byref(d), C.byref(h)))
        self._h = h
        ctx._children.add(self)
        self.n_faces = keep["indices"].size // 3
        self.n_vertices = keep["vertices"].shape[0]
        self.n_objects = keep["objects"].shape[0]

    @staticmethod
    def prepare(scene: dict, pinned: bool = False):
        """The scene's arrays laid out as the C ABI takes them (tpt_scene_desc + the arrays it points into): for callers that
        upload the same host arrays every frame (update_prepared).  pinned: the arrays live in page-locked memory (torch), so
        the uploads are truly asynchronous."""
        if not pinned:
            return DeviceScene._desc(scene)
        import torch
        keep_t, staged = [], dict(scene)
        for k in ("indices", "vertices", "normals", "materials", "objects", "vert_trans", "normal_trans", "lights"):
            a = np.ascontiguousarray(scene[k])
            if a.size == 0:
                continue
            t = torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).pin_memory()
            keep_t.append(t)
            staged[k] = t.numpy().view(a.dtype).reshape(a.shape)
        d, keep = DeviceScene._desc(staged)
        keep["_pinned"] = keep_t
        return d, keep

    def update_prepared(self, prepared):
        check(lib.tpt_scene_update(self._h, C.byref(prepared[0])))

    def update(self, scene: dict):
        """tpt_scene_update: re-upload host arrays of the same sizes into the existing device arrays (no allocation, no sync)."""
        d, keep = self._desc(scene)
        check(lib.tpt_scene_update(self._h, C.byref(d)))

    def close(self):
        if self._h:
            lib.tpt_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_transforms(self, vert_trans, normal_trans):
        vt = np.ascontiguousarray(vert_trans, np.float32).reshape(self.n_objects, 16)
        nt = np.ascontiguousarray(normal_trans, np.float32).reshape(self.n_objects, 16)
        check(lib.tpt_scene_set_transforms(self._h, _ptr(vt), _ptr(nt)))

    def world(self):
        wv = np.zeros((self.n_vertices, 3), np.float32)
        wn = np.zeros((self.n_vertices, 3), np.float32)
        check(lib.tpt_scene_export_world(self._h, _ptr(wv), _ptr(wn)))
        return wv, wn


class BVH:
    """include/bvh.cuh:60-68.  construct() = `transform` kernel + BVH::construct of the reference."""

    def __init__(self, scene: DeviceScene):
        self.scene = scene
        self.stats = None

    def refit(self, stats: bool = True):
        """After DeviceScene.set_transforms: reuse the hierarchy of the last construct(), redo transform + leaf records + boxes."""
        return self.construct(stats, _fn=lib.tpt_bvh_refit)

    def construct(self, stats: bool = True, _fn=None):
        build = _fn or lib.tpt_bvh_build
        if stats:
            st = capi.BvhStats()
            check(build(self.scene._h, C.byref(st)))
            self.stats = dict(build_ms=st.build_ms, n_faces=st.n_faces, n_nodes=st.n_nodes, max_depth=st.max_depth,
                              sah_cost=st.sah_cost, bounds_min=tuple(st.bounds_min), bounds_max=tuple(st.bounds_max))
        else:
            check(build(self.scene._h, None))
        return self.stats

    @property
    def m_nodes(self) -> np.ndarray:
        """Nodes in the reference's 36-byte layout (internal [0,F-1), leaves [F-1,2F-1))."""
        n = 2 * self.scene.n_faces - 1
        out = np.zeros(n, REF_NODE_DTYPE)
        check(lib.tpt_bvh_export_reference_nodes(self.scene._h, _ptr(out), n))
        return out

    def order(self):
        """(keys, faces) of the sorted leaves: m_keys after the sort (include/bvh.cuh:67, src/bvh.cu:326) and each leaf's face id."""
        keys, faces = np.zeros(self.scene.n_faces, np.uint64), np.zeros(self.scene.n_faces, np.uint32)
        check(lib.tpt_bvh_export_order(self.scene._h, _ptr(keys), _ptr(faces)))
        return keys, faces

    def trace(self, rays: np.ndarray, any_hit: bool = False, counted: bool = False, out: np.ndarray | None = None) -> np.ndarray:
        """Host buffers in, host buffers out (`out`: caller's HIT_DTYPE array, e.g. pinned memory)."""
        rays = np.ascontiguousarray(rays)
        if rays.dtype != RAY_DTYPE:
            rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6).view(RAY_DTYPE).reshape(-1)
        hits = np.zeros(rays.shape[0], HIT_DTYPE) if out is None else out
        fn = lib.tpt_trace_rays_counted if counted else lib.tpt_trace_rays
        check(fn(self.scene._h, _ptr(rays), rays.shape[0], _ptr(hits), int(any_hit), 0))
        return hits


    def trace_device(self, rays_ptr: int, n: int, hits_ptr: int, any_hit: bool = False, counted: bool = False):
        """Same entry with DEVICE buffers (n tpt_ray in, n tpt_hit out), asynchronous on the context stream."""
        fn = lib.tpt_trace_rays_counted if counted else lib.tpt_trace_rays
        check(fn(self.scene._h, C.c_void_p(rays_ptr), n, C.c_void_p(hits_ptr), int(any_hit), 1))


class EnvLight:
    """include/env_light.cuh:9-19: RGBA8 radiance texture, row 0 = bottom scanline; or a linear float RGB
    map [h, w, 3] for HDR environments."""

    def __init__(self, ctx: Context, rgba8: np.ndarray | None):
        self.ctx = ctx
        if rgba8 is None:
            check(lib.tpt_env_clear(ctx._h))
        elif np.asarray(rgba8).dtype.kind == "f":
            img = np.ascontiguousarray(rgba8, np.float32)
            assert img.ndim == 3 and img.shape[2] == 3
            check(lib.tpt_env_set_rgb32f(ctx._h, _ptr(img), img.shape[1], img.shape[0]))
        else:
            img = np.ascontiguousarray(rgba8, np.uint8)
            assert img.ndim == 3 and img.shape[2] == 4
            check(lib.tpt_env_set_rgba8(ctx._h, _ptr(img), img.shape[1], img.shape[0]))
        self.shape = None if rgba8 is None else (int(rgba8.shape[0]), int(rgba8.shape[1]))

    def sample(self, n: int, seed: int = 1) -> np.ndarray:
        """EnvLightSample (include/env_light.cuh:21-52; no call site in the reference): n x (dir.xyz, pdf per steradian)
        drawn on the device from the map's luminance*sin(theta) alias tables."""
        out = np.zeros((n, 4), np.float32)
        check(lib.tpt_env_sample(self.ctx._h, n, seed, _ptr(out)))
        return out

    def sampler_tables(self):
        """(prob, alias, mass) of the alias tables, each (height, width), row 0 = bottom."""
        h, w = self.shape
        prob, alias, mass = np.zeros((h, w), np.float32), np.zeros((h, w), np.uint32), np.zeros((h, w), np.float32)
        check(lib.tpt_env_sampler_tables(self.ctx._h, _ptr(prob), _ptr(alias), _ptr(mass)))
        return prob, alias, mass


class Accum:
    def __init__(self, ctx: Context, width: int, height: int, device_ptr: int | None = None):
        h = C.c_void_p()
        if device_ptr is None:
            check(lib.tpt_accum_create(ctx._h, width, height, C.byref(h)))
        else:
            check(lib.tpt_accum_wrap(ctx._h, C.c_void_p(device_ptr), width, height, C.byref(h)))
        self._h = h
        self.ctx = ctx
        ctx._children.add(self)
        self.width, self.height = width, height

    def close(self):
        if self._h:
            lib.tpt_accum_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def clear(self):
        check(lib.tpt_accum_clear(self._h))

    def add(self, other: "Accum"):
        check(lib.tpt_accum_add(self._h, other._h))

    def device_ptr(self) -> int:
        return int(lib.tpt_accum_device_ptr(self._h) or 0)

    def to_f32(self) -> np.ndarray:
        out = np.zeros((self.height, self.width, 3), np.float32)
        check(lib.tpt_accum_export_f32(self._h, _ptr(out)))
        return out

    def to_i64(self) -> np.ndarray:
        out = np.zeros((self.height, self.width, 3), np.int64)
        check(lib.tpt_accum_export_i64(self._h, _ptr(out)))
        return out

    def from_i64(self, a: np.ndarray):
        a = np.ascontiguousarray(a, np.int64).reshape(self.height, self.width, 3)
        check(lib.tpt_accum_import_i64(self._h, _ptr(a)))

    def resolve_bgra8(self, spp_total: int, fb: np.ndarray | None = None) -> np.ndarray:
        if fb is None:
            fb = np.zeros((self.height, self.width, 4), np.uint8)
        check(lib.tpt_resolve_bgra8(self._h, spp_total, _ptr(fb), 0))
        return fb


def render_config(width, height, spp, max_depth=8, seed=1, sample_offset=0, sample_stride=1, counters=False, paths_in_flight=0,
                  stage_times=False, env_nee=False, pixel_candidates=True, rough_metals=False, refit=False, specular_continuation=True):
    c = capi.RenderConfig()
    c.width, c.height, c.spp, c.max_depth = width, height, spp, max_depth
    c.seed = seed
    c.sample_offset, c.sample_stride = sample_offset, sample_stride
    c.flags = (capi.TPT_RENDER_COUNTERS if counters else 0) | (capi.TPT_RENDER_STAGE_TIMES if stage_times else 0) | \
              (capi.TPT_RENDER_ENV_NEE if env_nee else 0) | (0 if pixel_candidates else capi.TPT_RENDER_NO_PIXEL_CANDIDATES) | \
              (capi.TPT_RENDER_ROUGH_METALS if rough_metals else 0) | (capi.TPT_RENDER_REFIT if refit else 0) | \
              (0 if specular_continuation else capi.TPT_RENDER_NO_SPECULAR_CONTINUATION)
    c.paths_in_flight = paths_in_flight
    return c


class PathTracer:
    """include/path_tracer.h:16-35.  The reference hard-codes 1920x1080 (include/vkEngine.h:24), 64 spp
    (src/path_tracer.cu:559) and depth 8 (:17); here they are constructor/call arguments."""

    def __init__(self, ctx: Context, width: int = 1920, height: int = 1080, env_rgba8: np.ndarray | None = None):
        self.ctx = ctx
        self.m_width, self.m_height = width, height
        self.envLight = EnvLight(ctx, env_rgba8)
        self.m_radiance = Accum(ctx, width, height)

    def render_into(self, scene: DeviceScene, camera: Camera, spp: int, max_depth: int = 8, seed: int = 1, sample_offset: int = 0,
                    sample_stride: int = 1, counters: bool = False, paths_in_flight: int = 0, accum: Accum | None = None,
                    stage_times: bool = False, env_nee: bool = False, pixel_candidates: bool = True, rough_metals: bool = False,
                    specular_continuation: bool = True):
        cfg = render_config(self.m_width, self.m_height, spp, max_depth, seed, sample_offset, sample_stride, counters, paths_in_flight,
                            stage_times, env_nee, pixel_candidates, rough_metals, specular_continuation=specular_continuation)
        acc = accum or self.m_radiance
        check(lib.tpt_render(scene._h, C.byref(camera.c), C.byref(cfg), acc._h))

    def doTrace(self, scene: DeviceScene, camera: Camera, framebuffer: np.ndarray, nSamplesPerPixel: int, max_depth: int = 8, seed: int = 1,
                refit: bool = False):
        """One frame exactly as src/path_tracer.cu:491-554: transform + BVH rebuild, clear, trace, copyToFB."""
        cfg = render_config(self.m_width, self.m_height, nSamplesPerPixel, max_depth, seed, refit=refit)
        check(lib.tpt_do_trace(scene._h, C.byref(camera.c), C.byref(cfg), self.m_radiance._h, _ptr(framebuffer)))
        return framebuffer
