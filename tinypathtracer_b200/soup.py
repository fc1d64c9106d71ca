"""Synthetic instanced triangle soup — BASELINE.json configs[3]: 8 192 instances of ball.gltf's
sphere mesh (1 216 triangles / 660 vertices) + 3 211 unit boxes (12 triangles) = 10 000 004 triangles, each
instance with a uniform random translation in [-100, 100]^3, scale U[0.5, 2] and a uniform random rotation
(normalised Gaussian quaternion), generated with numpy's PCG64 seeded 0x5EED.  All |coordinates| < 256.

Host-side input generation only (the arrays of a DeviceScene, one object per instance): the transforms are
applied on the GPU by the build's transform kernel, like any other scene.
"""
from __future__ import annotations

import numpy as np

from .api import LIGHT_DTYPE, MATERIAL_DTYPE, MTL_INTERVAL_DTYPE

SEED = 0x5EED


def _unit_box():
    """24 vertices (4 per face, per-face normals), 12 triangles, centred, edge 1."""
    v, n, idx = [], [], []
    for axis in range(3):
        for sign in (-1.0, 1.0):
            u, w = (axis + 1) % 3, (axis + 2) % 3
            base = len(v)
            for a, b in ((-1, -1), (1, -1), (1, 1), (-1, 1)):
                p = [0.0, 0.0, 0.0]
                p[axis], p[u], p[w] = 0.5 * sign, 0.5 * a, 0.5 * b
                v.append(p)
                nn = [0.0, 0.0, 0.0]
                nn[axis] = sign
                n.append(nn)
            idx += [base, base + 1, base + 2, base, base + 2, base + 3]
    return np.array(v, np.float32), np.array(n, np.float32), np.array(idx, np.uint32)


def _trs(t, q, s):
    """Column-major TRS matrices (include/transform.h:28-33) and inverse-transpose normal matrices, batched."""
    x, y, z, w = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.empty((len(t), 3, 3), np.float64)
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - z * w); R[:, 0, 2] = 2 * (x * z + y * w)
    R[:, 1, 0] = 2 * (x * y + z * w); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - x * w)
    R[:, 2, 0] = 2 * (x * z - y * w); R[:, 2, 1] = 2 * (y * z + x * w); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    M = np.zeros((len(t), 4, 4), np.float64)
    M[:, :3, :3] = R * s[:, None, None]
    M[:, :3, 3] = t
    M[:, 3, 3] = 1.0
    N = np.zeros_like(M)
    N[:, :3, :3] = R / s[:, None, None]     # inverse transpose of R*s for a uniform scale
    N[:, 3, 3] = 1.0
    # column-major flattening: element (row r, col c) at index c*4 + r
    return (np.ascontiguousarray(M.transpose(0, 2, 1)).reshape(-1, 16).astype(np.float32),
            np.ascontiguousarray(N.transpose(0, 2, 1)).reshape(-1, 16).astype(np.float32))


def make_soup(ball_scene: dict, n_balls: int = 8192, n_boxes: int = 3211, seed: int = SEED) -> dict:
    rng = np.random.Generator(np.random.PCG64(seed))
    bv = np.asarray(ball_scene["vertices"], np.float32).reshape(-1, 3)
    bn = np.asarray(ball_scene["normals"], np.float32).reshape(-1, 3)
    bi = np.asarray(ball_scene["indices"], np.uint32).reshape(-1)
    xv, xn, xi = _unit_box()
    n_obj = n_balls + n_boxes
    t = rng.uniform(-100.0, 100.0, (n_obj, 3))
    s = rng.uniform(0.5, 2.0, n_obj)
    q = rng.normal(size=(n_obj, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    vert_trans, normal_trans = _trs(t, q, s)

    nbv, nbf = bv.shape[0], bi.size // 3
    nxv, nxf = xv.shape[0], xi.size // 3
    vertices = np.concatenate([np.tile(bv, (n_balls, 1)), np.tile(xv, (n_boxes, 1))])
    normals = np.concatenate([np.tile(bn, (n_balls, 1)), np.tile(xn, (n_boxes, 1))])
    ball_idx = (bi[None, :].astype(np.int64) + (np.arange(n_balls, dtype=np.int64) * nbv)[:, None]).reshape(-1)
    box_idx = (xi[None, :].astype(np.int64) + (n_balls * nbv + np.arange(n_boxes, dtype=np.int64) * nxv)[:, None]).reshape(-1)
    indices = np.concatenate([ball_idx, box_idx]).astype(np.uint32)
    objects = np.zeros(n_obj, MTL_INTERVAL_DTYPE)
    objects["begin"][:n_balls] = np.arange(n_balls) * nbf
    objects["begin"][n_balls:] = n_balls * nbf + np.arange(n_boxes) * nxf
    objects["mtl"] = np.arange(n_obj) % 4
    materials = np.zeros(4, MATERIAL_DTYPE)
    materials["base"] = [[0.8, 0.8, 0.8], [0.8, 0.3, 0.3], [0.3, 0.8, 0.3], [0.3, 0.3, 0.8]]
    c2w = np.eye(4, dtype=np.float32)
    c2w[2, 3] = 260.0       # camera at (0, 0, 260) looking down -Z
    return dict(indices=indices, vertices=vertices, normals=normals, materials=materials, objects=objects, vert_trans=vert_trans,
                normal_trans=normal_trans, lights=np.zeros(0, LIGHT_DTYPE), vfov=np.float32(0.7), aspect=np.float32(16.0 / 9.0),
                near=np.float32(0.1), c2w=np.ascontiguousarray(c2w.T).reshape(16))


def secondary_rays(rays: np.ndarray, hits: np.ndarray, world_verts: np.ndarray, indices: np.ndarray, seed: int = SEED) -> np.ndarray:
    """One cosine-distributed ray per primary hit, leaving the hit point about the geometric normal of the hit
    triangle (flipped towards the incoming ray); seeded, generated once, so that every implementation traces the
    identical ray set.  rays/hits: structured arrays (RAY_DTYPE / HIT_DTYPE); world_verts from DeviceScene.world()."""
    rng = np.random.Generator(np.random.PCG64(seed + 1))
    m = hits["prim"] >= 0
    din = rays["dir"][m]
    o = rays["origin"][m] + din * hits["t"][m][:, None]
    tri = np.asarray(indices).reshape(-1, 3)[hits["prim"][m]]
    v0, v1, v2 = world_verts[tri[:, 0]], world_verts[tri[:, 1]], world_verts[tri[:, 2]]
    n = np.cross(v1 - v0, v2 - v0)
    n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-30)
    n *= np.where((n * din).sum(1, keepdims=True) > 0.0, -1.0, 1.0)
    a = np.where(np.abs(n[:, :1]) < 0.9, np.array([[1.0, 0.0, 0.0]], np.float32), np.array([[0.0, 1.0, 0.0]], np.float32))
    tx = np.cross(n, a)
    tx /= np.linalg.norm(tx, axis=1, keepdims=True)
    ty = np.cross(n, tx)
    u1, u2 = rng.random(len(o), dtype=np.float32), rng.random(len(o), dtype=np.float32)
    r, phi = np.sqrt(u1), 2.0 * np.pi * u2
    d = tx * (r * np.cos(phi))[:, None] + ty * (r * np.sin(phi))[:, None] + n * np.sqrt(np.maximum(0.0, 1.0 - u1))[:, None]
    out = np.zeros(len(o), rays.dtype)
    out["origin"] = (o + n * 1e-3).astype(np.float32)
    out["dir"] = d.astype(np.float32)
    return out
