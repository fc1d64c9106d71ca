"""Generates tests/golden/ref_host_vectors.npz: outputs of the REFERENCE's own host-callable functions
(oracle/_ref/libref_host.so, compiled from /root/reference by oracle/ref_build.py) on seeded inputs.  The CPU test
suite checks the oracle restatement against these vectors bit for bit, so the oracle stays pinned to the reference
even where /root/reference is not mounted (the GPU box).

    python oracle/ref_build.py libref_host.so && python tests/golden/make_ref_vectors.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import gltf_scene, pyoracle as orc, pyref  # noqa: E402


def main():
    rng = np.random.default_rng(20261104)
    out = {}
    x = np.concatenate([rng.uniform(1e-6, 1e4, 900), [1.0, 2.0, 4.0, 0.25, 1e-12, 3e38]]).astype(np.float32)
    out["frsqrt_in"] = x
    out["frsqrt_out"] = np.array([pyref.frsqrt(float(v)) for v in x], np.float32)
    v = rng.normal(size=(500, 3)).astype(np.float32) * rng.uniform(0.01, 50, (500, 1)).astype(np.float32)
    out["normalize_in"] = v
    out["normalize_out"] = np.stack([pyref.normalize(a) for a in v])
    pts = np.concatenate([rng.uniform(-300, 300, (1500, 3)), rng.uniform(-1, 1, (400, 3)), rng.uniform(-1e-3, 1e-3, (80, 3)),
                          [[0, 0, 0], [256, -256, 255.999], [1e6, -1e6, 0.5], [1, -1, 2], [6e-10, -6e-10, 9e-10]]]).astype(np.float32)
    # (0 < |x| < 2^-31 is left out: floatTo21Int shifts by >= 32 there, src/bvh.cu:39 — undefined on the host,
    #  0 on the GPU; the oracle restates the GPU's result)
    out["morton_in"] = pts
    out["morton_out"] = np.array([pyref.morton(*map(float, p)) for p in pts], np.int64)
    # ray / triangle, ray / box
    tris = rng.uniform(-2, 2, (2000, 9)).astype(np.float32)
    rays = np.concatenate([rng.uniform(-3, 3, (2000, 3)), rng.normal(size=(2000, 3))], axis=1).astype(np.float32)
    # aim half of the rays at their triangle so that many tests hit, some through vertices / edges
    bary = rng.dirichlet([1, 1, 1], 1000).astype(np.float32)
    bary[:60] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 60)]
    tgt = (tris[:1000].reshape(-1, 3, 3) * bary[:, :, None]).sum(axis=1)
    rays[:1000, 3:] = tgt - rays[:1000, :3]
    res = [pyref.ray_hit_triangle(r, t) for r, t in zip(rays, tris)]
    out["tri_rays"], out["tri_tris"] = rays, tris
    out["tri_ok"] = np.array([a for a, _ in res], np.uint8)
    out["tri_out"] = np.stack([b for _, b in res])
    lo = rng.uniform(-2, 1, (2000, 3)).astype(np.float32)
    hi = (lo + rng.uniform(0, 2, (2000, 3))).astype(np.float32)
    hi[:200, rng.integers(0, 3)] = lo[:200, rng.integers(0, 3)]                      # flat boxes
    brays = np.concatenate([rng.uniform(-3, 3, (2000, 3)), rng.normal(size=(2000, 3))], axis=1).astype(np.float32)
    brays[:300, 3:] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 300)] * rng.choice([-1.0, 1.0], (300, 1))   # axis parallel: inf / NaN slabs
    brays[:100, 0] = lo[:100, 0]                                                        # origin on a slab plane
    out["box_rays"], out["box_lo"], out["box_hi"] = brays, lo, hi
    out["box_ok"] = np.array([pyref.ray_hit_bbox(r, a, b) for r, a, b in zip(brays, lo, hi)], np.uint8)
    # BSDF pieces
    d = rng.normal(size=(1000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    n = rng.normal(size=(1000, 3)).astype(np.float32)
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    ior = rng.choice([1.33, 1.5, 2.0, 0.7], 1000).astype(np.float32)
    out["bsdf_d"], out["bsdf_n"], out["bsdf_ior"] = d, n, ior
    out["reflect_out"] = np.stack([pyref.reflect(a, b) for a, b in zip(d, n)])
    out["refract_out"] = np.stack([pyref.refract(a, b, float(c)) for a, b, c in zip(d, n, ior)])
    cs = rng.uniform(0, 1, 1000).astype(np.float32)
    et = rng.choice([0.5, 2.0, 1 / 1.5, 1.5], 1000).astype(np.float32)
    out["schlick_in"] = np.stack([cs, et], 1)
    out["schlick_out"] = np.array([pyref.schlick(float(a), float(b)) for a, b in zip(cs, et)], np.float32)
    # delta lights
    lights = np.zeros(900, orc.LIGHT_DTYPE)
    lights["type"] = np.repeat([0, 1, 2], 300)
    lights["color"] = rng.uniform(0, 1, (900, 3))
    lights["intensity"] = rng.uniform(0.1, 30, 900)
    lights["a"] = rng.uniform(-6, 6, (900, 3))
    bdir = rng.normal(size=(900, 3))
    lights["b"] = bdir / np.linalg.norm(bdir, axis=1, keepdims=True)
    lights["cos_outer"] = rng.uniform(0.5, 0.95, 900)
    lights["inv_cone_diff"] = 1.0 / rng.uniform(0.01, 0.04, 900)
    lp = rng.uniform(-6, 6, (900, 3)).astype(np.float32)
    out["light_rec"] = lights.view(np.uint8).reshape(900, 52)
    out["light_p"] = lp
    out["light_out"] = np.stack([pyref.light_sample(L, p) for L, p in zip(lights, lp)])
    dd = rng.normal(size=(1000, 3)).astype(np.float32) * rng.uniform(0.9, 1.1, (1000, 1)).astype(np.float32)
    dd[:6] = [[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]]
    out["uv_in"] = dd
    out["uv_out"] = np.stack([pyref.vec2uv(a) for a in dd])
    rgb = rng.uniform(-0.5, 40, (500, 3)).astype(np.float32)
    out["uchar_in"] = rgb
    out["uchar_out"] = np.stack([pyref.to_uchar(a, 16) for a in rgb])
    t = rng.uniform(-5, 5, (100, 3)).astype(np.float32)
    q = rng.normal(size=(100, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    q[:3] = 0
    s = rng.uniform(0.2, 3, (100, 3)).astype(np.float32)
    out["trs_t"], out["trs_q"], out["trs_s"] = t, q, s
    out["trs_out"] = np.stack([pyref.trs_to_mat4(a, b, c) for a, b, c in zip(t, q, s)])
    out["nm_out"] = np.stack([pyref.normal_matrix(m) for m in out["trs_out"]])
    # traverseBVH (src/path_tracer.cu:61-107) on box.gltf over the oracle-built hierarchy in the reference's layout
    scene = gltf_scene.load_scene(ROOT / "tests" / "golden" / "scenes" / "box.npz")
    wv, _ = orc.transform(scene)
    nodes, _, rc = orc.build_reference_bvh(wv, scene["indices"])
    assert rc == 0
    prim = orc.primary_rays(scene, 160, 90)
    o = rng.uniform(wv.min(0) - 0.3, wv.max(0) + 0.3, (6000, 3)).astype(np.float32)
    dr = rng.normal(size=(6000, 3)).astype(np.float32)
    trays = np.concatenate([prim, np.concatenate([o, dr], 1)]).astype(np.float32)
    hits = pyref.traverse(nodes, wv, scene["indices"], trays)
    out["trav_rays"] = trays
    out["trav_hits"] = hits.view(np.uint8).reshape(-1, 16)
    np.savez_compressed(Path(__file__).resolve().parent / "ref_host_vectors.npz", **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
