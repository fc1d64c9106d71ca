"""BASELINE.json configs[3] at its own size: the 10 000 004-triangle instanced soup (tinypathtracer_b200/soup.py).

This is the only place where the 16-bit-per-axis Morton key branch (csrc/bvh.cu: taken from 2^20 triangles on), the own radix
sort at scale and a large refit frontier run, so the tree is checked structurally (checkBVHNodes' invariant, reference
include/debug_utils.h:51-83, + exact boxes + every face once) and 20 000 randomly chosen rays of EACH of the two bench ray
sets (3840 x 2160 pixel-centre primary rays; one cosine secondary ray per primary hit) are compared with the brute-force
oracle over all 10 M triangles (oracle mode 2 = brute_force() evaluated block-wise with vector arithmetic, bit-identical to
mode 0): primitive ids exact, t / u / v bit-equal — the bar of every other traversal test here (reference rule:
src/path_tracer.cu:61-107, :83; geometry_queries.h:66-86)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N_CHECK = 20000


def test_soup_10m_tree_and_hits_vs_bruteforce(tpt, ctx, scenes, orc):
    from tinypathtracer_b200.soup import make_soup, secondary_rays
    scene = make_soup(scenes["ball"])
    idx = np.asarray(scene["indices"], np.uint32)
    F = idx.size // 3
    assert F == 10_000_004
    ds = tpt.DeviceScene(ctx, scene)
    bvh = tpt.BVH(ds)
    st = bvh.construct(stats=True)
    assert st["n_faces"] == F and st["n_nodes"] == F - 1 and st["max_depth"] < 96
    wv, _ = ds.world()
    owv, _ = orc.transform(scene)
    assert np.array_equal(wv, owv)                       # the transform kernel at 5.5 M vertices: bit-exact vs the oracle

    # ---- structure (vectorised form of test_bvh_structure)
    nodes = bvh.m_nodes
    n_int = F - 1
    a, b, parent = nodes["a"][:n_int].astype(np.int64), nodes["b"][:n_int].astype(np.int64), nodes["parent"].astype(np.int64)
    assert a.min() >= 0 and b.min() >= 0 and a.max() < 2 * F - 1 and b.max() < 2 * F - 1
    cnt = np.bincount(np.concatenate([a, b]), minlength=2 * F - 1)
    assert cnt[0] == 0 and (cnt[1:] == 1).all()          # every node but the root is the child of exactly one node
    me = np.arange(n_int)
    assert (parent[a] == me).all() and (parent[b] == me).all()
    mn, mx = nodes["mn"], nodes["mx"]
    assert (mn[:n_int] == np.minimum(mn[a], mn[b])).all() and (mx[:n_int] == np.maximum(mx[a], mx[b])).all()
    leaf_f = nodes["a"][n_int:].astype(np.int64)
    assert (np.sort(leaf_f) == np.arange(F)).all()       # every face in exactly one leaf
    tri = wv[idx.reshape(-1, 3)[leaf_f]]
    assert (mn[n_int:] == tri.min(axis=1)).all() and (mx[n_int:] == tri.max(axis=1)).all()
    del nodes, mn, mx, tri, cnt, parent, a, b

    # ---- the two ray sets of bench.py --workload soup, traced whole; a random sample of each against brute force
    W, H = 3840, 2160
    rays = ctx.primary_rays(tpt.Camera.from_scene(scene), W, H)
    hits = bvh.trace(rays)
    assert 0.2 < (hits["prim"] >= 0).mean() < 1.0
    sec = secondary_rays(rays, hits, wv, idx)
    sec_hits = bvh.trace(sec)
    rng = np.random.default_rng(77)
    for name, r, h in (("primary", rays, hits), ("secondary", sec, sec_hits)):
        pick = np.sort(rng.choice(r.shape[0], N_CHECK, replace=False))
        r6 = np.concatenate([r["origin"][pick], r["dir"][pick]], axis=1)
        ref = orc.trace_rays(wv, idx, r6, mode=2)
        got = h[pick]
        assert (ref["prim"] >= 0).sum() > N_CHECK // 10, name
        bad = np.flatnonzero(got["prim"] != ref["prim"])
        assert bad.size == 0, f"{name}: {bad.size} of {N_CHECK} primitive ids differ from brute force, first at ray {pick[bad[:5]]}"
        for f in ("t", "u", "v"):
            assert np.array_equal(got[f].view(np.uint32), ref[f].view(np.uint32)), f"{name}: {f} not bit-equal"
    ds.close()
