"""BASELINE.json configs[3] at its own size: the 10 000 004-triangle instanced soup (tinypathtracer_b200/soup.py).

This is the only place where the 16-bit-per-axis Morton key branch (csrc/bvh.cu: taken from 2^20 triangles on), the own radix
sort at scale and a large refit frontier run, so the tree is checked structurally (checkBVHNodes' invariant, reference
include/debug_utils.h:51-83, + exact boxes + every face once) and 20 000 randomly chosen rays of EACH of the two bench ray
sets (3840 x 2160 pixel-centre primary rays; one cosine secondary ray per primary hit) are compared with the brute-force
oracle over all 10 M triangles (oracle mode 2 = brute_force() evaluated block-wise with vector arithmetic, bit-identical to
mode 0): primitive ids exact, t / u / v bit-equal — the bar of every other traversal test here (reference rule:
src/path_tracer.cu:61-107, :83; geometry_queries.h:66-86) — except for edge-grazing hits, which are classified, counted and
bounded (see the comment in the test)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N_CHECK = 20000


def test_soup_10m_tree_and_hits_vs_bruteforce(tpt, ctx, scenes, orc, capsys):
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

    # ---- the own radix sort at scale (48-bit keys, 6 passes, 2 442 tiles): the unique stable order
    keys, faces = bvh.order()
    dk = np.diff(keys.astype(np.int64))
    assert (dk >= 0).all() and (np.diff(faces.astype(np.int64))[dk == 0] > 0).all()
    assert keys.max() < (1 << 48)
    unsorted = np.empty(F, np.uint64)
    unsorted[faces] = keys
    assert np.array_equal(np.argsort(unsorted, kind="stable").astype(np.uint32), faces)
    del keys, faces, dk, unsorted

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
    report = []
    for name, r, h in (("primary", rays, hits), ("secondary", sec, sec_hits)):
        pick = np.sort(rng.choice(r.shape[0], N_CHECK, replace=False))
        r6 = np.concatenate([r["origin"][pick], r["dir"][pick]], axis=1)
        ref = orc.trace_rays(wv, idx, r6, mode=2)
        got = h[pick]
        assert (ref["prim"] >= 0).sum() > N_CHECK // 10, name
        same = got["prim"] == ref["prim"]
        for f in ("t", "u", "v"):   # wherever the ids agree: bit-equal
            assert np.array_equal(got[f][same].view(np.uint32), ref[f][same].view(np.uint32)), f"{name}: {f} not bit-equal"
        # Where they differ, the brute-force hit must be an EDGE-GRAZING one.  At this scale (triangles of ~0.1 units seen from
        # ~300) fp32 Möller-Trumbore barycentrics carry an error of ~5e-4, so the arithmetic accepts rays that pass a
        # silhouette edge on the outside by up to ~3e-3 units, while the ray misses the triangle's exact bounding box by more
        # than the slab test's rounding slack: a tree traversal (the reference's own included: its rayHitBBox has no slack at
        # all) cannot see such a hit.  SURVEY §8a's harness rule: a mismatch is legitimate iff the oracle confirms that the
        # disputed triangle is hit only within the edge tolerance and that the product's own hit is a true hit.
        n_graze = 0
        for i in np.flatnonzero(~same):
            g, rr = int(got["prim"][i]), int(ref["prim"][i])
            assert rr >= 0, f"{name} ray {pick[i]}: the product reports a hit ({g}) where brute force finds none"
            robust, _ = orc.retest_triangle(wv, idx, rr, r6[i], -2e-3)     # still a hit 2e-3 INSIDE the edges?
            assert not robust, f"{name} ray {pick[i]}: missed triangle {rr}, a hit well inside its edges (t={ref['t'][i]})"
            if g >= 0:
                ok, t_g = orc.retest_triangle(wv, idx, g, r6[i], 0.0)
                assert ok and np.float32(t_g) == got["t"][i] and got["t"][i] >= ref["t"][i], f"{name} ray {pick[i]}: own hit {g} not confirmed"
            n_graze += 1
        assert n_graze <= 1e-3 * N_CHECK, f"{name}: {n_graze} edge-grazing differences in {N_CHECK} rays"
        report.append(f"{name}: {int(same.sum())}/{N_CHECK} rays identical to brute force over {F} triangles (id, t, u, v bit-equal), "
                      f"{n_graze} edge-grazing")
    with capsys.disabled():
        print("\n[soup 10M] " + "; ".join(report))
    ds.close()
