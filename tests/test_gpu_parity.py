"""Parity of the CUDA path (through the C ABI) against the CPU oracle.  Run with `pytest -m gpu` on a B200.

Bars (BASELINE.json north_star): primitive ids bit-exact for deterministic first hits under the stated
tie-break rule, hit distance t within 1e-5 relative, image rel-MSE <= 1e-3 at high spp.
"""
import numpy as np
import pytest

from conftest import SCENES, synthetic_sky

pytestmark = pytest.mark.gpu

T_REL_TOL = 1e-5      # north_star: "hit distance t within 1e-5 relative"


def classify_hit_mismatches(orc, wv, indices, rays, got, ref):
    """SURVEY §8a tie-break rule.  A primitive-id mismatch is legitimate only if the oracle confirms that the
    ray hits BOTH triangles (edge tolerance 1e-5 in barycentrics) at |dt| <= 1e-5 * t  (shared edge /
    coincident geometry), or that one side's hit is an edge-grazing hit the other rounds to a miss.
    Returns (n_ties, n_grazing, n_bad)."""
    bad = np.nonzero(got["prim"] != ref["prim"])[0]
    ties = grazing = wrong = 0
    for i in bad:
        g, r = int(got["prim"][i]), int(ref["prim"][i])
        if g >= 0 and r >= 0:
            ok, t_g = orc.retest_triangle(wv, indices, g, rays[i], 1e-5)
            if ok and abs(t_g - ref["t"][i]) <= T_REL_TOL * abs(ref["t"][i]):
                ties += 1
            elif ok and t_g > ref["t"][i]:
                wrong += 1          # a strictly farther triangle was reported
            elif ok:
                grazing += 1        # the oracle's own test rejects g only by rounding at an edge
            else:
                wrong += 1
        else:
            h = g if g >= 0 else r   # one side hit, the other missed: must be an edge-grazing hit
            ok, _ = orc.retest_triangle(wv, indices, h, rays[i], 1e-5)
            strict, _ = orc.retest_triangle(wv, indices, h, rays[i], 0.0)
            if ok:
                grazing += 1
            else:
                wrong += 1
    return ties, grazing, wrong


@pytest.fixture(scope="module")
def built(tpt, ctx, scenes, orc):
    out = {}
    for name in SCENES:
        ds = tpt.DeviceScene(ctx, scenes[name])
        bvh = tpt.BVH(ds)
        bvh.construct()
        out[name] = (ds, bvh)
    yield out
    for ds, _ in out.values():
        ds.close()


@pytest.mark.parametrize("name", SCENES)
def test_transform_matches_oracle(name, built, scenes, orc):
    """`transform` kernel (reference src/path_tracer.cu:239-263): world vertices / frsqrt-normalised normals."""
    ds, _ = built[name]
    wv, wn = ds.world()
    ov, on = orc.transform(scenes[name])
    used = np.zeros(wv.shape[0], bool)
    used[scenes[name]["indices"]] = True
    # the geometric core is evaluated with one IEEE rounding per operation in the reference's order: bit-exact
    assert np.array_equal(wv[used], ov[used])
    assert np.array_equal(wn[used], on[used])


@pytest.mark.parametrize("name", SCENES)
def test_bvh_structure(name, built, scenes, orc):
    """checkBVHNodes' invariant (reference include/debug_utils.h:51-83) + box containment + every face once."""
    ds, bvh = built[name]
    nodes = bvh.m_nodes
    F = ds.n_faces
    n_int = F - 1
    child_count = np.zeros(2 * F - 1, int)
    for i in range(n_int):
        for c in (nodes["a"][i], nodes["b"][i]):
            assert 0 <= c < 2 * F - 1
            child_count[c] += 1
            assert nodes["parent"][c] == i
            assert (nodes["mn"][i] <= nodes["mn"][c]).all() and (nodes["mx"][i] >= nodes["mx"][c]).all()
        both_mn = np.minimum(nodes["mn"][nodes["a"][i]], nodes["mn"][nodes["b"][i]])
        both_mx = np.maximum(nodes["mx"][nodes["a"][i]], nodes["mx"][nodes["b"][i]])
        assert (nodes["mn"][i] == both_mn).all() and (nodes["mx"][i] == both_mx).all()
    assert child_count[0] == 0 and (child_count[1:] == 1).all()
    fids = np.sort(nodes["a"][n_int:])
    assert (fids == np.arange(F)).all()
    wv, _ = ds.world()
    tri = wv[scenes[name]["indices"].reshape(-1, 3)]
    leaf = nodes[n_int:]
    assert (leaf["mn"] == tri[leaf["a"]].min(axis=1)).all() and (leaf["mx"] == tri[leaf["a"]].max(axis=1)).all()


def _check_stable_order(keys, faces, F):
    """What a stable sort of (key, face) pairs that start in face order must give: keys ascend, equal keys keep ascending
    face ids, every face once — which determines the permutation uniquely (it is numpy's stable argsort of the keys)."""
    assert keys.shape == (F,) and faces.shape == (F,)
    dk = np.diff(keys.astype(np.int64))              # keys are < 2^63
    assert (dk >= 0).all()
    eq = dk == 0
    assert (np.diff(faces.astype(np.int64))[eq] > 0).all()
    assert (np.sort(faces) == np.arange(F)).all()
    unsorted = np.empty(F, np.uint64)
    unsorted[faces] = keys
    assert np.array_equal(np.argsort(unsorted, kind="stable").astype(np.uint32), faces)


@pytest.mark.parametrize("name", SCENES)
def test_radix_sort_is_the_stable_sort(name, built):
    """Own radix sort (csrc/sort.cuh; replaces thrust::sort_by_key, reference src/bvh.cu:326): single-block path."""
    ds, bvh = built[name]
    keys, faces = bvh.order()
    _check_stable_order(keys, faces, ds.n_faces)


@pytest.mark.parametrize("nb,nx", [(3, 40), (64, 25), (300, 7)])
def test_radix_sort_onesweep_path(nb, nx, tpt, ctx, scenes):
    """4 128 / 78 124 / 364 884 triangles: the onesweep path with 63-bit keys (8 passes), full and partial tiles, and many
    equal keys (instances of one mesh overlap)."""
    from tinypathtracer_b200.soup import make_soup
    scene = make_soup(scenes["ball"], nb, nx)
    ds = tpt.DeviceScene(ctx, scene)
    bvh = tpt.BVH(ds)
    for _ in range(2):                                # the second build reuses (and must re-zero) the sort's scratch
        bvh.construct()
        keys, faces = bvh.order()
        _check_stable_order(keys, faces, ds.n_faces)
    ds.close()


@pytest.mark.parametrize("name", SCENES)
def test_primary_hits_vs_bruteforce(name, tpt, ctx, built, scenes, orc):
    """Deterministic primary rays through pixel centres: id exact (ties classified), t within 1e-5 rel."""
    ds, bvh = built[name]
    W, H = 480, 270
    rays = orc.primary_rays(scenes[name], W, H)
    gpu_rays = ctx.primary_rays(tpt.Camera.from_scene(scenes[name]), W, H)
    g6 = np.concatenate([gpu_rays["origin"], gpu_rays["dir"]], axis=1)
    assert np.array_equal(g6, rays)   # sampleRays incl. the frsqrt normalisation: bit-exact
    wv, _ = orc.transform(scenes[name])
    got = bvh.trace(rays)
    ref = orc.trace_rays(wv, scenes[name]["indices"], rays, mode=0)
    same = got["prim"] == ref["prim"]
    ties, grazing, wrong = classify_hit_mismatches(orc, wv, scenes[name]["indices"], rays, got, ref)
    assert wrong == 0, f"{wrong} real primitive-id mismatches (ties {ties}, grazing {grazing})"
    # Möller–Trumbore is evaluated op-for-op like the oracle and exact-t ties go to the lowest id on both sides,
    # so in fact nothing may differ at all: id, t, u, v bit-exact (the 1e-5 bar of north_star is met with 0).
    assert ties == 0 and grazing == 0 and same.all()
    assert np.array_equal(got["t"], ref["t"]) and np.array_equal(got["u"], ref["u"]) and np.array_equal(got["v"], ref["v"])
    miss = ref["prim"] < 0
    assert (got["t"][miss] == np.float32(3.4028234663852886e38)).all()   # HitStatus() default hitDist = FLT_MAX


@pytest.mark.parametrize("name", ["box", "ball", "box2"])
def test_random_rays_vs_bruteforce(name, built, scenes, orc):
    """Incoherent rays from inside the scene bounds, incl. axis-parallel directions (inf/NaN slabs in the reference)."""
    ds, bvh = built[name]
    wv, _ = orc.transform(scenes[name])
    rng = np.random.default_rng(1234)
    lo, hi = wv.min(axis=0) - 0.5, wv.max(axis=0) + 0.5
    n = 60000
    o = rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:300] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 300)] * rng.choice([-1.0, 1.0], size=(300, 1)).astype(np.float32)
    rays = np.concatenate([o, d], axis=1).astype(np.float32)
    got = bvh.trace(rays)
    ref = orc.trace_rays(wv, scenes[name]["indices"], rays, mode=0)
    ties, grazing, wrong = classify_hit_mismatches(orc, wv, scenes[name]["indices"], rays, got, ref)
    assert wrong == 0 and ties == 0 and grazing == 0, f"mismatches: wrong {wrong}, ties {ties}, grazing {grazing}"
    assert np.array_equal(got["prim"], ref["prim"]) and np.array_equal(got["t"], ref["t"])
    # any-hit agrees with "closest hit exists" (the reference's shadow-ray visibility, src/path_tracer.cu:278-279)
    anyh = bvh.trace(rays, any_hit=True)
    assert np.array_equal(anyh["prim"] >= 0, ref["prim"] >= 0)


def test_counted_trace_matches_uncounted(built, scenes, orc, ctx):
    ds, bvh = built["box"]
    rays = orc.primary_rays(scenes["box"], 160, 90)
    ctx.reset_counters()
    a = bvh.trace(rays)
    b = bvh.trace(rays, counted=True)
    assert (a == b).all()
    c = ctx.counters()
    assert c["node_visits"] > rays.shape[0] and c["tri_tests"] > 0


def _relmse(a, b):
    return float(((a.astype(np.float64) - b) ** 2).sum() / max((b.astype(np.float64) ** 2).sum(), 1e-30))


@pytest.mark.parametrize("name,depth,use_env", [("box", 8, False), ("tir", 32, False), ("ball", 8, True), ("square", 8, False),
                                                ("light", 8, True), ("box2", 8, True), ("box1", 8, True)])
def test_render_same_rng_vs_oracle(name, depth, use_env, tpt, ctx, built, scenes, orc):
    """Same counter-based RNG on both sides: the per-pixel estimates agree except where a 1-ulp difference
    flips a discrete decision.  Small image, low spp, compared on the linear radiance mean."""
    ds, _ = built[name]
    W, H, spp = 96, 54, 32
    env = synthetic_sky(256, 128) if use_env else None
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=env)
    pt.render_into(ds, tpt.Camera.from_scene(scenes[name]), spp=spp, max_depth=depth, seed=11)
    img = pt.m_radiance.to_f32() / spp
    ref = orc.render(scenes[name], W, H, spp, max_depth=depth, seed=11, env=env) / spp
    assert np.isfinite(img).all()
    err = _relmse(img, ref)
    assert err <= 2e-2, f"{name}: rel-MSE at {spp} spp with shared RNG = {err:.3e}"
    assert abs(img.mean() - ref.mean()) <= 0.02 * max(ref.mean(), 1e-3)


@pytest.mark.parametrize("name,depth,use_env", [("box", 8, False), ("tir", 32, False), ("ball", 8, True)])
def test_converged_image_relmse(name, depth, use_env, tpt, ctx, built, scenes, orc):
    """north_star: converged-image relative MSE <= 1e-3 — GPU at high spp vs the oracle at high spp with a
    DIFFERENT seed (independent estimates of the same expectation)."""
    ds, _ = built[name]
    W, H = 48, 27
    env = synthetic_sky(256, 128) if use_env else None
    spp_gpu = spp_cpu = {"box": 8192, "tir": 65536, "ball": 2048}[name]   # enough that MC noise sits below the 1e-3 bar
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=env)
    pt.render_into(ds, tpt.Camera.from_scene(scenes[name]), spp=spp_gpu, max_depth=depth, seed=21)
    img = pt.m_radiance.to_f32() / spp_gpu
    ref = orc.render(scenes[name], W, H, spp_cpu, max_depth=depth, seed=99, env=env) / spp_cpu
    err = _relmse(img, ref)
    assert err <= 1e-3, f"{name}: converged rel-MSE {err:.3e}"


def test_slices_are_bit_identical(tpt, ctx, built, scenes):
    """Multi-GPU contract: 8 fixed spp slices rendered separately and summed give the same int64
    accumulation buffer as one render of all samples — on any number of GPUs, in any order."""
    ds, _ = built["box2"]
    cam = tpt.Camera.from_scene(scenes["box2"])
    W, H, S, per = 80, 45, 8, 4
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=synthetic_sky(128, 64))
    pt.render_into(ds, cam, spp=S * per, seed=5)
    whole = pt.m_radiance.to_i64()
    total = tpt.Accum(ctx, W, H)
    part = tpt.Accum(ctx, W, H)
    for s in reversed(range(S)):
        part.clear()
        pt.render_into(ds, cam, spp=per, seed=5, sample_offset=s, sample_stride=S, accum=part)
        total.add(part)
    assert (total.to_i64() == whole).all()
    # and a different pool size (different batching / scheduling) changes nothing either
    pt.m_radiance.clear()
    pt.render_into(ds, cam, spp=S * per, seed=5, paths_in_flight=W * H * 3)
    assert (pt.m_radiance.to_i64() == whole).all()
    assert ctx.counters()["queue_overflow"] == 0


@pytest.mark.parametrize("name", ["box", "ball", "light", "tir", "box2", "square"])
def test_pixel_candidates_are_bit_identical_to_per_sample_traversal(name, tpt, ctx, built, scenes):
    """k_primary tests each camera sample against its pixel's candidate triangles (one pyramid walk per pixel) instead
    of walking the tree per sample; the closest hit, and with it every later decision of the path, must be the same bits."""
    ds, _ = built[name]
    cam = tpt.Camera.from_scene(scenes[name])
    W, H, spp = 200, 120, 24
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=synthetic_sky(128, 64))
    pt.render_into(ds, cam, spp=spp, seed=9)
    a = pt.m_radiance.to_i64()
    pt.m_radiance.clear()
    pt.render_into(ds, cam, spp=spp, seed=9, pixel_candidates=False)
    assert (pt.m_radiance.to_i64() == a).all()
    assert a.any()
    assert ctx.counters()["queue_overflow"] == 0


def test_resolve_matches_reference_bytes(tpt, ctx, built, scenes, orc):
    """copyToFB (reference src/path_tracer.cu:451-471): truncation, BGR order, Y flip, alpha untouched."""
    ds, _ = built["box"]
    W, H, spp = 64, 36, 8
    pt = tpt.PathTracer(ctx, W, H)
    pt.render_into(ds, tpt.Camera.from_scene(scenes["box"]), spp=spp, seed=3)
    fb = np.full((H, W, 4), 77, np.uint8)
    pt.m_radiance.resolve_bgra8(spp, fb)
    want = orc.resolve_bgra8(pt.m_radiance.to_f32(), spp, np.full((H, W, 4), 77, np.uint8))
    assert (fb == want).all()
    assert (fb[..., 3] == 77).all()


def test_do_trace_one_call(tpt, ctx, built, scenes):
    ds, _ = built["box"]
    W, H = 64, 36
    pt = tpt.PathTracer(ctx, W, H)
    fb = np.zeros((H, W, 4), np.uint8)
    pt.doTrace(ds, tpt.Camera.from_scene(scenes["box"]), fb, 4)
    assert fb[..., :3].max() > 0
    fb2 = np.zeros((H, W, 4), np.uint8)
    pt.doTrace(ds, tpt.Camera.from_scene(scenes["box"]), fb2, 4)
    assert (fb == fb2).all()    # deterministic: fixed seed, rebuilt BVH, exact accumulation


def test_errors_are_loud(tpt, ctx, scenes):
    ds = tpt.DeviceScene(ctx, scenes["tir"])
    bvh = tpt.BVH(ds)
    with pytest.raises(tpt.TptError):
        bvh.trace(np.zeros((4, 6), np.float32))          # trace before build
    pt = tpt.PathTracer(ctx, 32, 18)
    with pytest.raises(tpt.TptError):
        pt.render_into(ds, tpt.Camera.from_scene(scenes["tir"]), spp=1)   # render before build
    bvh.construct()
    with pytest.raises(tpt.TptError):
        pt.render_into(ds, tpt.Camera.from_scene(scenes["tir"]), spp=0)
    bad = dict(scenes["tir"])
    bad["indices"] = bad["indices"].copy()
    bad["indices"][0] = 10 ** 6
    with pytest.raises(tpt.TptError):
        tpt.DeviceScene(ctx, bad)
    ds.close()


def test_degenerate_inputs(tpt, ctx, scenes, orc):
    """Adversarial builds the reference cannot survive: 1 and 2 triangles, all-equal Morton keys,
    runs of >= 3 duplicates, coordinates >= 256."""
    base = scenes["tir"]

    def make(tris):
        tris = np.asarray(tris, np.float32).reshape(-1, 3, 3)
        F = tris.shape[0]
        s = dict(base)
        s["indices"] = np.arange(3 * F, dtype=np.uint32)
        s["vertices"] = tris.reshape(-1, 3).copy()
        s["normals"] = np.tile(np.array([[0, 0, 1]], np.float32), (3 * F, 1))
        s["objects"] = np.array([(0, 0)], base["objects"].dtype)
        s["vert_trans"] = np.eye(4, dtype=np.float32).reshape(1, 16)
        s["normal_trans"] = np.eye(4, dtype=np.float32).reshape(1, 16)
        return s

    one = [[[-1, -1, 0], [1, -1, 0], [0, 1, 0]]]
    stack = [[[-1, -1, -k], [1, -1, -k], [0, 1, -k]] for k in range(9)]              # distinct keys along z
    same = [[[-1, -1, 0], [1, -1, 0], [0, 1, 0]]] * 7                                  # 7 identical triangles: equal keys
    far = [[[300 + k, -1, 0], [302 + k, -1, 0], [301 + k, 1, 0]] for k in range(5)]    # beyond the reference's +-256 grid
    rays = np.array([[0, 0, 5, 0, 0, -1], [301.2, 0, 5, 0, 0, -1], [0, 0, -20, 0, 0, 1], [50, 50, 5, 0, 0, -1]], np.float32)
    for tris in (one, one * 2, stack, same, far, stack + same + far):
        s = make(tris)
        ds = tpt.DeviceScene(ctx, s)
        bvh = tpt.BVH(ds)
        bvh.construct()
        got = bvh.trace(rays)
        ref = orc.trace_rays(s["vertices"], s["indices"], rays, mode=0)
        assert (got["prim"] == ref["prim"]).all(), (got, ref)      # equal-t ties -> lowest primitive id on both sides
        assert np.allclose(got["t"][ref["prim"] >= 0], ref["t"][ref["prim"] >= 0], rtol=1e-6)
        # the renderer's three camera-ray forms agree on them too: per-thread candidates (< 32 spp), warp-cooperative
        # candidates (>= 32 spp), per-sample traversal; equal-t duplicates must resolve to the same primitive in all of them
        cam = tpt.Camera.from_scene(s)
        pt = tpt.PathTracer(ctx, 64, 36, env_rgba8=synthetic_sky(64, 32))
        imgs = []
        for spp_chunks, cand in (([40], True), ([40], False), ([8] * 5, True)):
            pt.m_radiance.clear()
            off = 0
            for n in spp_chunks:
                pt.render_into(ds, cam, spp=n, seed=3, sample_offset=off, pixel_candidates=cand)
                off += n
            imgs.append(pt.m_radiance.to_i64())
        assert (imgs[0] == imgs[1]).all() and (imgs[0] == imgs[2]).all() and imgs[0].any()
        ds.close()


def _moved(scene, angle, shift):
    """New per-object matrices: every object rotated about y by its own multiple of `angle` and shifted, on top of its
    glTF transform; normal matrix = inverse transpose of the upper 3x3 (src/mesh.cu:371-384)."""
    s = dict(scene)
    O = scene["vert_trans"].shape[0]
    vts, nts = [], []
    for o in range(O):
        M = scene["vert_trans"][o].reshape(4, 4).T.astype(np.float64)          # column-major -> math layout
        a = angle * (o + 1)
        R = np.array([[np.cos(a), 0, np.sin(a), shift * o], [0, 1, 0, -shift * (o % 3)], [-np.sin(a), 0, np.cos(a), 0.5 * shift], [0, 0, 0, 1]])
        M2 = R @ M
        N = np.eye(4)
        N[:3, :3] = np.linalg.inv(M2[:3, :3]).T
        vts.append(M2.T.reshape(16))
        nts.append(N.T.reshape(16))
    s["vert_trans"] = np.ascontiguousarray(np.stack(vts), np.float32)
    s["normal_trans"] = np.ascontiguousarray(np.stack(nts), np.float32)
    return s


@pytest.mark.parametrize("name", ["box", "box2", "ball"])
def test_refit_equals_rebuild(name, tpt, ctx, scenes, orc):
    """tpt_bvh_refit: new object matrices on the same mesh, hierarchy of the first build kept.  World
    vertices bit-exact against the oracle's transform of the moved scene, first hits bit-exact against brute force,
    node invariants hold, and a frame rendered on the refitted tree has the same accumulator bits as one rendered on a
    tree rebuilt from scratch (hits do not depend on the tree shape)."""
    ds = tpt.DeviceScene(ctx, scenes[name])
    bvh = tpt.BVH(ds)
    with pytest.raises(tpt.TptError):
        bvh.refit()                                   # nothing to reuse yet
    bvh.construct()
    cam = tpt.Camera.from_scene(scenes[name])
    W, H = 160, 90
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=synthetic_sky(128, 64))
    rng = np.random.default_rng(7)
    for step, (angle, shift) in enumerate([(0.05, 0.02), (0.4, 0.3), (0.0, 0.0)]):
        moved = _moved(scenes[name], angle, shift)
        ds.set_transforms(moved["vert_trans"], moved["normal_trans"])
        with pytest.raises(tpt.TptError):
            bvh.trace(np.zeros((4, 6), np.float32))   # stale boxes are never traced
        st = bvh.refit()
        assert st["n_faces"] == ds.n_faces
        wv, wn = ds.world()
        ov, on = orc.transform(moved)
        used = np.zeros(wv.shape[0], bool)
        used[moved["indices"]] = True
        assert np.array_equal(wv[used], ov[used]) and np.array_equal(wn[used], on[used])
        # structure: same children as before, boxes contain their triangles again
        nodes = bvh.m_nodes
        F = ds.n_faces
        tri = ov[moved["indices"].reshape(-1, 3)]
        leaves = nodes[F - 1:]
        face = leaves["a"]
        assert sorted(face.tolist()) == list(range(F))
        assert np.array_equal(leaves["mn"], tri[face].min(axis=1)) and np.array_equal(leaves["mx"], tri[face].max(axis=1))
        root = nodes[0]
        assert np.array_equal(root["mn"], tri.min(axis=(0, 1))) and np.array_equal(root["mx"], tri.max(axis=(0, 1)))
        # first hits: pixel-centre camera rays + incoherent rays, against brute force
        rays = orc.primary_rays(moved, W, H)
        lo, hi = ov[used].min(axis=0) - 0.5, ov[used].max(axis=0) + 0.5
        n = 20000
        o = rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)
        d = rng.normal(size=(n, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        rays = np.concatenate([rays, np.concatenate([o, d], axis=1)]).astype(np.float32)
        got = bvh.trace(rays)
        ref = orc.trace_rays(ov, moved["indices"], rays, mode=0)
        for k in ("prim", "t", "u", "v"):
            assert np.array_equal(got[k], ref[k]), (step, k)
        # a frame on the refitted tree == the same frame on a tree built from scratch for the moved scene
        pt.m_radiance.clear()
        pt.render_into(ds, cam, spp=16, seed=11)
        a = pt.m_radiance.to_i64()
        ds2 = tpt.DeviceScene(ctx, moved)
        tpt.BVH(ds2).construct()
        pt.m_radiance.clear()
        pt.render_into(ds2, cam, spp=16, seed=11)
        assert (pt.m_radiance.to_i64() == a).all() and a.any()
        ds2.close()
    assert ctx.counters()["queue_overflow"] == 0
    ds.close()


def test_live_pixel_batches_change_nothing(tpt, ctx, built, scenes):
    """A frame too big for one worst-case batch (1080p x 128 spp = 265 M camera samples > 2^27) is sized by its LIVE pixels
    (k_tile_live: tir.gltf has ~10 %) and runs as one batch; forcing the old worst-case batches gives the same accumulator
    bits, nothing overflows, and a changed camera re-measures."""
    ds, _ = built["tir"]
    W, H, spp = 1920, 1080, 128
    cam = tpt.Camera.from_scene(scenes["tir"])
    pt = tpt.PathTracer(ctx, W, H)
    n0 = ctx.counters()["kernel_launches"]
    pt.render_into(ds, cam, spp=spp, max_depth=32, seed=3)
    one = ctx.counters()["kernel_launches"] - n0
    a = pt.m_radiance.to_i64()
    pt.m_radiance.clear()
    n0 = ctx.counters()["kernel_launches"]
    pt.render_into(ds, cam, spp=spp, max_depth=32, seed=3, paths_in_flight=W * H * 32)
    four = ctx.counters()["kernel_launches"] - n0
    assert (pt.m_radiance.to_i64() == a).all() and a.any()
    # k_tile_live + k_primary + 32 waves against 4 x (k_primary + 32 waves)   (TPT_WAVE_LOOP=1: 4 against 12)
    assert one <= 35 and four >= 3 * one
    # another camera (closer: more live pixels) must not reuse the first camera's table
    sc = scenes["tir"]
    c2w = np.array(sc["c2w"], np.float32).reshape(16).copy()
    c2w[12:15] *= np.float32(0.6)                            # column-major: the translation column
    cam2 = tpt.Camera(sc["vfov"], sc["aspect"], sc["near"], c2w)
    pt.m_radiance.clear()
    pt.render_into(ds, cam2, spp=spp, max_depth=32, seed=3)
    b = pt.m_radiance.to_i64()
    pt.m_radiance.clear()
    pt.render_into(ds, cam2, spp=spp, max_depth=32, seed=3, paths_in_flight=W * H * 32)
    assert (pt.m_radiance.to_i64() == b).all() and not (a == b).all()
    assert ctx.counters()["queue_overflow"] == 0


@pytest.mark.parametrize("name,depth", [("light", 8), ("tir", 32), ("box", 8), ("box2", 8), ("box1", 8), ("ball", 8)])
def test_specular_continuation_is_bit_identical_to_queued_paths(name, depth, tpt, ctx, built, scenes):
    """k_primary follows mirror / glass bounces of a camera path in registers (while-while traversal, no queue round trip)
    until the first diffuse vertex, which is queued with the number of vertices it is ahead of its wave.  Same hits, same
    RNG dimensions, exact integer accumulation: the accumulator and the ray counts must equal those of the queued form,
    in both camera-ray forms (per-thread < 32 spp, warp-cooperative >= 32 spp)."""
    ds, _ = built[name]
    cam = tpt.Camera.from_scene(scenes[name])
    W, H = 200, 120
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=synthetic_sky(128, 64))
    for spp in (8, 40):
        res = []
        for cont in (True, False):
            pt.m_radiance.clear()
            ctx.reset_counters()
            pt.render_into(ds, cam, spp=spp, max_depth=depth, seed=13, specular_continuation=cont)
            c = ctx.counters()
            res.append((pt.m_radiance.to_i64(), c["rays_extension"], c["rays_shadow"], c["rays_direct"], c["kernel_launches"]))
            assert c["queue_overflow"] == 0
        assert (res[0][0] == res[1][0]).all() and res[0][0].any()
        assert res[0][1:4] == res[1][1:4], (res[0][1:], res[1][1:])


def test_specular_continuation_with_delta_lights(tpt, ctx, scenes, orc):
    """No fixture has both a delta light and a mirror / glass surface; box.gltf plus a point light has.  The shadow rays of the
    vertices passed in registers are traced there too: accumulator and counters equal the queued form, and the image agrees
    with the oracle (shared RNG)."""
    s = dict(scenes["box"])
    L = np.zeros(1, s["lights"].dtype)
    L["type"] = 0                                     # point light inside the box
    L["color"] = (1.0, 0.9, 0.8)
    L["intensity"] = 0.6
    wv, _ = orc.transform(s)
    L["a"] = (wv.min(axis=0) + wv.max(axis=0)) * 0.5 + np.array([0.1, 0.2, 0.05], np.float32)
    s["lights"] = L
    ds = tpt.DeviceScene(ctx, s)
    tpt.BVH(ds).construct()
    cam = tpt.Camera.from_scene(s)
    W, H, spp = 96, 54, 32
    pt = tpt.PathTracer(ctx, W, H)
    out = []
    for cont in (True, False):
        pt.m_radiance.clear()
        ctx.reset_counters()
        pt.render_into(ds, cam, spp=spp, seed=17, specular_continuation=cont)
        c = ctx.counters()
        out.append((pt.m_radiance.to_i64(), c["rays_extension"], c["rays_shadow"], c["rays_direct"]))
    assert (out[0][0] == out[1][0]).all() and out[0][1:] == out[1][1:] and out[0][2] > 0
    img = pt.m_radiance.to_f32() / spp
    ref = orc.render(s, W, H, spp, seed=17) / spp
    assert _relmse(img, ref) <= 2e-2
    ds.close()


def test_host_pointer_trace_is_pipelined_in_chunks_with_the_same_hits(built, scenes, orc):
    """tpt_trace_rays with host buffers cuts batches above 2^20 rays into chunks (copy-in, traversal and copy-out of
    neighbouring chunks overlap on three streams); the hits must be those of the same rays traced in single-chunk calls."""
    ds, bvh = built["box"]
    wv, _ = orc.transform(scenes["box"])
    rng = np.random.default_rng(99)
    n = 2_600_000                                            # 3 chunks, the last one ragged
    lo, hi = wv.min(axis=0) - 0.5, wv.max(axis=0) + 0.5
    o = rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d], axis=1).astype(np.float32)
    whole = bvh.trace(rays)
    parts = np.concatenate([bvh.trace(rays[i:i + 650_000]) for i in range(0, n, 650_000)])
    for k in ("prim", "t", "u", "v"):
        assert np.array_equal(whole[k], parts[k]), k
    ref = orc.trace_rays(wv, scenes["box"]["indices"], rays[:20000], mode=0)
    assert np.array_equal(whole["prim"][:20000], ref["prim"]) and np.array_equal(whole["t"][:20000], ref["t"])
    assert (whole["prim"] >= 0).mean() > 0.2


def test_do_trace_with_refit_flag_equals_rebuild(tpt, ctx, scenes):
    """TPT_RENDER_REFIT: tpt_do_trace builds on the first frame and refits afterwards; frames equal those of per-frame rebuilds,
    also after the objects have moved."""
    s = scenes["box2"]
    W, H = 96, 54
    cam = tpt.Camera.from_scene(s)
    a, b = tpt.DeviceScene(ctx, s), tpt.DeviceScene(ctx, s)
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=synthetic_sky(64, 32))
    for step, (angle, shift) in enumerate([(0.0, 0.0), (0.2, 0.1), (0.5, 0.25)]):
        moved = _moved(s, angle, shift)
        fa, fb = np.zeros((H, W, 4), np.uint8), np.zeros((H, W, 4), np.uint8)
        a.set_transforms(moved["vert_trans"], moved["normal_trans"])
        b.set_transforms(moved["vert_trans"], moved["normal_trans"])
        pt.doTrace(a, cam, fa, 8, seed=4, refit=True)      # frame 0 builds (no hierarchy yet), later frames refit
        pt.doTrace(b, cam, fb, 8, seed=4)
        assert (fa == fb).all() and fa[..., :3].any(), step
    a.close()
    b.close()
