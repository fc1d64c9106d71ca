"""CPU tests of the oracle: hand-derived known answers for every function on the path, so the
checker itself is checked before any GPU number is trusted.  Reference citations are in oracle/oracle.cpp."""
import math

import numpy as np
import pytest


def test_frsqrt_one_newton_step(orc):
    # Quake rsqrt with ONE Newton step: relative error up to ~0.175 %, never exact (vec.h:25-38)
    xs = np.array([1e-6, 0.25, 1.0, 2.0, 3.0, 7.5, 100.0, 12345.0], np.float32)
    rel = np.array([abs(orc.frsqrt(float(x)) * math.sqrt(float(x)) - 1.0) for x in xs])
    assert rel.max() < 1.76e-3
    assert rel.max() > 1e-4                      # i.e. it is NOT an exact rsqrt: t-parity depends on reproducing it
    assert orc.frsqrt(1.0) == pytest.approx(0.99830717, abs=1e-7)
    assert orc.frsqrt(4.0) == pytest.approx(0.49915358, abs=1e-7)


def test_morton_known_values(orc):
    # bvh.cu:23-46: fixed point x*2^23 for |x|<256, bias INT32_MAX, top 21 bits; x lowest, z highest (bvh.cu:60)
    def q(x):   # 21-bit cell of one coordinate
        return orc.morton(x, 0.0, 0.0) & 0x1249249249249249

    def spread(v):
        out = 0
        for b in range(21):
            out |= ((v >> b) & 1) << (3 * b)
        return out

    zero_cell = (0x7FFFFFFF & 0xFFFFF800) >> 11           # value 0 + INT32_MAX
    assert q(0.0) == spread(zero_cell)
    # +1.0 -> 2^23 + INT32_MAX overflows int32 on the GPU (wraps): bits = 0x807FFFFF
    assert q(1.0) == spread((0x807FFFFF & 0xFFFFF800) >> 11)
    assert q(-1.0) == spread(((0x7FFFFFFF - (1 << 23)) & 0xFFFFF800) >> 11)
    # grid cell is 2^-12: two points in the same cell share a key, neighbours differ
    # (the INT32_MAX bias puts cell boundaries at x*2^23 == 1 mod 2^11, so start 2 units above 0.5)
    x1 = 0.5 + 2.0 ** -22
    assert orc.morton(x1, 0.5, 0.5) == orc.morton(x1 + 2.0 ** -14, 0.5, 0.5)
    assert orc.morton(x1, 0.5, 0.5) != orc.morton(x1 + 2.0 ** -11, 0.5, 0.5)
    assert orc.morton(0.5, 0.5, 0.5) != orc.morton(x1, 0.5, 0.5)
    # saturation at |x| >= 256 (exponent >= 8): every far coordinate collapses to one cell
    assert orc.morton(256.0, 0, 0) == orc.morton(1e6, 0, 0)
    assert orc.morton(-256.0, 0, 0) == orc.morton(-1e6, 0, 0)
    # axis interleave: y is bit 1, z is bit 2
    assert orc.morton(0, 1.0, 0) & 0x1249249249249249 == spread(zero_cell)
    assert (orc.morton(0, 1.0, 0) >> 1) & 0x1249249249249249 == spread((0x807FFFFF & 0xFFFFF800) >> 11)
    assert (orc.morton(0, 0, 1.0) >> 2) & 0x1249249249249249 == spread((0x807FFFFF & 0xFFFFF800) >> 11)
    assert orc.morton(3.0, -7.0, 11.0) < 2 ** 63


def test_ray_triangle_edges(orc):
    tri = [0, 0, 0, 1, 0, 0, 0, 1, 0]
    ok, o = orc.ray_hit_triangle([0.25, 0.25, 1, 0, 0, -1], tri)
    assert ok and o[0] == pytest.approx(1.0) and o[1] == pytest.approx(0.25) and o[2] == pytest.approx(0.25)
    # two-sided: hit from behind too (no culling, geometry_queries.h:66-86)
    ok, o = orc.ray_hit_triangle([0.25, 0.25, -1, 0, 0, 1], tri)
    assert ok and o[0] == pytest.approx(1.0)
    # exactly on an edge / a vertex is inside (u<0 || v<0 || u+v>1 rejects only strict outside)
    assert orc.ray_hit_triangle([0.5, 0.0, 1, 0, 0, -1], tri)[0]
    assert orc.ray_hit_triangle([0.0, 0.0, 1, 0, 0, -1], tri)[0]
    assert orc.ray_hit_triangle([0.5, 0.5, 1, 0, 0, -1], tri)[0]
    assert not orc.ray_hit_triangle([0.51, 0.5, 1, 0, 0, -1], tri)[0]
    # parallel ray: denom == 0 exactly -> miss
    assert not orc.ray_hit_triangle([0.25, 0.25, 1, 1, 0, 0], tri)[0]
    # a hit behind the origin is reported with negative distance (the t > 2e-4 filter is the caller's)
    ok, o = orc.ray_hit_triangle([0.25, 0.25, 1, 0, 0, 1], tri)
    assert ok and o[0] == pytest.approx(-1.0)
    # non-unit direction: distance is in units of the direction's length
    ok, o = orc.ray_hit_triangle([0.25, 0.25, 1, 0, 0, -2], tri)
    assert ok and o[0] == pytest.approx(0.5)


def test_ray_bbox_is_a_line_test(orc):
    mn, mx = [0, 0, 0], [1, 1, 1]
    assert orc.ray_hit_bbox([0.5, 0.5, 5, 0, 0, -1], mn, mx)
    assert orc.ray_hit_bbox([0.5, 0.5, 5, 0, 0, 1], mn, mx)        # box entirely BEHIND the origin still passes
    assert not orc.ray_hit_bbox([2.5, 0.5, 5, 0, 0, -1], mn, mx)
    # axis-parallel ray on a face plane: 0*inf = NaN slabs; fmaxf/fminf ignore the NaN, the box passes
    assert orc.ray_hit_bbox([0.0, 0.5, 5, 0, 0, -1], mn, mx)
    # flat box (axis-aligned quad)
    assert orc.ray_hit_bbox([0.5, 5, 0.5, 0, -1, 0], [0, 0, 0], [1, 0, 1])


def test_closest_hit_rule(orc):
    # t <= 2e-4 is rejected (self-intersection guard, path_tracer.cu:83); ties keep the lowest id in brute force
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, -1], [1, 0, -1], [0, 1, -1]], np.float32)
    idx = np.arange(9, dtype=np.uint32)
    rays = np.array([[0.2, 0.2, 1, 0, 0, -1], [0.2, 0.2, 1e-4, 0, 0, -1], [0.2, 0.2, 3e-4, 0, 0, -1]], np.float32)
    h = orc.trace_rays(v, idx, rays, mode=0)
    assert h["prim"][0] == 0 and h["t"][0] == pytest.approx(1.0)       # coincident triangles 0 and 1: lowest id
    assert h["prim"][1] == 2 and h["t"][1] == pytest.approx(1.0001, rel=1e-5)   # t=1e-4 hits rejected, next surface taken
    assert h["prim"][2] == 0 and h["t"][2] == pytest.approx(3e-4, rel=1e-3)


def test_refract_schlick(orc):
    n = [0, 0, 1]
    # normal incidence from outside (dir . n < 0): eta = 1/ior, straight through
    d, cos_i, eta, tir = orc.refract([0, 0, -1], n, 2.0)
    assert not tir and eta == pytest.approx(0.5) and cos_i == pytest.approx(1.0)
    np.testing.assert_allclose(d, [0, 0, -1], atol=1e-6)
    assert orc.schlick(1.0, 0.5) == pytest.approx(((1 - 0.5) / (1 + 0.5)) ** 2)
    # leaving the glass (dir . n > 0): eta = ior = 2, critical angle 30 degrees
    s29, s31 = math.sin(math.radians(29.0)), math.sin(math.radians(31.0))
    _, _, eta, tir = orc.refract([s29, 0, math.cos(math.radians(29.0))], n, 2.0)
    assert eta == pytest.approx(2.0) and not tir
    _, _, _, tir = orc.refract([s31, 0, math.cos(math.radians(31.0))], n, 2.0)
    assert tir
    # Snell: sin(t) = eta * sin(i)
    d, _, _, _ = orc.refract([math.sin(0.3), 0, -math.cos(0.3)], n, 1.5)
    assert d[0] / np.linalg.norm(d) == pytest.approx(math.sin(0.3) / 1.5, rel=1e-5)
    # grazing: Schlick -> 1
    assert orc.schlick(0.0, 0.5) == pytest.approx(1.0)
    np.testing.assert_allclose(orc.reflect([1, 0, -1], n), [1, 0, 1], atol=1e-7)


def test_hemisphere_cosine(orc):
    n = np.array([0.0, 0.0, 1.0], np.float32)
    d = orc.hemisphere_cosine(0.25, 1.0, n)              # cos(theta) = sqrt(1) = 1 -> along the normal
    np.testing.assert_allclose(d, n, atol=1e-6)
    d = orc.hemisphere_cosine(0.25, 0.25, n)             # cos(theta) = 0.5
    assert float(d @ n) == pytest.approx(0.5, abs=1e-6)
    # n.z == 0 branch uses xBase = (0,0,1)
    d = orc.hemisphere_cosine(1.0, 0.25, [1, 0, 0])
    assert float(d[0]) == pytest.approx(0.5, abs=1e-6)
    # distribution: E[cos] = 2/3 for a cosine-weighted hemisphere
    rng = np.random.default_rng(0)
    cs = [float(orc.hemisphere_cosine(float(a), float(b), [0.3, -0.5, 0.81]) @ np.array([0.3, -0.5, 0.81])) for a, b in rng.uniform(1e-6, 1, (4000, 2))]
    assert np.mean(cs) / np.linalg.norm([0.3, -0.5, 0.81]) ** 2 == pytest.approx(2 / 3, abs=0.02)


def _light(orc, typ, **kw):
    rec = np.zeros((), orc.LIGHT_DTYPE)
    rec["type"] = typ
    rec["color"] = (1, 1, 1)
    rec["intensity"] = 1
    for k, v in kw.items():
        rec[k] = v
    return rec


def test_delta_light_falloff(orc):
    # CalcDistAttenuation (delta_light.h:25-33): 1/(d^2+1) * saturate(1-(d^2*0.01)^2)^2 -> 0 at d >= 10
    L = _light(orc, 0, a=(0, 0, 0), intensity=3.0)
    rad, direction, dist = orc.light_sample(L, [0, 0, 0.0])     # d = 0: direction is NaN (0/0), radiance 3
    assert dist == 0.0 and rad[0] == pytest.approx(3.0)
    rad, direction, dist = orc.light_sample(L, [0, 0, -1.0])
    assert dist == pytest.approx(1.0) and rad[0] == pytest.approx(3.0 * 0.5 * (1 - 1e-4) ** 2, rel=1e-6)
    np.testing.assert_allclose(direction, [0, 0, 1], atol=1e-6)
    rad, _, _ = orc.light_sample(L, [0, 0, -10.0])
    assert rad[0] == 0.0
    rad, _, _ = orc.light_sample(L, [0, 0, -25.0])
    assert rad[0] == 0.0
    # directional: distance 0 -> no attenuation, direction = -light direction (delta_light.h:56-63)
    D = _light(orc, 1, a=(0, -1, 0), intensity=2.0)
    rad, direction, dist = orc.light_sample(D, [5, 5, 5])
    assert dist == 0.0 and rad[1] == pytest.approx(2.0)
    np.testing.assert_allclose(direction, [0, 1, 0], atol=1e-7)
    # spot: square(saturate(cos - cosOuter) * invDiff) (delta_light.h:73-85); on-axis, inner 0.9/outer 0.8
    S = _light(orc, 2, a=(0, 1, 0), b=(0, -1, 0), cos_outer=0.8, inv_cone_diff=1.0 / (0.9 - 0.8))
    rad, _, dist = orc.light_sample(S, [0, 0, 0])
    assert rad[0] == pytest.approx(((1.0 - 0.8) * 10.0) ** 2 * 0.5 * (1 - 1e-4) ** 2, rel=1e-5)   # not clamped to 1 above the inner cone
    rad, _, _ = orc.light_sample(S, [5, 0, 0])                    # far outside the cone
    assert rad[0] == 0.0


def test_vec2uv_and_env_lookup(orc):
    # env_light.cuh:72-78: u = atan2(z,x)/2pi (+1 if <0), v = 1 - acos(y)/pi
    np.testing.assert_allclose(orc.vec2uv([1, 0, 0]), [0.0, 0.5], atol=1e-6)
    np.testing.assert_allclose(orc.vec2uv([0, 0, 1]), [0.25, 0.5], atol=1e-6)
    np.testing.assert_allclose(orc.vec2uv([-1, 0, 1e-9]), [0.5, 0.5], atol=1e-6)
    np.testing.assert_allclose(orc.vec2uv([0, 0, -1]), [0.75, 0.5], atol=1e-6)
    assert orc.vec2uv([0, 1, 0])[1] == pytest.approx(1.0)
    assert orc.vec2uv([0, -1, 0])[1] == pytest.approx(0.0)
    assert orc.vec2uv([0, 1.7, 0])[1] == pytest.approx(1.0)      # clamp of |d.y| > 1 (non-unit directions)
    img = np.zeros((4, 8, 4), np.uint8)
    img[3, :, 0] = 255      # top row (row 0 = bottom scanline) red
    img[0, :, 2] = 255      # bottom row blue
    img[1, 2, 1] = 51       # v in [0.25,0.5), u in [0.25,0.375)
    np.testing.assert_allclose(orc.sample_env(img, [0, 1, 0]), [1, 0, 0])       # v = 1 -> clamps to the last row
    np.testing.assert_allclose(orc.sample_env(img, [0, -1, 0]), [0, 0, 1])
    d = [math.cos(2 * math.pi * 0.3) * math.sin(math.pi * 0.6), math.cos(math.pi * 0.6), math.sin(2 * math.pi * 0.3) * math.sin(math.pi * 0.6)]
    np.testing.assert_allclose(orc.sample_env(img, d), [0, 51 / 255.0, 0], atol=1e-7)


def test_resolve_truncates_flips_and_keeps_alpha(orc):
    rad = np.zeros((2, 2, 3), np.float32)
    rad[0, 0] = (2 * 0.999, 2 * 0.5, 2 * 2.0)     # bottom-left, spp = 2
    rad[1, 1] = (-1.0, float("nan"), 2 * 0.0039)
    fb = orc.resolve_bgra8(rad, 2, np.full((2, 2, 4), 9, np.uint8))
    assert tuple(fb[1, 0]) == (255, 127, 254, 9)        # row flipped; B,G,R; 0.999*255 = 254.7 -> 254 (truncation)
    assert tuple(fb[0, 1]) == (0, 255, 0, 9)            # NaN -> 255 through the reference's min/max templates; <0 -> 0


def test_rng_spec(orc):
    # golden values of the counter-based generator (DESIGN.md "RNG spec"); the CUDA side must reproduce them
    a = orc.rng_pair(1, 0, 0, 0)
    b = orc.rng_pair(1, 0, 0, 1)
    c = orc.rng_pair(1, 1, 0, 0)
    d = orc.rng_pair(2, 0, 0, 0)
    vals = np.stack([a, b, c, d])
    assert ((vals > 0) & (vals <= 1)).all()
    assert len({tuple(v) for v in vals.tolist()}) == 4
    # pinned numbers (computed once from the spec with Python integers, see test below)
    M = (1 << 64) - 1

    def mix(z):
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        return z ^ (z >> 31)

    def pair(seed, pixel, sample, j):
        G = 0x9E3779B97F4A7C15
        key = mix(mix((seed + G) & M) ^ ((pixel << 32) | sample))
        z = key if j == 0 else mix((key + j * G) & M)
        u = lambda x: np.float32(((x >> 8) + 1)) * np.float32(2.0 ** -24)
        return u(z >> 32), u(z & 0xFFFFFFFF)

    for args in [(1, 0, 0, 0), (1, 0, 0, 1), (1, 1, 0, 0), (2, 0, 0, 0), (0xDEADBEEF, 2073599, 4095, 17)]:
        np.testing.assert_array_equal(orc.rng_pair(*args), np.array(pair(*args), np.float32))
    # uniformity smoke: mean of 20000 draws
    xs = np.array([orc.rng_pair(7, p, 3, 2) for p in range(10000)])
    assert xs.mean() == pytest.approx(0.5, abs=0.01)
    assert np.corrcoef(xs[:, 0], xs[:, 1])[0, 1] == pytest.approx(0.0, abs=0.03)


def test_trs_and_normal_matrix(orc):
    m = orc.trs_to_mat4([1, 2, 3], [0, 0, 0, 0], [1, 1, 1]).reshape(4, 4)    # Quat() default (all zero) is identity rotation
    np.testing.assert_allclose(m.T, [[1, 0, 0, 1], [0, 1, 0, 2], [0, 0, 1, 3], [0, 0, 0, 1]])
    s = math.sqrt(0.5)
    m = orc.trs_to_mat4([0, 0, 0], [0, 0, s, s], [2, 1, 1])                   # 90 deg about z, scale x by 2
    p = orc.mat4_mul_point(m, [1, 0, 0], 1.0)
    np.testing.assert_allclose(p, [0, 2, 0], atol=1e-6)
    nm = orc.normal_matrix(m).reshape(4, 4).T
    want = np.linalg.inv(m.reshape(4, 4).T[:3, :3]).T
    np.testing.assert_allclose(nm[:3, :3], want, atol=1e-6)


def test_reference_bvh_restated(orc, scenes):
    """Restated hierarchy (bvh.cu:150-217) is a valid tree on the bundled scenes (duplicate key PAIRS only), and
    reference-order traversal equals brute force there."""
    for name, s in scenes.items():
        wv, _ = orc.transform(s)
        nodes, keys, rc = orc.build_reference_bvh(wv, s["indices"])
        assert rc == 0
        assert (np.diff(keys) >= 0).all()
        F = s["indices"].size // 3
        seen = np.zeros(2 * F - 1, int)
        for i in range(F - 1):
            seen[nodes["a"][i]] += 1
            seen[nodes["b"][i]] += 1
        assert seen[0] == 0 and (seen[1:] == 1).all(), name
        rays = orc.primary_rays(s, 96, 54)
        a = orc.trace_rays(wv, s["indices"], rays, mode=0)
        b = orc.trace_rays(wv, s["indices"], rays, nodes=nodes, mode=1)
        assert (a["prim"] == b["prim"]).mean() > 0.999
        same = a["prim"] == b["prim"]
        assert (a["t"][same] == b["t"][same]).all()


def test_reference_bvh_breaks_on_triple_keys(orc):
    """>= 3 equal Morton keys make the reference's hierarchy malformed (no index tie-break)."""
    tri = np.array([[0, 0, 0], [1e-5, 0, 0], [0, 1e-5, 0]], np.float32)
    verts = np.concatenate([tri + np.float32(k) * np.float32(1e-6) for k in range(4)] + [tri + 3.0, tri - 3.0]).astype(np.float32)
    idx = np.arange(verts.shape[0], dtype=np.uint32)
    nodes, keys, rc = orc.build_reference_bvh(verts, idx)
    F = idx.size // 3
    seen = np.zeros(2 * F - 1, int)
    for i in range(F - 1):
        seen[nodes["a"][i]] += 1
        seen[nodes["b"][i]] += 1
    assert rc == 1 or not (seen[1:] == 1).all()


def test_estimator_closed_form_furnace(orc, scenes):
    """Closed box idea in the open: a single diffuse quad under a constant white environment.  Every camera path
    hits the quad once (baseColor c) and then escapes: L = c * env exactly, per sample, since atten/pdf == c."""
    s = dict(scenes["tir"])
    quad = np.array([[-50, 0, -50], [50, 0, -50], [50, 0, 50], [-50, 0, 50]], np.float32)
    s["vertices"] = quad
    s["normals"] = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    s["indices"] = np.array([0, 2, 1, 0, 3, 2], np.uint32)
    s["objects"] = np.array([(0, 0)], s["objects"].dtype)
    mats = s["materials"][:1].copy()
    mats["base"] = (0.5, 0.25, 1.0)
    mats["emission"] = 0
    mats["eta"] = 0
    mats["metallic"] = 0
    s["materials"] = mats
    s["vert_trans"] = np.eye(4, dtype=np.float32).reshape(1, 16)
    s["normal_trans"] = np.eye(4, dtype=np.float32).reshape(1, 16)
    s["lights"] = s["lights"][:0]
    s["c2w"] = orc.trs_to_mat4([0, 1, 0], [-math.sin(0.6), 0, 0, math.cos(0.6)], [1, 1, 1])   # looking down at the plane
    env = np.full((4, 8, 4), 255, np.uint8)
    img = orc.render(s, 16, 9, 4, max_depth=8, seed=1, env=env) / 4.0
    np.testing.assert_allclose(img, np.broadcast_to(np.array([0.5, 0.25, 1.0], np.float32), img.shape), rtol=2e-6)


@pytest.mark.parametrize("name", ["box", "ball", "tir"])
def test_blocked_vector_brute_force_equals_scalar_brute_force(name, orc, scenes):
    """Oracle mode 2 (brute_force_blocked: the authority for the 10 M-triangle soup, tests/test_gpu_soup_full.py) must be the
    scalar brute force (mode 0) bit for bit — ids with the lowest-id tie rule, t, u, v — including rays through shared edges
    and vertices (pixel-centre primary rays hit the quads' diagonals) and rays that miss."""
    s = scenes[name]
    wv, _ = orc.transform(s)
    rng = np.random.default_rng(9)
    prim = orc.primary_rays(s, 96, 54)
    lo, hi = wv.min(0), wv.max(0)
    o = (lo + (hi - lo) * rng.random((3000, 3))).astype(np.float32)
    d = rng.normal(size=(3000, 3)).astype(np.float32)
    # rays aimed exactly at vertices: exact-t ties between the triangles that share them
    v = wv[rng.integers(0, wv.shape[0], 500)]
    aimed = np.concatenate([np.broadcast_to(prim[0, :3], (500, 3)), v - prim[0, :3]], axis=1).astype(np.float32)
    rays = np.concatenate([prim, np.concatenate([o, d], axis=1), aimed])
    a = orc.trace_rays(wv, s["indices"], rays, mode=0)
    b = orc.trace_rays(wv, s["indices"], rays, mode=2)
    assert (a["prim"] >= 0).any() and (a["prim"] < 0).any()
    assert a.tobytes() == b.tobytes()
