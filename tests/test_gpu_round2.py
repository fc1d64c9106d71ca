"""Round-2 additions to the render path: the cooperative multi-level wave kernel (k_wave_loop), the warp-cooperative tile walk
in front of the per-pixel candidate lists (tile_collect) and tpt_scene_update.  The first two are pure re-schedulings of the
same integer-accumulated terms, so the bar is bit-identical accumulation buffers against the per-level / per-pixel forms
(switched back on with the tuning variables TPT_NO_WAVE_LOOP / TPT_NO_TILE_WALK, which the library reads once per process —
hence the subprocesses).  k_wave_loop and path_tail are opt-in (TPT_WAVE_LOOP, TPT_PATH_TAIL_MAX): measured, they do not pay
(DESIGN.md §7), but they stay correct."""
import hashlib
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]

_SNIPPET = r"""
import sys, hashlib, json
sys.path.insert(0, {root!r})
import numpy as np
import tinypathtracer_b200 as tpt
from tinypathtracer_b200.scene_io import load_scene_npz
from tinypathtracer_b200.sky import synthetic_sky
out = {{}}
ctx = tpt.Context(0)
for name, W, H, spp, depth, env, pif in {cases!r}:
    scene = load_scene_npz({root!r} + "/tests/golden/scenes/" + name + ".npz")
    ds = tpt.DeviceScene(ctx, scene)
    tpt.BVH(ds).construct(stats=False)
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=synthetic_sky(256, 128) if env else None)
    pt.render_into(ds, tpt.Camera.from_scene(scene), spp=spp, max_depth=depth, seed=11, paths_in_flight=pif)
    out[name + str(depth) + str(pif)] = hashlib.sha256(pt.m_radiance.to_i64().tobytes()).hexdigest()
    c = ctx.counters()
    assert c["queue_overflow"] == 0
    ds.close()
print(json.dumps(out))
"""

CASES = [("box", 96, 54, 64, 8, False, 0), ("ball", 160, 90, 64, 8, True, 0), ("tir", 96, 54, 64, 32, False, 0), ("light", 96, 54, 32, 8, True, 0),
         ("square", 64, 36, 32, 8, False, 0), ("box2", 64, 36, 16, 2, True, 0), ("box", 64, 36, 8, 1, False, 0),
         ("box", 96, 54, 64, 8, False, 20000)]      # several batches (form A and B of k_primary)


def _run(extra_env):
    import json
    env = dict(os.environ)
    env.update(extra_env)
    r = subprocess.run([sys.executable, "-c", _SNIPPET.format(root=str(ROOT), cases=CASES)], capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


def test_wave_loop_and_tile_walk_keep_every_accumulator_bit():
    base = _run({})                                   # one launch per wave, tile walk on
    assert _run({"TPT_NO_TILE_WALK": "1"}) == base
    loop = _run({"TPT_WAVE_LOOP": "1"})               # waves 2.. in one cooperative launch
    assert loop == base
    # ... with the path-by-path finish of small levels (path_tail) forced for every level after wave 1
    assert _run({"TPT_WAVE_LOOP": "1", "TPT_PATH_TAIL_MAX": "4000000000"}) == base


def test_scene_update_equals_a_new_scene(tpt, ctx, scenes, orc):
    """tpt_scene_update re-uploads into the existing allocations: the rebuilt tree and its hits are those of a scene created
    from the same arrays; sizes must match; a refit is refused until the scene has been rebuilt."""
    base = scenes["box"]
    moved = dict(base)
    moved["vertices"] = (np.asarray(base["vertices"], np.float32) * np.float32(1.25) + np.float32(0.125)).astype(np.float32)
    mats = np.array(base["materials"], copy=True)
    mats["base"][0] = (0.25, 0.5, 0.75)
    moved["materials"] = mats
    ds = tpt.DeviceScene(ctx, base)
    bvh = tpt.BVH(ds)
    bvh.construct()
    ds.update(moved)
    with pytest.raises(tpt.TptError):
        bvh.refit()                                   # no hierarchy for the new geometry yet
    rays = orc.primary_rays(moved, 160, 90)
    with pytest.raises(tpt.TptError):
        bvh.trace(rays)                               # not built
    bvh.construct()
    fresh = tpt.DeviceScene(ctx, moved)
    fb = tpt.BVH(fresh)
    fb.construct()
    a, b = bvh.trace(rays), fb.trace(rays)
    assert a.tobytes() == b.tobytes() and (a["prim"] >= 0).any()
    wv, _ = orc.transform(moved)
    assert np.array_equal(ds.world()[0], wv)
    ref = orc.trace_rays(wv, moved["indices"], rays, mode=0)
    assert a.tobytes() == ref.tobytes()
    # the material table went up too: same image as the fresh scene, bit for bit
    cam = tpt.Camera.from_scene(moved)
    imgs = []
    for d in (ds, fresh):
        pt = tpt.PathTracer(ctx, 64, 36)
        pt.render_into(d, cam, spp=8, max_depth=4, seed=3)
        imgs.append(pt.m_radiance.to_i64())
    assert (imgs[0] == imgs[1]).all() and imgs[0].any()
    with pytest.raises(tpt.TptError, match="sizes differ"):
        ds.update(scenes["ball"])
    ds.close()
    fresh.close()


_TRACE_SNIPPET = r"""
import sys, hashlib, json
sys.path.insert(0, {root!r})
import numpy as np
import tinypathtracer_b200 as tpt
from tinypathtracer_b200.scene_io import load_scene_npz
from tinypathtracer_b200.soup import make_soup
out = {{}}
ctx = tpt.Context(0)
ball = load_scene_npz({root!r} + "/tests/golden/scenes/ball.npz")
for name in ["box", "box1", "box2", "tir", "ball", "light", "square", "soup"]:
    scene = make_soup(ball, 64, 25) if name == "soup" else load_scene_npz({root!r} + "/tests/golden/scenes/" + name + ".npz")
    ds = tpt.DeviceScene(ctx, scene)
    bvh = tpt.BVH(ds)
    bvh.construct(stats=False)
    wv, _ = ds.world()
    lo, hi = wv.min(0), wv.max(0)
    rng = np.random.default_rng(5)
    n = 200000
    o = (lo - 0.25 * (hi - lo) + 1.5 * (hi - lo) * rng.random((n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    rays = np.concatenate([o, d], axis=1)
    prim = ctx.primary_rays(tpt.Camera.from_scene(scene), 320, 180)
    rays = np.concatenate([rays, np.concatenate([prim["origin"], prim["dir"]], axis=1)])
    ctx.reset_counters()
    h = bvh.trace(rays, counted=True)
    c = ctx.counters()
    a = bvh.trace(rays, any_hit=True)
    out[name] = [hashlib.sha256(h.tobytes()).hexdigest(), hashlib.sha256((a["prim"] >= 0).tobytes()).hexdigest(), int((h["prim"] >= 0).sum()),
                 c["node_visits"], c["tri_tests"]]
    ds.close()
print(json.dumps(out))
"""


def test_wide_and_binary_walks_agree(capsys):
    """tpt_trace_rays over the 4-wide compressed nodes (k_collapse4 + trav_node_step4) and over the binary nodes: the same
    (prim, t, u, v) bits for 257 600 rays per scene (incoherent + primary) on the seven fixtures and a 78 124-triangle soup, the
    same any-hit answers, and fewer node visits."""
    import json

    def run(extra):
        env = dict(os.environ)
        env["TPT_WIDE_MIN_FACES"] = "2"
        env.update(extra)
        r = subprocess.run([sys.executable, "-c", _TRACE_SNIPPET.format(root=str(ROOT))], capture_output=True, text=True, env=env, timeout=900)
        assert r.returncode == 0, r.stderr[-3000:]
        return json.loads(r.stdout.strip().splitlines()[-1])

    wide, binary = run({}), run({"TPT_NO_WIDE": "1"})
    lines = []
    for name in wide:
        assert wide[name][:3] == binary[name][:3], name
        assert wide[name][2] > 0
        lines.append(f"{name}: node visits {binary[name][3]} -> {wide[name][3]}, triangle tests {binary[name][4]} -> {wide[name][4]}")
    assert wide["soup"][3] < 0.7 * binary["soup"][3]
    with capsys.disabled():
        print("\n[wide vs binary] " + "; ".join(lines))


_BUILD_SNIPPET = r"""
import sys, hashlib, json
sys.path.insert(0, {root!r})
import numpy as np
import tinypathtracer_b200 as tpt
from tinypathtracer_b200.scene_io import load_scene_npz
from tinypathtracer_b200.soup import make_soup
out = {{}}
ctx = tpt.Context(0)
ball = load_scene_npz({root!r} + "/tests/golden/scenes/ball.npz")
for name in ["box", "box1", "box2", "tir", "ball", "light", "square", "soup_small", "soup"]:
    scene = make_soup(ball, 64, 25) if name == "soup" else make_soup(ball, 3, 40) if name == "soup_small" else \
        load_scene_npz({root!r} + "/tests/golden/scenes/" + name + ".npz")
    ds = tpt.DeviceScene(ctx, scene)
    bvh = tpt.BVH(ds)
    st = bvh.construct(stats=True)
    nodes = bvh.m_nodes                        # the reference's node layout: boxes, children, parents, leaf face ids
    keys, faces = bvh.order()
    # a refit on the same topology reads the parents / range ends the build left behind
    bvh.refit()
    again = bvh.m_nodes
    out[name] = [hashlib.sha256(nodes.tobytes()).hexdigest(), hashlib.sha256(faces.tobytes()).hexdigest(), hashlib.sha256(again.tobytes()).hexdigest(),
                 int(st["n_faces"]), int(st["max_depth"])]
    ds.close()
print(json.dumps(out))
"""


def test_bottom_up_hierarchy_is_the_search_based_tree():
    """A full build emits the hierarchy in the bottom-up walk (k_leaves_build / k_build_upper); TPT_BUILD_KARRAS=1 runs the
    search-based k_hierarchy + refit walk instead.  Same nodes (boxes, children, parents, leaf order) byte for byte on every
    fixture, a 3 672-triangle and a 78 124-triangle soup (many blocks, many equal keys), and a refit over either gives the same."""
    import json

    def run(extra):
        env = dict(os.environ)
        env.update(extra)
        r = subprocess.run([sys.executable, "-c", _BUILD_SNIPPET.format(root=str(ROOT))], capture_output=True, text=True, env=env, timeout=900)
        assert r.returncode == 0, r.stderr[-3000:]
        return json.loads(r.stdout.strip().splitlines()[-1])

    walk, search = run({}), run({"TPT_BUILD_KARRAS": "1"})
    assert walk == search
    for name, v in walk.items():
        assert v[0] == v[2], name      # refit of unchanged geometry reproduces the build's nodes


def test_lazy_normal_records_of_large_scenes(tpt, ctx, scenes, orc):
    """Scenes of 65 536 triangles and more leave the per-leaf shading-normal records out of the build (k_leaf_normals writes them
    when a render first needs them): the image must be the oracle's, the first and a later render of the same build identical,
    and a rebuild + render identical again."""
    from tinypathtracer_b200.soup import make_soup
    scene = make_soup(scenes["ball"], 64, 25)            # 78 124 triangles
    c2w = np.eye(4, dtype=np.float32)
    c2w[2, 3] = 130.0
    scene["c2w"] = np.ascontiguousarray(c2w.T).reshape(16)
    ds = tpt.DeviceScene(ctx, scene)
    bvh = tpt.BVH(ds)
    bvh.construct()
    cam = tpt.Camera.from_scene(scene)
    W, H, spp = 64, 36, 8
    from conftest import synthetic_sky
    env = synthetic_sky(256, 128)
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=env)
    imgs = []
    for rebuild in (False, False, True):
        if rebuild:
            bvh.construct()
        pt.m_radiance.clear()
        pt.render_into(ds, cam, spp=spp, max_depth=4, seed=9)
        imgs.append(pt.m_radiance.to_i64())
    assert (imgs[0] == imgs[1]).all() and (imgs[0] == imgs[2]).all() and imgs[0].any()
    img = pt.m_radiance.to_f32() / spp
    oimg = orc.render(scene, W, H, spp, max_depth=4, seed=9, env=env) / spp
    relmse = float(((img - oimg) ** 2).sum() / (oimg ** 2).sum())
    assert relmse < 1e-3, relmse          # same RNG: per-sample agreement up to rounding-flipped decisions
    ds.close()


def test_eight_contexts_share_one_gpu(tpt, scenes):
    """The wavefront queue pools are sized from what the device can spare (min(40 GiB, 60 % of free memory) per context, looked up at
    the context's first render), not from a constant: eight contexts on ONE GPU each render a frame whose worst case wants the full
    budget (box.gltf at 1920 x 1080: every pixel live) — the later ones get smaller batches, nothing fails, and every context
    produces the same accumulator bits."""
    import hashlib
    scene = scenes["box"]
    ctxs = [tpt.Context(0) for _ in range(8)]
    shas, keep = [], []
    for c in ctxs:
        ds = tpt.DeviceScene(c, scene)
        tpt.BVH(ds).construct(stats=False)
        pt = tpt.PathTracer(c, 1920, 1080)
        pt.render_into(ds, tpt.Camera.from_scene(scene), spp=96, max_depth=8, seed=21)
        shas.append(hashlib.sha256(pt.m_radiance.to_i64().tobytes()).hexdigest())
        assert c.counters()["queue_overflow"] == 0
        keep.append((ds, pt))          # pools stay allocated while the next context sizes its own
    assert len(set(shas)) == 1
    for c in ctxs:
        c.close()


def test_reference_style_entry_points(tpt, ctx, scenes, orc):
    """The boundary forms the reference itself uses: BVH::construct on WORLD-SPACE device arrays (include/bvh.cuh:60-68 as called
    at src/path_tracer.cu:541-542) -> tpt_bvh_build_from_world, and doTrace with a DEVICE framebuffer (include/path_tracer.h:34,
    src/vkEngine.cu:244) -> tpt_do_trace_device.  Same hits / same bytes as the host-array forms."""
    import ctypes as C
    import torch
    from tinypathtracer_b200._capi import check, lib, RenderConfig
    scene = scenes["box"]
    ds = tpt.DeviceScene(ctx, scene)
    bvh = tpt.BVH(ds)
    bvh.construct()
    wv, _ = ds.world()
    idx = np.ascontiguousarray(scene["indices"], np.uint32)
    used = np.zeros(wv.shape[0], bool)
    used[idx] = True
    wv = np.where(used[:, None], wv, 0.0).astype(np.float32)     # vertices no face references are never transformed
    d_wv, d_idx = torch.from_numpy(wv).cuda(), torch.from_numpy(idx.astype(np.int32)).cuda()
    h = C.c_void_p()
    for _ in range(2):                                           # second call: the scene is reused
        check(lib.tpt_bvh_build_from_world(ctx._h, C.c_void_p(d_wv.data_ptr()), wv.shape[0], C.c_void_p(d_idx.data_ptr()), idx.size // 3, C.byref(h), None))
    rays = orc.primary_rays(scene, 160, 90)
    want = bvh.trace(rays)
    got = np.zeros(rays.shape[0], tpt.HIT_DTYPE)
    r32 = np.ascontiguousarray(rays, np.float32)
    check(lib.tpt_trace_rays(h, r32.ctypes.data_as(C.c_void_p), rays.shape[0], got.ctypes.data_as(C.c_void_p), 0, 0))
    assert got.tobytes() == want.tobytes() and (got["prim"] >= 0).any()
    with pytest.raises(tpt.TptError):
        check(lib.tpt_bvh_build_from_world(ctx._h, C.c_void_p(d_wv.data_ptr()), wv.shape[0] - 1, C.c_void_p(d_idx.data_ptr()), idx.size // 3, C.byref(h), None))
    lib.tpt_scene_destroy(h)
    # device framebuffer
    W, H, spp = 96, 54, 8
    pt = tpt.PathTracer(ctx, W, H)
    cam = tpt.Camera.from_scene(scene)
    fb_host = np.full((H, W, 4), 7, np.uint8)
    pt.doTrace(ds, cam, fb_host, spp, max_depth=8, seed=4)
    fb_dev = torch.full((H, W, 4), 7, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    from tinypathtracer_b200.api import render_config
    cfg = render_config(W, H, spp, 8, 4)
    check(lib.tpt_do_trace_device(ds._h, C.byref(cam.c), C.byref(cfg), pt.m_radiance._h, C.c_void_p(fb_dev.data_ptr())))
    ctx.synchronize()
    assert np.array_equal(fb_dev.cpu().numpy(), fb_host) and (fb_host[..., 3] == 7).all() and fb_host[..., :3].any()
    assert lib.tpt_pointer_is_device(C.c_void_p(fb_dev.data_ptr())) == 1
    assert lib.tpt_pointer_is_device(fb_host.ctypes.data_as(C.c_void_p)) == 0
    ds.close()
