"""tpt_trace_rays_static — the allocation-free variant of tpt_trace_rays (include/tpt.h).

No counterpart in the reference: added for callers that cannot call cudaMalloc.
Its scratch (work cursor + host staging) is a file-scope __device__ array that all calls share and that keeps the last
call's contents, so the checks are: the same (prim, t, u, v) bits as the existing entry point on two calls in a row (the
second smaller and different, so stale staging or a stale cursor would show), with host and with device buffers, and a
clear refusal above the stated maximum.  Collected last on purpose.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _incoherent_rays(scene, orc, n, seed):
    wv, _ = orc.transform(scene)
    lo, hi = wv.min(0), wv.max(0)
    rng = np.random.default_rng(seed)
    o = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return np.concatenate([o, d], axis=1)


def _static(tpt, scene_h, rays, any_hit=False):
    from tinypathtracer_b200._capi import check, lib
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    hits = np.zeros(rays.shape[0], tpt.HIT_DTYPE)
    check(lib.tpt_trace_rays_static(scene_h, rays.ctypes.data_as(C.c_void_p), rays.shape[0], hits.ctypes.data_as(C.c_void_p), int(any_hit), 0))
    return hits


def test_static_trace_matches_tpt_trace_rays_on_two_calls_in_a_row(tpt, scenes, orc):
    from tinypathtracer_b200._capi import TPT_TRACE_STATIC_MAX_RAYS, lib
    ctx = tpt.Context(0)          # a context of its own: passes alone in a fresh process and after the other tests
    try:
        scene = scenes["box"]
        ds = tpt.DeviceScene(ctx, scene)
        bvh = tpt.BVH(ds)
        bvh.construct(stats=False)
        first = _incoherent_rays(scene, orc, TPT_TRACE_STATIC_MAX_RAYS, seed=101)     # the maximum size
        second = _incoherent_rays(scene, orc, 12345, seed=202)                       # smaller, different: stale scratch would show
        for rays in (first, second):
            want = bvh.trace(rays)
            got = _static(tpt, ds._h, rays)
            assert (want["prim"] >= 0).mean() > 0.5
            for f in ("prim", "t", "u", "v"):   # bit-exact, like every traversal test of this repo
                assert np.array_equal(got[f].view(np.uint32), want[f].view(np.uint32)), f
        # any-hit form: same occlusion answer
        want = bvh.trace(second, any_hit=True)
        got = _static(tpt, ds._h, second, any_hit=True)
        assert np.array_equal(got["prim"] >= 0, want["prim"] >= 0)
        # device buffers, used in place (only the cursor comes from the static array)
        import torch
        r_t = torch.from_numpy(second).cuda()
        h_t = torch.zeros((second.shape[0], 4), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        for _ in range(2):
            h_t.zero_()
            torch.cuda.synchronize()
            assert lib.tpt_trace_rays_static(ds._h, C.c_void_p(r_t.data_ptr()), second.shape[0], C.c_void_p(h_t.data_ptr()), 0, 1) == 0
            ctx.synchronize()
            got = h_t.cpu().numpy().view(tpt.HIT_DTYPE).reshape(-1)
            want = bvh.trace(second)
            for f in ("prim", "t", "u", "v"):
                assert np.array_equal(got[f].view(np.uint32), want[f].view(np.uint32)), f
        # beyond the maximum: refused with a message, nothing traced
        too_many = np.zeros((TPT_TRACE_STATIC_MAX_RAYS + 1, 6), np.float32)
        hits = np.zeros(too_many.shape[0], tpt.HIT_DTYPE)
        rc = lib.tpt_trace_rays_static(ds._h, too_many.ctypes.data_as(C.c_void_p), too_many.shape[0], hits.ctypes.data_as(C.c_void_p), 0, 0)
        assert rc == -1 and b"TPT_TRACE_STATIC_MAX_RAYS" in lib.tpt_last_error()
        # and n == 0 is a no-op
        assert lib.tpt_trace_rays_static(ds._h, None, 0, None, 0, 0) == 0
        ds.close()
    finally:
        ctx.close()
