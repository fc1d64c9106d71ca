"""tpt_comm_init_all / tpt_accum_allreduce: the one exchange step of the multi-GPU path for a host that drives
all GPUs from ONE process — every GPU reduces its share of the int64 accumulation buffers by reading its peers' memory, then
fetches the other shares; no library collective.  The reduction is exact, so after it every buffer holds the bits of one render
of all samples.  With a single GPU the same code runs between two contexts on that GPU (peer copies degenerate to local ones);
on a box with >= 2 GPUs the test uses one context per GPU."""
import ctypes as C

import numpy as np
import pytest

from conftest import synthetic_sky

pytestmark = pytest.mark.gpu


def _device_count():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("n", [2, 3, 8])
def test_allreduce_of_slices_equals_one_render(n, tpt, scenes, capsys):
    ngpu = _device_count()
    if ngpu >= n:
        ctxs = tpt.comm_init_all(n)                       # one context per GPU, peer access enabled
        form = f"REAL PEERS, one context per GPU 0..{n - 1} (peer loads / cudaMemcpyPeerAsync over NVLink)"
    else:
        ctxs = [tpt.Context(0) for _ in range(n)]          # n contexts on the one GPU there is
        form = (f"{ngpu} GPU(s) visible -> DEGENERATE form, {n} contexts on GPU 0 (peer accesses are local); the NVLink form runs in "
                "bench.py --gpus N (native_allreduce_sha256)")
    with capsys.disabled():     # visible under `pytest -q` too: a pass on a one-GPU box says nothing about the NVLink path
        print(f"\n[native all-reduce] n={n}: {form}")
    s = scenes["box2"]
    cam = tpt.Camera.from_scene(s)
    W, H, per = 81, 45, 4                                  # 81 * 45 * 3 elements: odd, and not a multiple of n
    env = synthetic_sky(64, 32)
    pts, dss = [], []
    for r, ctx in enumerate(ctxs):
        ds = tpt.DeviceScene(ctx, s)
        tpt.BVH(ds).construct()
        pt = tpt.PathTracer(ctx, W, H, env_rgba8=env)
        pt.render_into(ds, cam, spp=per, seed=5, sample_offset=r, sample_stride=n)     # rank r: sample indices = r (mod n)
        pts.append(pt)
        dss.append(ds)
    parts = [pt.m_radiance.to_i64() for pt in pts]
    tpt.accum_allreduce([pt.m_radiance for pt in pts])
    whole = tpt.PathTracer(ctxs[0], W, H, env_rgba8=env)
    whole.render_into(dss[0], cam, spp=per * n, seed=5)
    want = whole.m_radiance.to_i64()
    assert (sum(parts) == want).all() and want.any()
    for pt in pts:
        assert (pt.m_radiance.to_i64() == want).all()      # every GPU holds the total, bit for bit
    # the next frame may start at once: clear + render + all-reduce again gives the same bits
    for r, (pt, ds) in enumerate(zip(pts, dss)):
        pt.m_radiance.clear()
        pt.render_into(ds, cam, spp=per, seed=5, sample_offset=r, sample_stride=n)
    tpt.accum_allreduce([pt.m_radiance for pt in pts])
    for pt in pts:
        assert (pt.m_radiance.to_i64() == want).all()
    for ctx in ctxs:
        ctx.close()


def test_allreduce_argument_errors(tpt, ctx):
    a = tpt.Accum(ctx, 32, 18)
    b = tpt.Accum(ctx, 16, 18)
    tpt.accum_allreduce([a])                               # n = 1: nothing to do
    with pytest.raises(tpt.TptError):
        tpt.accum_allreduce([a, b])                        # sizes differ
    with pytest.raises(tpt.TptError):
        tpt.accum_allreduce([a, a])                        # the same buffer twice
    with pytest.raises(tpt.TptError):
        tpt.comm_init_all(_device_count() + 1)             # more GPUs than there are
    a.close()
    b.close()
