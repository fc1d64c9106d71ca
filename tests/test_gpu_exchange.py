"""tpt_exchange_* (one process PER GPU): the fused "sum the accumulators + copyToFB" pass over peer memory.
Every rank renders its sample slice, then all call resolve(): the root's framebuffer must hold, byte for byte, what one
context that rendered ALL samples gets from tpt_resolve_bgra8.  The ranks are real processes (CUDA IPC mappings cannot be
opened inside the process that exported them); with fewer GPUs than ranks they share GPU 0 — the mappings and the flag
handshake are the same, the loads are local instead of NVLink.  bench.py --gpus N runs the same path on N real GPUs."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parents[1]
W, H, PER, SEED = 81, 45, 4, 5       # 81 * 45 pixels: odd, so the last rank's part ends in half a 128-bit word


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _sky():
    sys.path.insert(0, str(ROOT / "tests"))
    from conftest import synthetic_sky
    return synthetic_sky(64, 32)


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, str(ROOT))
    import tinypathtracer_b200 as tpt
    from tinypathtracer_b200.scene_io import load_scene_npz
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    ctx = tpt.Context(dev)
    s = load_scene_npz(ROOT / "tests" / "golden" / "scenes" / "box2.npz")
    cam = tpt.Camera.from_scene(s)
    ds = tpt.DeviceScene(ctx, s)
    tpt.BVH(ds).construct()
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=_sky())
    acc = tpt.Accum(ctx, W, H)
    x = tpt.FrameExchange(ctx, acc, rank, world, alpha=0)
    x.connect_over(dist)
    frames = []
    for frame, root in enumerate((0, -1, world - 1)):          # three frames in a row; to rank 0, to everybody, to the last rank
        acc.clear()
        pt.render_into(ds, cam, spp=PER, seed=SEED + frame, sample_offset=rank, sample_stride=world, accum=acc)
        x.resolve(PER * world, root)
        x.status()                                             # synchronises; raises on a handshake timeout
        frames.append(x.framebuffer())
        dist.barrier()                                         # nobody starts the next frame's comparison copy early
    np.save(Path(out_dir) / f"fb_{rank}.npy", np.stack(frames))
    dist.barrier()
    x.close()                                                  # every rank closes its mappings before any accumulator goes away
    dist.barrier()
    acc.close()
    ctx.close()
    dist.destroy_process_group()


def _expected(tpt, ctx, scenes, frame):
    s = scenes["box2"]
    cam = tpt.Camera.from_scene(s)
    ds = tpt.DeviceScene(ctx, s)
    tpt.BVH(ds).construct()
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=_sky())
    return ds, cam, pt


@pytest.mark.parametrize("world", [2, 3])
def test_exchange_equals_one_render_resolved(world, tpt, ctx, scenes, tmp_path, capsys):
    import torch
    import torch.multiprocessing as mp
    ngpu = torch.cuda.device_count()
    with capsys.disabled():
        print(f"\n[frame exchange] world={world}: " + ("REAL PEERS (one GPU per rank, NVLink)" if ngpu >= world else
              f"{ngpu} GPU(s) visible -> ranks share GPU 0 (IPC mappings + flag handshake are the same, loads are local)"))
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    ds, cam, pt = _expected(tpt, ctx, scenes, 0)
    got = [np.load(tmp_path / f"fb_{r}.npy") for r in range(world)]
    for frame, root in enumerate((0, -1, world - 1)):
        whole = tpt.Accum(ctx, W, H)
        pt.render_into(ds, cam, spp=PER * world, seed=SEED + frame, accum=whole)
        want = whole.resolve_bgra8(PER * world)
        assert want[..., :3].any()
        holders = range(world) if root < 0 else [root]
        for r in holders:
            np.testing.assert_array_equal(got[r][frame], want, err_msg=f"frame {frame}, rank {r}")
        whole.close()


def _worker_missing_rank(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, str(ROOT))
    os.environ["TPT_EXCHANGE_TIMEOUT_MS"] = "400"              # read when the exchange is created
    import tinypathtracer_b200 as tpt
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    ctx = tpt.Context(dev)
    acc = tpt.Accum(ctx, 32, 18)
    x = tpt.FrameExchange(ctx, acc, rank, world)
    x.connect_over(dist)
    acc.clear()
    x.resolve(1, 0)                                            # frame 1: everybody takes part
    x.status()
    dist.barrier()
    outcome = "skipped"
    if rank == 0:                                              # frame 2: rank 1 never calls resolve
        x.resolve(1, 0)
        try:
            x.status()
            outcome = "no error"
        except tpt.TptError as e:
            outcome = "timeout reported: " + str(e)
    (Path(out_dir) / f"outcome_{rank}.txt").write_text(outcome)
    dist.barrier()
    x.close()
    dist.barrier()
    acc.close()
    ctx.close()
    dist.destroy_process_group()


def test_a_rank_that_never_arrives_is_an_error_not_a_hang(tmp_path):
    """The handshake polls peer memory from kernels; a peer that never signals must cost a time-out and an error from
    tpt_exchange_status, never a GPU that spins for ever."""
    import time
    import torch.multiprocessing as mp
    t0 = time.time()
    mp.spawn(_worker_missing_rank, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "outcome_0.txt").read_text().startswith("timeout reported"), (tmp_path / "outcome_0.txt").read_text()
    assert time.time() - t0 < 120


def test_exchange_world_one_is_a_plain_resolve(tpt, ctx, scenes):
    s = scenes["box2"]
    cam = tpt.Camera.from_scene(s)
    ds = tpt.DeviceScene(ctx, s)
    tpt.BVH(ds).construct()
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=_sky())
    acc = tpt.Accum(ctx, W, H)
    pt.render_into(ds, cam, spp=8, seed=3, accum=acc)
    want = acc.resolve_bgra8(8)
    x = tpt.FrameExchange(ctx, acc, 0, 1, alpha=0)
    for _ in range(2):                                         # twice: the completion counter is reset by the kernel itself
        x.resolve(8, 0)
        x.status()
        got = x.framebuffer()
        np.testing.assert_array_equal(got, want)
    x.close()
    acc.close()


def test_exchange_argument_errors(tpt, ctx):
    import torch
    acc = tpt.Accum(ctx, 32, 18)
    with pytest.raises(tpt.TptError):
        tpt.FrameExchange(ctx, acc, 2, 2)                      # rank out of range
    with pytest.raises(tpt.TptError):
        tpt.FrameExchange(ctx, acc, 0, 17)                     # more ranks than the flag block holds
    t = torch.zeros((18, 32, 3), dtype=torch.int64, device="cuda")
    wrapped = tpt.Accum(ctx, 32, 18, device_ptr=t.data_ptr())
    with pytest.raises(tpt.TptError):
        tpt.FrameExchange(ctx, wrapped, 0, 2)                  # peers map the allocation itself: it must be the library's own
    x = tpt.FrameExchange(ctx, acc, 0, 2)
    with pytest.raises(tpt.TptError):
        x.resolve(4, 0)                                        # not connected
    with pytest.raises(tpt.TptError):
        x.connect(b"\0" * 10)                                  # wrong size
    with pytest.raises(tpt.TptError):
        x.connect([x.handle, x.handle])                        # second handle is not rank 1's
    x.close()
    wrapped.close()
    acc.close()
