"""The N > 1 host logic on CPU: world_size-2 `gloo` processes deal the 8 fixed spp slices exactly like bench.py does
(slice s = sample indices s, s+8, ...; slice -> rank s mod N), accumulate their samples as int64 fixed point
and all-reduce with SUM.  The result must be bit-identical to one process summing every sample: integer sums do not
depend on the partition or on the reduction order.  Per-sample radiance comes from the CPU oracle (the checker), since
no GPU is present here; the GPU counterpart is tests/test_gpu_parity.py::test_slices_are_bit_identical."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
W, H, SPP, FRAC = 24, 16, 16, 36


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fixed(img):
    return np.rint(img.astype(np.float64) * float(1 << FRAC)).astype(np.int64)


def _sample_images(indices):
    sys.path.insert(0, str(ROOT))
    from oracle import gltf_scene, pyoracle as orc
    scene = gltf_scene.load_scene(ROOT / "tests" / "golden" / "scenes" / "box2.npz")
    wv, wn = orc.transform(scene)
    nodes, _, _ = orc.build_reference_bvh(wv, scene["indices"])
    acc = np.zeros((H, W, 3), np.int64)
    for i in indices:
        acc += _fixed(orc.render(scene, W, H, 1, max_depth=4, seed=11, sample_offset=int(i), prepared=(wv, wn, nodes)))
    return acc


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, str(ROOT))
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench", ROOT / "bench.py")
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    slices = bench.my_slices(rank, world)
    # the one-call form bench.py uses (offset = rank, stride = world) covers exactly the union of this rank's slices
    mine = [i for i in range(SPP) if i % world == rank]
    assert sorted(mine) == sorted(i for s in slices for i in range(s, SPP, bench.N_SLICES))
    acc = torch.from_numpy(_sample_images(mine))
    dist.all_reduce(acc, op=dist.ReduceOp.SUM)
    if rank == 0:
        np.save(Path(out_dir) / "reduced.npy", acc.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_slices_partition_the_samples():
    sys.path.insert(0, str(ROOT))
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench", ROOT / "bench.py")
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    for world in (1, 2, 4, 8):
        seen = sorted(s for r in range(world) for s in bench.my_slices(r, world))
        assert seen == list(range(bench.N_SLICES))
        for r in range(world):
            assert all(s % world == r for s in bench.my_slices(r, world))


def test_two_ranks_reduce_to_the_single_rank_buffer(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    reduced = np.load(tmp_path / "reduced.npy")
    single = _sample_images(range(SPP))
    assert reduced.dtype == np.int64 and reduced.any()
    np.testing.assert_array_equal(reduced, single)
