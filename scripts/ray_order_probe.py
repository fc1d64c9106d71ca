"""Probe: how much does the ORDER of the rays matter to tpt_trace_rays on the 10 M-triangle soup?  (torch-side permutations only;
decides whether a device-side ray sort inside tpt_trace_rays is worth building.)  python scripts/ray_order_probe.py [--small]"""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import tinypathtracer_b200 as tpt  # noqa: E402
from tinypathtracer_b200.api import primary_rays_device  # noqa: E402
from tinypathtracer_b200.soup import make_soup, secondary_rays  # noqa: E402
from oracle import gltf_scene  # noqa: E402

small = "--small" in sys.argv
nb, nx = (64, 25) if small else (8192, 3211)
scene = make_soup(gltf_scene.load_scene(ROOT / "tests" / "golden" / "scenes" / "ball.npz"), nb, nx)
W, H = 3840, 2160
ctx = tpt.Context(0)
stream = torch.cuda.current_stream()
ctx.set_stream(stream.cuda_stream)
ds = tpt.DeviceScene(ctx, scene)
bvh = tpt.BVH(ds)
st = bvh.construct()
cam = tpt.Camera.from_scene(scene)
n = W * H
rays = torch.empty((n, 6), dtype=torch.float32, device="cuda")
hits = torch.empty((n, 4), dtype=torch.float32, device="cuda")
primary_rays_device(ctx, cam, W, H, rays.data_ptr())
bvh.trace_device(rays.data_ptr(), n, hits.data_ptr())
torch.cuda.synchronize()
rh = rays.cpu().numpy().view(tpt.RAY_DTYPE).reshape(-1)
hh = hits.cpu().numpy().view(tpt.HIT_DTYPE).reshape(-1)
wv, _ = ds.world()
sec = torch.from_numpy(secondary_rays(rh, hh, wv, scene["indices"]).view(np.float32).reshape(-1, 6)).cuda()


def timed(r, reps=10):
    h = torch.empty((r.shape[0], 4), dtype=torch.float32, device="cuda")
    for _ in range(3):
        bvh.trace_device(r.data_ptr(), r.shape[0], h.data_ptr())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        bvh.trace_device(r.data_ptr(), r.shape[0], h.data_ptr())
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def spread(v, bits):   # interleave helper: spread `bits` bits of v to every third position
    out = torch.zeros_like(v)
    for b in range(bits):
        out |= ((v >> b) & 1) << (3 * b)
    return out


def morton_key(r, ob=7, db=3):
    lo = torch.tensor(st["bounds_min"], device="cuda")
    hi = torch.tensor(st["bounds_max"], device="cuda")
    q = (((r[:, :3] - lo) / (hi - lo)).clamp(0, 0.999999) * (1 << ob)).long()
    d = ((r[:, 3:] * 0.5 + 0.5).clamp(0, 0.999999) * (1 << db)).long()
    ko = spread(q[:, 0], ob) | (spread(q[:, 1], ob) << 1) | (spread(q[:, 2], ob) << 2)
    kd = spread(d[:, 0], db) | (spread(d[:, 1], db) << 1) | (spread(d[:, 2], db) << 2)
    return ko, kd


print(f"faces {st['n_faces']}  build {st['build_ms']:.3f} ms")
print(f"primary, row-major (as given): {timed(rays):.3f} ms  ({n / timed(rays) / 1e3:.0f} Mrays/s)")
ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
tile = ((ys // 4) * (W // 8) + xs // 8) * 32 + (ys % 4) * 8 + xs % 8
perm = torch.argsort(tile.reshape(-1))
print(f"primary, 8x4 tiles:            {timed(rays[perm].contiguous()):.3f} ms")
_, kd = morton_key(rays, 7, 8)
print(f"primary, direction morton 24b: {timed(rays[torch.argsort(kd)].contiguous()):.3f} ms")
t_sec = timed(sec)
print(f"secondary, as given:           {t_sec:.3f} ms  ({sec.shape[0] / t_sec / 1e3:.0f} Mrays/s)")
for ob, db in ((5, 0), (7, 0), (10, 0), (7, 1), (7, 2), (7, 3), (10, 2)):
    ko, kd = morton_key(sec, ob, max(db, 1))
    key = (ko << (3 * db)) | (kd if db else 0)
    p = torch.argsort(key)
    print(f"secondary, origin {3 * ob}b + dir {3 * db}b: {timed(sec[p].contiguous()):.3f} ms")
print(f"secondary, random order:       {timed(sec[torch.randperm(sec.shape[0], device='cuda')].contiguous()):.3f} ms")
