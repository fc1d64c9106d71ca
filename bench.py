#!/usr/bin/env python
"""bench.py — headline benchmark of the path-tracing hot path (BASELINE.json metric: Mrays/s & samples/s at
1920x1080; BVH build ms).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload ball|light|box|tir]

A *step* is one whole frame of the workload, as PathTracer::doTrace does it in the reference
(src/path_tracer.cu:491-554): object->world transform + full BVH rebuild + clear + `spp` samples per pixel
+ resolve to BGRA8.  Default workload = BASELINE.json configs[1]: ball.gltf, 1920x1080, 256 spp, depth 8,
synthetic sky environment (the shipped JPG is a missing blob).  Inputs are the flattened scene fixtures under
tests/golden/scenes (no /root/reference at run time); the camera, materials and lights are the glTF's.

N > 1 (torchrun, one rank per GPU): the 8 fixed spp slices of SURVEY §8(e) are dealt round-robin to the
ranks (strong scaling: total spp fixed); every rank holds a replicated scene + BVH; one NCCL all-reduce (SUM,
int64 fixed point — exact, so the image is bit-identical for any N) combines the accumulation buffers.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

WORKLOADS = {
    # name: (scene fixture, width, height, spp, max_depth, env, seed)   — BASELINE.json configs
    "ball": ("ball", 1920, 1080, 256, 8, "sky", 2),     # configs[1] (a)
    "light": ("light", 1920, 1080, 256, 8, "sky", 2),   # configs[1] (b)
    "box": ("box", 512, 512, 64, 8, None, 1),           # configs[0]
    "tir": ("tir", 1920, 1080, 256, 32, None, 3),       # configs[2]
    "box2_4k": ("box2", 3840, 2160, 4096, 8, "sky", 5),  # configs[4]
}
N_SLICES = 8
# traverseBVH calls per camera sample of the REFERENCE on each workload, counted by its counting build (oracle/_ref/tpt_ref_headless_count,
# `bench.py --impl reference` prints them as rays_per_sample; profiles/r02_reference_ball.json).  The product traverses fewer rays for the
# same samples (pixel-pyramid culling, no "direct" rays in emitter-free scenes), so its own Mrays/s is over a smaller ray population;
# mrays_per_s_at_reference_ray_count = these x the product's Msamples/s is the like-for-like figure for north_star's Mrays/s wording.
REFERENCE_RAYS_PER_SAMPLE = {"ball": 1.5419, "box": 4.4084, "tir": 1.1797, "light": 1.3210}


def synthetic_sky(width=4096, height=2048):
    from tinypathtracer_b200.sky import synthetic_sky as sky
    return sky(width, height)


def load_scene(name):
    from tinypathtracer_b200.scene_io import load_scene_npz
    return load_scene_npz(ROOT / "tests" / "golden" / "scenes" / f"{name}.npz")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None
        self.i0 = 0

    def mark(self):
        """The timed region starts now: rows from here on (and the one just before, taken under the warm-up's load) are the ones reported.
        The sampler itself is started earlier — nvidia-smi needs ~0.1 s before its first row, longer than a short timed region."""
        self.i0 = max(len(self.rows) - 1, 0)

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "25", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        t0 = time.perf_counter()
        while len(self.rows) <= self.i0 + 1 and time.perf_counter() - t0 < 0.3:   # a very short region: wait for the row that was being taken
            time.sleep(0.01)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for r in self.rows[self.i0:]:
            try:
                sm.append(float(r[0]))
                smax.append(float(r[1]))
            except Exception:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def my_slices(rank: int, world: int):
    """8 fixed slices (slice s = sample indices s, s+8, ...), slice -> GPU map s mod N."""
    return [s for s in range(N_SLICES) if s % world == rank]


def _ncu_capture(fixture: str, kernel: str):
    """Counters of the committed `ncu --set full` capture of `kernel` on this workload (profiles/, newest round first):
    {duration_ms, dram_bytes, issue_active_pct, lanes_per_inst, l1_wavefront_pct, file} or None."""
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    for rnd in ("r02", "r01"):
        f = ROOT / "profiles" / f"{rnd}_ncu_full_{kernel}_{fixture}.csv"
        if not f.exists():
            continue
        try:
            rows = {r[0]: r for r in (ln.split(",") for ln in f.read_text().splitlines()) if len(r) >= 3}
            col = 2          # launch1
            val = lambda k: float(rows[k][col])
            return {"duration_ms": val("gpu__time_duration.sum") * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(rows["gpu__time_duration.sum"][1], 1.0),
                    "dram_bytes": sum(val(k) * unit[rows[k][1]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum")),
                    "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                    "lanes_per_inst": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
                    "l1_wavefront_pct": val("l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed")
                    if "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed" in rows else None,
                    "file": str(f.relative_to(ROOT))}
        except Exception:
            continue
    return None


HOST_GROUP = None   # gloo group of the ranks (host-side barriers, exchange handles); set by run_ours when world > 1


class _DevicePtr:
    """A device buffer the library owns, as something torch.as_tensor can view (CUDA array interface)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


class FrameBench:
    """One workload of WORKLOADS on this rank's GPU: device-resident frames (`value`), end-to-end frames through host buffers
    (`e2e`), and the counters the roofline object is computed from.

    Frames are independent (the reference's loop renders one after the other from scratch, src/vkEngine.cu:294-302), so
    `--frames-in-flight` (default: auto = 4) of them are kept in flight: frame i runs on lane i mod 4 — a context, stream, device scene,
    queue pools, accumulator and framebuffer of its own — and the latency-bound end of one frame (its deep waves hold a
    handful of rays and cost the latency of the slowest, the all-reduce, the resolve, the read-back) overlaps the start of
    the next.  Every frame still does everything: build, clear, all its samples, the exchange between the ranks, resolve.

    The exchange at world > 1 (--exchange): "peer" (default) = tpt_exchange_resolve, the library's own fused kernel — every rank
    sums its 1/N of the pixels out of all ranks' accumulators over NVLink peer mappings, resolves them and stores the BGRA8
    pixels into rank 0's framebuffer; "nccl" = torch.distributed all-reduce of the int64 buffers, then every rank resolves the
    whole frame (round 1's path; still what `accum_sha256` is computed from, and the byte-for-byte check of the peer path)."""

    def __init__(self, args, workload, spp_override=0, lanes=None):
        import torch
        import torch.distributed as dist
        import tinypathtracer_b200 as tpt
        self.torch, self.dist, self.tpt, self.args = torch, dist, tpt, args
        self.rank, self.world, self.local = dist_env()
        self.fixture, self.W, self.H, self.spp, self.depth, self.envkind, self.seed = WORKLOADS[workload]
        if spp_override:
            self.spp = spp_override
        assert self.spp % N_SLICES == 0, "spp must be a multiple of the 8 fixed slices"
        assert N_SLICES % self.world == 0, "the 8 fixed slices are dealt to 1, 2, 4 or 8 ranks"
        self.per_slice = self.spp // N_SLICES
        self.scene = load_scene(self.fixture)
        self.env = synthetic_sky() if self.envkind == "sky" else None
        W, H = self.W, self.H
        want = lanes if lanes is not None else args.frames_in_flight
        self.tune = not want      # auto: lanes are built for the larger choice, prime() measures both and keeps the faster
        if not want:
            # What a frame cannot shrink with its sample count (deep waves of a few rays, the exchange, kernel tails) is hidden by the other
            # frames in flight, and the shorter the frame the more of them it takes — ball.gltf, ms per step with 2 | 4 frames in flight:
            # 256 spp 25.61 | 25.54, 128 spp 13.03 | 12.97, 64 spp 6.70 | 6.62, 32 spp (one of 8 ranks) 3.65 | 3.51 (6: no better).  A frame
            # whose waves are all large has nothing to hide and only pays for the extra queues (box.gltf: 9.16 | 9.37).  So the choice is
            # MEASURED: prime() times a few untimed frames with 2 and with 4 and keeps the faster (rank 0 decides for all ranks).  Very
            # large jobs (4K, 4096 spp) keep 2: four sets of queue pools would not fit.
            want = 4 if W * H * (self.spp // self.world) <= (1 << 30) else 2
            self.tune = want == 4
        self.tuning = None
        self.n_lanes = max(1, want)
        self.n_built = self.n_lanes
        self.main = torch.cuda.current_stream()
        self.streams = [self.main] + [torch.cuda.Stream() for _ in range(self.n_lanes - 1)]
        self.ctxs, self.pts, self.acc_ts, self.accs, self.dscenes, self.fb_dev, self.fb_host = [], [], [], [], [], [], []
        for k in range(self.n_lanes):
            c = tpt.Context(self.local)
            c.set_stream(self.streams[k].cuda_stream)          # kernels, torch events and NCCL of a lane all on one stream
            self.ctxs.append(c)
            self.pts.append(tpt.PathTracer(c, W, H, env_rgba8=self.env))
            t = torch.zeros((H, W, 3), dtype=torch.int64, device="cuda")     # accumulation buffer owned by torch (NCCL operand)
            self.acc_ts.append(t)
            self.accs.append(tpt.Accum(c, W, H, device_ptr=t.data_ptr()))
            self.dscenes.append(tpt.DeviceScene(c, self.scene))
            self.fb_dev.append(torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda"))
            self.fb_host.append(torch.zeros((H, W, 4), dtype=torch.uint8).pin_memory())
        self.ctx, self.pt, self.acc_t, self.acc, self.dscene, self.stream = (self.ctxs[0], self.pts[0], self.acc_ts[0], self.accs[0],
                                                                             self.dscenes[0], self.main)
        self.exchange = "none" if self.world == 1 else args.exchange
        self.exchange_note = None
        self.xaccs, self.xchg, self.fb_view = [], [], []
        if self.exchange == "peer":
            try:
                if os.environ.get("TPT_BENCH_FAIL_EXCHANGE_ON_RANK") == str(self.rank):   # test hook for the fallback below
                    raise tpt.TptError("forced by TPT_BENCH_FAIL_EXCHANGE_ON_RANK")
                for k in range(self.n_lanes):
                    a = tpt.Accum(self.ctxs[k], W, H)                # the library's own allocation: peers map it through CUDA IPC
                    x = tpt.FrameExchange(self.ctxs[k], a, self.rank, self.world, alpha=0)
                    self.xaccs.append(a)
                    self.xchg.append(x)
                ok = 1
            except tpt.TptError as e:
                ok, self.exchange_note = 0, str(e)
            # every collective first, then the local part that can fail: no rank is ever left alone in an all_gather
            mine = [x.handle for x in self.xchg] if ok else None
            everyone = [None] * self.world
            dist.all_gather_object(everyone, mine, group=HOST_GROUP)
            if all(h is not None for h in everyone):
                try:
                    for k, x in enumerate(self.xchg):
                        x.connect([everyone[r][k] for r in range(self.world)])
                except tpt.TptError as e:
                    ok, self.exchange_note = 0, str(e)
            else:
                ok = 0
            flags = [None] * self.world
            dist.all_gather_object(flags, ok, group=HOST_GROUP)
            if not all(flags):                                       # e.g. no IPC in this container: say so and use the library collective
                if self.rank == 0:
                    print(f"bench.py: peer exchange unavailable ({self.exchange_note or 'another rank failed'}); using NCCL", file=sys.stderr)
                self.exchange, self.exchange_note = "nccl", self.exchange_note or "another rank could not set up the peer exchange"
                for x in self.xchg:
                    x.close()
                for a in self.xaccs:
                    a.close()
                self.xchg, self.xaccs = [], []
            else:
                self.fb_view = [torch.as_tensor(_DevicePtr(x.framebuffer_ptr(), (H, W, 4)), device="cuda") for x in self.xchg]
        self.cam = tpt.Camera.from_scene(self.scene)
        self.slices = my_slices(self.rank, self.world)
        self.copy_stream = torch.cuda.Stream()
        self.workload_name = f"{self.fixture}.gltf {W}x{H} {self.spp}spp depth{self.depth} env={self.envkind or 'black'} seed={self.seed}"

    def close(self):
        self.torch.cuda.synchronize()
        for x in self.xchg:
            x.status()                                               # raises if a handshake ever timed out
            x.close()
        if self.xchg:
            self.dist.barrier(group=HOST_GROUP)                      # every rank has closed its mappings before any accumulator is freed
        for a in self.xaccs:
            a.close()
        for k in range(self.n_built):
            self.dscenes[k].close()
            self.accs[k].close()
            self.ctxs[k].close()

    def frame(self, i=0, dscene=None, counters=False, stage_times=False, flags_nee=False, exchange=None):
        """One doTrace (reference src/path_tracer.cu:491-554) with device-resident inputs, on lane i mod n_lanes."""
        import ctypes as C
        from tinypathtracer_b200._capi import lib, check
        tpt, dist, torch = self.tpt, self.dist, self.torch
        k = i % self.n_lanes
        dscene = dscene or self.dscenes[k]
        exchange = exchange or self.exchange
        if exchange == "peer":
            with torch.cuda.stream(self.streams[k]):
                tpt.BVH(dscene).construct(stats=False)
                self.xaccs[k].clear()
                self.pts[k].render_into(dscene, self.cam, spp=self.per_slice * len(self.slices), max_depth=self.depth, seed=self.seed,
                                        sample_offset=self.rank, sample_stride=self.world, accum=self.xaccs[k], counters=counters,
                                        stage_times=stage_times, env_nee=flags_nee, paths_in_flight=self.args.paths_in_flight)
                self.xchg[k].resolve(self.spp, 0)                    # handshake + reduce-scatter + resolve + pixels to rank 0, one pass
            return
        with torch.cuda.stream(self.streams[k]):
            tpt.BVH(dscene).construct(stats=False)
            self.accs[k].clear()
            # this rank's slices {s : s mod N == rank} are exactly the sample indices i with i mod N == rank (N divides 8):
            # one call with offset = rank, stride = N renders them all (integer accumulation: grouping cannot change a bit)
            self.pts[k].render_into(dscene, self.cam, spp=self.per_slice * len(self.slices), max_depth=self.depth, seed=self.seed,
                                    sample_offset=self.rank, sample_stride=self.world, accum=self.accs[k], counters=counters, stage_times=stage_times,
                                    env_nee=flags_nee, paths_in_flight=self.args.paths_in_flight)
            if self.world > 1:
                dist.all_reduce(self.acc_ts[k], op=dist.ReduceOp.SUM)
            check(lib.tpt_resolve_bgra8(self.accs[k]._h, self.spp, C.c_void_p(self.fb_dev[k].data_ptr()), 1))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, fn, steps, finish=None):
        """K calls of fn bracketed by barrier + synchronize, CUDA events on the launching stream (the lanes' streams start after
        the first event and the second waits for all of them), max over ranks (ms total)."""
        torch, dist = self.torch, self.dist
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.main)
        for s in self.streams[1:]:
            s.wait_event(e0)
        t0 = time.perf_counter()
        for i in range(steps):
            fn(i)
        self.host_ms = (time.perf_counter() - t0) * 1e3 / max(steps, 1)   # what the host spends enqueueing one step (Python + driver)
        if finish:
            finish()
        for s in self.streams[1:]:
            ev = torch.cuda.Event()
            ev.record(s)
            self.main.wait_event(ev)
        e1.record(self.main)
        self.barrier()
        ms = e0.elapsed_time(e1)
        if self.world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def counters_sum(self):
        out = None
        for c in self.ctxs:
            d = c.counters()
            out = d if out is None else {k: out[k] + d[k] for k in d}
        return out

    def reset_counters(self):
        for c in self.ctxs:
            c.reset_counters()

    def prime(self, **kw):
        """One untimed frame on EVERY lane before the warm-up steps: a lane's first frame allocates its queue pools (tens of ms for
        tens of GB) and measures the live-pixel map; with more lanes than warm-up steps that would land in the timed region."""
        for k in range(self.n_lanes):
            self.frame(k, **kw)
        self.torch.cuda.synchronize()
        if self.tune and self.n_lanes == 4:
            built, res = self.n_lanes, {}
            for n in (2, 4):
                self.n_lanes = n
                for i in range(n):
                    self.frame(i, **kw)
                res[n] = self.timed(lambda i: self.frame(i, **kw), 8) / 8
            pick = [min(res, key=res.get)]
            if self.world > 1:
                self.dist.broadcast_object_list(pick, src=0, group=HOST_GROUP)
            self.n_lanes, self.tune = int(pick[0]), False
            self.tuning = {"ms_per_step_with_2": round(res[2], 3), "ms_per_step_with_4": round(res[4], 3), "lanes_built": built}

    def run_value(self, steps, warmup, **kw):
        self.prime(**kw)
        for i in range(warmup):
            self.frame(i, **kw)
        self.reset_counters()
        ms = self.timed(lambda i: self.frame(i, **kw), steps) / steps
        ctr = self.counters_sum()
        if ctr["queue_overflow"] or ctr["dropped_nonfinite"] > 1e-6 * max(ctr["samples"], 1):
            raise SystemExit(f"invalid run: queue_overflow={ctr['queue_overflow']} dropped_nonfinite={ctr['dropped_nonfinite']}")
        return ms, ctr

    def run_e2e(self, steps):
        """The same frame end to end through the C ABI: every step uploads the scene's host arrays (tpt_scene_update: H2D into
        the lane's existing device arrays) and reads the framebuffer back into pinned host memory.  The read-back of a frame
        runs on a copy stream while the next frames are built and traced (a framebuffer per lane); the timed region ends when
        the last read-back has landed."""
        torch = self.torch
        ds = [self.tpt.DeviceScene(self.ctxs[k], self.scene) for k in range(self.n_lanes)]   # allocations made once, outside the timed region
        done = [torch.cuda.Event() for _ in range(self.n_lanes)]
        copied = [torch.cuda.Event() for _ in range(self.n_lanes)]
        used = [False] * self.n_lanes

        desc = self.tpt.DeviceScene.prepare(self.scene, pinned=True)   # the caller's arrays in page-locked memory, as the C ABI takes them

        def step(i):
            k = i % self.n_lanes
            with torch.cuda.stream(self.streams[k]):
                ds[k].update_prepared(desc)                        # tpt_scene_update: H2D of every scene array
                if used[k]:
                    self.streams[k].wait_event(copied[k])          # the lane's framebuffer is free again once its last read-back is done
            self.frame(i, ds[k])
            done[k].record(self.streams[k])
            self.copy_stream.wait_event(done[k])
            with torch.cuda.stream(self.copy_stream):
                if self.exchange != "peer":
                    self.fb_host[k].copy_(self.fb_dev[k], non_blocking=True)   # D2H of the step's result
                elif self.rank == 0:
                    self.fb_host[k].copy_(self.fb_view[k], non_blocking=True)  # the frame lives in rank 0's framebuffer
                copied[k].record(self.copy_stream)
            used[k] = True

        def finish():
            for k in range(self.n_lanes):
                if used[k]:
                    self.main.wait_event(copied[k])

        for k in range(self.n_lanes):                              # every lane once, untimed (first-use allocations, staging buffers)
            step(k)
        finish()
        ms = self.timed(step, steps, finish) / steps
        self.torch.cuda.synchronize()
        for d in ds:
            d.close()
        return ms

    def accum_sha(self):
        import hashlib
        self.frame(0, exchange="nccl" if self.world > 1 else None)
        self.torch.cuda.synchronize()
        return hashlib.sha256(self.acc_ts[0].cpu().numpy().tobytes()).hexdigest() if self.rank == 0 else None

    def exchange_alone(self, reps=5):
        """The exchange by itself on otherwise idle GPUs: CUDA events around tpt_exchange_resolve (handshake + fused reduce + resolve
        + completion wait) after a barrier, max over ranks, best of `reps`; and the NCCL all-reduce + full resolve it replaces."""
        if self.world == 1:
            return None
        import ctypes as C
        from tinypathtracer_b200._capi import lib, check
        torch, dist = self.torch, self.dist
        out = {}
        for mode in (["peer"] if self.exchange == "peer" else []) + ["nccl"]:
            best = None
            for _ in range(reps + 1):
                self.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(self.main)
                if mode == "peer":
                    self.xchg[0].resolve(self.spp, 0)
                else:
                    dist.all_reduce(self.acc_ts[0], op=dist.ReduceOp.SUM)
                    check(lib.tpt_resolve_bgra8(self.accs[0]._h, self.spp, C.c_void_p(self.fb_dev[0].data_ptr()), 1))
                e1.record(self.main)
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                best = float(t.item()) if best is None else min(best, float(t.item()))   # the first repetition is a warm-up
            out[mode + "_ms"] = round(best, 4)
        n, px = self.world, self.W * self.H
        if "peer_ms" in out:
            out["peer_bytes_in_per_rank"] = (n - 1) * (px // n) * 24
            out["peer_gbs_in_per_rank"] = round(out["peer_bytes_in_per_rank"] / (out["peer_ms"] * 1e-3) / 1e9, 1)
        return out

    def exchange_check(self):
        """The peer exchange's framebuffer on rank 0 against all-reduce + tpt_resolve_bgra8 of the same frame: must be the same bytes."""
        if self.exchange != "peer":
            return None
        self.frame(0)
        self.frame(0, exchange="nccl")
        self.torch.cuda.synchronize()
        self.barrier()
        return bool((self.fb_view[0] == self.fb_dev[0]).all().item()) if self.rank == 0 else None

    def rays_traversed(self, ctr):
        # camera samples whose pixel pyramid misses every leaf box are answered without a traversal
        return ctr["rays_primary"] - ctr["rays_culled"] + ctr["rays_extension"] + ctr["rays_shadow"] + ctr["rays_direct"]


def native_allreduce_check(fb: "FrameBench"):
    """tpt_accum_allreduce (csrc/comm.cu: peer loads + cudaMemcpyPeerAsync over NVLink, no library collective) on REAL peers:
    rank 0 alone drives all `world` GPUs of the box from one process — one context per GPU, each renders its rank's samples —
    and the all-reduced buffer must carry the SHA-256 of the NCCL frames.  The other ranks wait at a host (gloo) barrier."""
    import hashlib
    tpt, torch = fb.tpt, fb.torch
    n = fb.world
    ctxs = tpt.comm_init_all(n)
    pts, dss = [], []
    for r, c in enumerate(ctxs):
        ds = tpt.DeviceScene(c, fb.scene)
        tpt.BVH(ds).construct(stats=False)
        pt = tpt.PathTracer(c, fb.W, fb.H, env_rgba8=fb.env)
        pts.append(pt)
        dss.append(ds)
    ms = None
    for it in range(3):
        for r, (pt, ds) in enumerate(zip(pts, dss)):
            pt.m_radiance.clear()
            pt.render_into(ds, fb.cam, spp=fb.per_slice * len(my_slices(r, n)), max_depth=fb.depth, seed=fb.seed, sample_offset=r, sample_stride=n)
        for c in ctxs:
            c.synchronize()
        t0 = time.perf_counter()
        tpt.accum_allreduce([pt.m_radiance for pt in pts])
        for c in ctxs:
            c.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
    shas = {hashlib.sha256(pt.m_radiance.to_i64().tobytes()).hexdigest() for pt in pts}
    for c in ctxs:
        c.close()
    assert len(shas) == 1, "tpt_accum_allreduce left different buffers on different GPUs"
    return shas.pop(), ms


def run_ours(args):
    import torch
    import torch.distributed as dist
    from tinypathtracer_b200.scene_io import scene_upload_bytes

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        host_group = dist.new_group(backend="gloo")
        global HOST_GROUP
        HOST_GROUP = host_group
    fb = FrameBench(args, args.workload, args.spp)
    W, H, spp, depth = fb.W, fb.H, fb.spp, fb.depth
    ctx = fb.ctx

    # ---- value: device-resident inputs -------------------------------------------------------------------
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    fb.prime()
    for i in range(args.warmup):
        fb.frame(i)
    fb.reset_counters()
    clocks.mark()
    ms_step = fb.timed(lambda i: fb.frame(i), args.steps) / args.steps
    host_ms_value = fb.host_ms
    clk = clocks.stop() if rank == 0 else {}
    ctr = fb.counters_sum()
    launches = ctr["kernel_launches"]
    if ctr["queue_overflow"] or ctr["dropped_nonfinite"] > 1e-6 * max(ctr["samples"], 1):
        raise SystemExit(f"invalid run: queue_overflow={ctr['queue_overflow']} dropped_nonfinite={ctr['dropped_nonfinite']}")
    samples_step = W * H * spp                       # whole job, all ranks
    sums = torch.tensor([fb.rays_traversed(ctr), ctr["rays_queued"]], device="cuda", dtype=torch.int64)
    if world > 1:
        dist.all_reduce(sums)
    rays_step, queued_step = int(sums[0].item()) / args.steps, int(sums[1].item()) / args.steps
    value = samples_step / (ms_step * 1e-3) / 1e6
    # one frame at a time on lane 0, for the record: what a single frame takes when nothing overlaps it
    ms_serial = fb.timed(lambda i: fb.frame(0), max(3, args.steps // 2)) / max(3, args.steps // 2)

    # ---- e2e: host arrays in, host framebuffer out, through the C ABI --------------------------------------
    ms_e2e = fb.run_e2e(args.steps)
    host_ms_e2e = fb.host_ms
    e2e_value = samples_step / (ms_e2e * 1e-3) / 1e6

    # ---- determinism evidence: SHA-256 of the all-reduced int64 accumulation buffer of one more frame (untimed)
    accum_sha = fb.accum_sha()
    exchange_ok = fb.exchange_check()
    exchange_alone = fb.exchange_alone()

    line = None
    if rank == 0:
        # ---- roofline pass (untimed for `value`): per-stage CUDA-event times + node/triangle counters --------
        tpt = fb.tpt
        torch.cuda.synchronize()
        ctx.reset_counters()
        bstats = tpt.BVH(fb.dscene).construct(stats=True)
        fb.acc.clear()
        rank_spp = fb.per_slice * len(fb.slices)
        fb.pt.render_into(fb.dscene, fb.cam, spp=rank_spp, max_depth=depth, seed=fb.seed, sample_offset=rank, sample_stride=world, accum=fb.acc,
                          stage_times=True, paths_in_flight=args.paths_in_flight)
        st = ctx.stage_times()
        c1 = ctx.counters()
        ctx.reset_counters()
        prof_spp = max(1, min(rank_spp, 4))
        fb.pt.render_into(fb.dscene, fb.cam, spp=prof_spp, max_depth=depth, seed=fb.seed, accum=fb.acc, counters=True, paths_in_flight=args.paths_in_flight)
        c2 = ctx.counters()
        ctx.synchronize()
        n_all, n_cnt = fb.rays_traversed(c1), fb.rays_traversed(c2)
        n_int, n_leaf = c2["node_visits"] / max(n_cnt, 1), c2["tri_tests"] / max(n_cnt, 1)
        trav_ms = st["primary"]["ms"] + st["wave"]["ms"]
        trav_launches = st["primary"]["launches"] + st["wave"]["launches"]
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        # The BVH of a bundled scene (<= 250 KB) lives in L1/L2: node and triangle fetches never reach HBM.  What does is the
        # wavefront records (88 B per queued ray, written once and read once) and the accumulator (24 B per pixel
        # and accumulation pass, 16 B per pixel for the resolve): THAT is the HBM figure of this workload.  The limiter is
        # instruction issue, reported from the committed ncu captures of this build's two kernels.
        batches = max(1, st["primary"]["launches"])
        hbm_bytes = 88.0 * queued_step / max(world, 1) + W * H * (24.0 * (1 + batches) + 16.0)
        hbm_ach = hbm_bytes / (ms_step * 1e-3) / 1e9
        caps = {k: _ncu_capture(fb.fixture, k) for k in ("k_primary", "k_wave")}
        issue = None
        if all(caps.values()):
            w = {"k_primary": st["primary"]["ms"], "k_wave": st["wave"]["ms"]}
            tot = sum(w.values()) or 1.0
            util = sum(w[k] * caps[k]["issue_active_pct"] * caps[k]["lanes_per_inst"] / 32.0 for k in caps) / tot
            issue = {"thread_issue_utilisation_pct": round(util, 2),
                     "formula": "sum over kernels of stage_ms x smsp__issue_active% x (thread instructions per warp instruction / 32) / sum stage_ms",
                     "kernels": {k: {"issue_active_pct": round(caps[k]["issue_active_pct"], 1), "lanes_per_inst": round(caps[k]["lanes_per_inst"], 1),
                                     "dram_bytes": round(caps[k]["dram_bytes"]), "capture_ms": round(caps[k]["duration_ms"], 3), "file": caps[k]["file"]}
                                 for k in caps}}
        traffic = round(sum(caps[k]["dram_bytes"] for k in caps)) if all(caps.values()) else None
        roofline = {"bound": "issue" if issue else "hbm", "kernel": "k_primary + k_wave (BVH traversal fused with shading)",
                    "achieved": issue["thread_issue_utilisation_pct"] if issue else round(hbm_ach, 2),
                    "peak": 100.0 if issue else peak, "unit": "% of peak thread-instruction issue (warp issue slots x 32 lanes)" if issue else "GB/s",
                    "frac": round(issue["thread_issue_utilisation_pct"] / 100.0, 4) if issue else round(hbm_ach / peak, 5),
                    "traffic": traffic,
                    "traffic_of": "dram__bytes_read + write of one k_primary launch + one k_wave (wave 1) launch = one batch, from the captures named under issue.kernels"
                    if traffic else None,
                    "issue": issue,
                    "hbm": {"achieved": round(hbm_ach, 2), "peak": peak, "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback (B200_PROFILING.md)",
                            "unit": "GB/s", "frac": round(hbm_ach / peak, 5), "algorithmic_bytes_per_step": round(hbm_bytes),
                            "what": "88 B per queued ray (record written once, read once) + accumulator clear / accumulate / resolve; node and "
                                    "triangle fetches are served by L1/L2 (the tree is %d KB)" % ((bstats["n_faces"] * 112) // 1024)},
                    "traversal_accounting": {"bytes_per_ray": round(56.0 * n_int + 40.0 * n_leaf + 88.0, 1), "n_int_per_ray": round(n_int, 2),
                                             "n_leaf_per_ray": round(n_leaf, 2),
                                             "cache_level_gbs": round((56.0 * n_int + 40.0 * n_leaf + 88.0) * n_all / (trav_ms * 1e-3) / 1e9, 1) if trav_ms > 0 else None,
                                             "note": "SURVEY §8(d)'s per-ray formula; these bytes are L1/L2 traffic here, not HBM, and are not a roofline fraction"},
                    "rays_per_launch": round(n_all / max(trav_launches, 1), 1), "avg_launch_us": round(1e3 * trav_ms / max(trav_launches, 1), 2),
                    "share_of_step": round(trav_ms / ms_step, 4), "stage_ms": {k: round(v["ms"], 3) for k, v in st.items()},
                    "stage_launches": {k: v["launches"] for k, v in st.items()},
                    "rays": {k: c1[k] for k in ("rays_primary", "rays_culled", "rays_candidates", "rays_extension", "rays_shadow", "rays_direct", "rays_queued")}}
        pool_gb = 2 * 48 * (c1["rays_queued"] and max(c1["rays_extension"], 1) or 1) / 1e9
        line = {
            "metric": "Msamples/s (whole frames: transform + BVH build + render + resolve)", "value": round(value, 3), "unit": "Msamples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": fb.workload_name, "parallelism": f"spp-slices x{world}", "slices": N_SLICES, "frames_in_flight": fb.n_lanes, "frames_in_flight_tuning": fb.tuning,
                       "exchange": {"none": "single GPU: tpt_resolve_bgra8", "nccl": "NCCL all-reduce of the int64 accumulators, then tpt_resolve_bgra8 on every rank",
                                    "peer": "tpt_exchange_resolve: fused reduce-scatter + resolve over NVLink peer mappings (CUDA IPC), pixels to rank 0"}[fb.exchange],
                       "l2": "inputs larger than L2: every step writes and reads %.1f GB of wavefront records (%d M queued rays x 96 B) through "
                             "HBM, far above the 126 MB L2; no flush between steps" % (96.0 * queued_step / max(world, 1) / 1e9, round(queued_step / 1e6))},
            "mrays_per_s": round(rays_step / (ms_step * 1e-3) / 1e6, 2), "rays_per_sample": round(rays_step / samples_step, 4),
            "rays_note": "rays actually traversed; the reference traverses every camera ray and every 'direct' ray (see reference arm's rays_per_sample)",
            "mrays_per_s_at_reference_ray_count": round(REFERENCE_RAYS_PER_SAMPLE[fb.fixture] * value, 2) if fb.fixture in REFERENCE_RAYS_PER_SAMPLE else None,
            "bvh_build_ms": round(bstats["build_ms"], 4), "bvh_faces": bstats["n_faces"],
            "host_enqueue_ms_per_step": {"value": round(host_ms_value, 3), "e2e": round(host_ms_e2e, 3)},   # rank 0's host time to enqueue one step
            "exchange_matches_allreduce": exchange_ok,   # peer exchange's framebuffer == all-reduce + resolve of the same frame, byte for byte (None: not used)
            "exchange_note": fb.exchange_note,
            "exchange_alone": exchange_alone,   # the exchange of one frame on idle GPUs (max over ranks, best of 5): peer kernel vs NCCL all-reduce + resolve
            "frame_latency_ms": round(ms_serial, 3),   # one frame alone on one lane (no overlap); ms_per_step = timed region / steps with frames_in_flight
            "e2e": {"value": round(e2e_value, 3), "unit": "Msamples/s", "ms_per_step": round(ms_e2e, 3),
                    "h2d_bytes_per_step": scene_upload_bytes(fb.scene), "d2h_bytes_per_step": W * H * 4,
                    "how": "tpt_scene_update (H2D of all scene arrays) + frame + D2H of the BGRA8 framebuffer into pinned memory; frames_in_flight "
                           "frames overlap, read-backs on a copy stream"},
            "gpu_launches": int(launches), "clocks": clk, "roofline": roofline,
            "accum_sha256": accum_sha,   # identical for 1, 2, 4 and 8 GPUs (exact integer accumulation + all-reduce)
        }
        del pool_gb
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, fb.fixture, W, H, depth, fb.env, fb.seed)
    # ---- N > 1: the library's own all-reduce over NVLink peers, driven by rank 0 alone ---------------------------
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier(group=host_group)
        if rank == 0:
            try:
                sha, ms = native_allreduce_check(fb)
                line["native_allreduce_sha256"] = sha
                line["native_allreduce_ms"] = round(ms, 3)
                line["native_allreduce_matches_nccl"] = bool(sha == accum_sha)
            except Exception as e:   # reported, never hidden
                line["native_allreduce_error"] = str(e)[:300]
        dist.barrier(group=host_group)
    # ---- the other BASELINE configs, a few steps each (one GPU only): driver-visible numbers for north_star's box / tir target
    if world == 1 and not args.no_also and args.workload == "ball" and not args.spp:
        also = {}
        fb.close()
        for name in ("box", "tir", "light"):
            try:
                o = FrameBench(args, name)
                ms, c = o.run_value(max(3, args.steps), args.warmup)
                ms_e = o.run_e2e(max(3, args.steps))
                n = o.W * o.H * o.spp
                also[name] = {"workload": o.workload_name, "ms_per_step": round(ms, 3), "value": round(n / (ms * 1e-3) / 1e6, 3), "unit": "Msamples/s",
                              "e2e_ms_per_step": round(ms_e, 3), "e2e_value": round(n / (ms_e * 1e-3) / 1e6, 3),
                              "mrays_per_s": round(o.rays_traversed(c) / max(3, args.steps) / (ms * 1e-3) / 1e6, 2),
                              "mrays_per_s_at_reference_ray_count": round(REFERENCE_RAYS_PER_SAMPLE.get(name, 0.0) * n / (ms * 1e-3) / 1e6, 2),
                              "accum_sha256": o.accum_sha()}
                o.close()
            except Exception as e:
                also[name] = {"error": str(e)[:300]}
        # configs[1] with the alias-table environment sampling north_star names (TPT_RENDER_ENV_NEE; off in the headline
        # because the reference never samples its CDF): same frame, one extra shadow ray per diffuse vertex
        try:
            o = FrameBench(args, "ball")
            ms, c = o.run_value(max(3, args.steps // 2), 2, flags_nee=True)
            also["ball_env_nee"] = {"workload": o.workload_name + " TPT_RENDER_ENV_NEE", "ms_per_step": round(ms, 3),
                                    "value": round(o.W * o.H * o.spp / (ms * 1e-3) / 1e6, 3), "unit": "Msamples/s"}
            o.close()
        except Exception as e:
            also["ball_env_nee"] = {"error": str(e)[:300]}
        if not args.no_also_soup:
            try:
                also["soup"] = soup_measure(args, small=False, steps=3, warmup=1, with_e2e=False)
            except Exception as e:
                also["soup"] = {"error": str(e)[:300]}
        line["also"] = also
    if world > 1:
        fb.close()                                                   # raises if the peer exchange ever timed out
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line))


def run_soup(args):
    print(json.dumps(soup_measure(args, small=args.soup_small, steps=args.steps, warmup=args.warmup, with_e2e=True)))


def soup_measure(args, small, steps, warmup, with_e2e):
    """BASELINE.json configs[3]: synthetic 10 000 004-triangle instanced soup — BVH build time and primary / secondary
    ray traversal throughput through tpt_trace_rays (device buffers for `value`, host buffers for `e2e`)."""
    import torch
    import tinypathtracer_b200 as tpt
    from tinypathtracer_b200.api import primary_rays_device
    from tinypathtracer_b200.soup import make_soup, secondary_rays

    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    nb, nx = (8192, 3211) if not small else (64, 25)
    scene = make_soup(load_scene("ball"), nb, nx)
    F = scene["indices"].size // 3
    W, H = 3840, 2160
    ctx = tpt.Context(local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    dscene = tpt.DeviceScene(ctx, scene)
    bvh = tpt.BVH(dscene)
    cam = tpt.Camera.from_scene(scene)
    n = W * H
    rays_t = torch.empty((n, 6), dtype=torch.float32, device="cuda")
    hits_t = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    # ---- BVH build: median of 5 after warm-up (CUDA events inside tpt_bvh_build)
    for _ in range(max(1, warmup)):
        bvh.construct(stats=True)
    builds = sorted(bvh.construct(stats=True)["build_ms"] for _ in range(5))
    build_ms = builds[2]
    bstats = bvh.stats
    # refit (tpt_bvh_refit: same matrices pushed again, hierarchy kept): transform + leaf records + boxes only
    dscene.set_transforms(scene["vert_trans"], scene["normal_trans"])
    bvh.refit(stats=True)
    refit_ms = sorted(bvh.refit(stats=True)["build_ms"] for _ in range(5))[2]
    # ---- ray sets: primary = pixel centres; secondary = one cosine ray per primary hit (seeded, host-generated once)
    primary_rays_device(ctx, cam, W, H, rays_t.data_ptr())
    bvh.trace_device(rays_t.data_ptr(), n, hits_t.data_ptr())
    torch.cuda.synchronize()
    rays_h = rays_t.cpu().numpy().view(tpt.RAY_DTYPE).reshape(-1)
    hits_h = hits_t.cpu().numpy().view(tpt.HIT_DTYPE).reshape(-1)
    wv, _ = dscene.world()
    sec_h = secondary_rays(rays_h, hits_h, wv, scene["indices"])
    ns = sec_h.shape[0]
    sec_t = torch.from_numpy(sec_h.view(np.float32).reshape(-1, 6)).cuda()
    sec_hits_t = torch.empty((ns, 4), dtype=torch.float32, device="cuda")

    def timed(fn, k):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(k):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / k

    def step():   # one pass of the hot path over the batch: both ray sets
        bvh.trace_device(rays_t.data_ptr(), n, hits_t.data_ptr())
        bvh.trace_device(sec_t.data_ptr(), ns, sec_hits_t.data_ptr())

    for _ in range(warmup):
        step()
    ctx.reset_counters()
    clocks = ClockSampler(local)
    clocks.start()
    ms_step = timed(step, steps)
    launches = ctx.counters()["kernel_launches"]
    timed(step, max(steps, 40))            # keep the GPU under the same load long enough for nvidia-smi's 100 ms sampler
    clk = clocks.stop()
    ms_pri = timed(lambda: bvh.trace_device(rays_t.data_ptr(), n, hits_t.data_ptr()), steps)
    ms_sec = timed(lambda: bvh.trace_device(sec_t.data_ptr(), ns, sec_hits_t.data_ptr()), steps)
    import hashlib
    torch.cuda.synchronize()
    hits_sha = hashlib.sha256(hits_t.cpu().numpy().tobytes() + sec_hits_t.cpu().numpy().tobytes()).hexdigest()
    e2e = None
    if with_e2e:
        # ---- e2e: host ray buffers in, host hits out, through tpt_trace_rays(device_ptrs = 0)
        def pinned(a):   # host staging in page-locked memory, as a caller who cares about PCIe throughput would hold it
            t = torch.from_numpy(a.view(np.float32).reshape(a.shape[0], -1).copy()).pin_memory()
            return t, t.numpy().view(a.dtype).reshape(-1)
        _k1, rays_p = pinned(rays_h)
        _k2, sec_p = pinned(sec_h)
        _k3, hits_p = pinned(np.zeros(n, tpt.HIT_DTYPE))
        _k4, shits_p = pinned(np.zeros(ns, tpt.HIT_DTYPE))
        bvh.trace(rays_p, out=hits_p)
        reps = max(1, steps // 2)
        t0 = time.perf_counter()
        for _ in range(reps):
            bvh.trace(rays_p, out=hits_p)
            bvh.trace(sec_p, out=shits_p)
        e2e_ms = (time.perf_counter() - t0) * 1e3 / reps
        assert np.array_equal(hits_p["prim"], hits_h["prim"])
        e2e = {"value": round((n + ns) / (e2e_ms * 1e-3) / 1e6, 2), "unit": "Mrays/s", "ms_per_step": round(e2e_ms, 3),
               "h2d_bytes_per_step": int((n + ns) * 24), "d2h_bytes_per_step": int((n + ns) * 16)}
    # ---- counted pass for the roofline: node / triangle visits per ray on the same tree
    ctx.reset_counters()
    bvh.trace_device(rays_t.data_ptr(), n, hits_t.data_ptr(), counted=True)
    c_pri = ctx.counters()
    ctx.reset_counters()
    bvh.trace_device(sec_t.data_ptr(), ns, sec_hits_t.data_ptr(), counted=True)
    c_sec = ctx.counters()
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # 56 B per internal node whose children were tested (BINARY nodes; a 4-wide visit tests four children and is counted
    # as two binary visits' worth of boxes: 2 x 48 B + 16 B of child indices = 112 B), 40 B per triangle tested, 88 B of record traffic
    wide = os.environ.get("TPT_NO_WIDE") is None and "TPT_WIDE_MIN_FACES" in os.environ and F >= int(os.environ["TPT_WIDE_MIN_FACES"])
    node_bytes = 112.0 if wide else 56.0
    bpr_sec = node_bytes * c_sec["node_visits"] / ns + 40.0 * c_sec["tri_tests"] / ns + 88.0
    bpr_pri = node_bytes * c_pri["node_visits"] / n + 40.0 * c_pri["tri_tests"] / n + 88.0
    ach = bpr_sec * ns / (ms_sec * 1e-3) / 1e9
    build_gbs = 408.0 * F / (build_ms * 1e-3) / 1e9
    # DRAM traffic and L1 pressure of the secondary-ray launch from the committed ncu --set full capture (full-size soup only)
    cap = None if small else _ncu_capture("soup", "k_trace_rays")
    line = {
        "metric": "Mrays/s (tpt_trace_rays, primary + secondary ray sets; BVH build ms beside it)", "value": round((n + ns) / (ms_step * 1e-3) / 1e6, 2),
        "unit": "Mrays/s", "n_gpus": 1, "steps": steps, "warmup": warmup, "ms_per_step": round(ms_step, 3), "higher_is_better": True,
        "scaling": "replicas only", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"synthetic soup {F} triangles ({nb} spheres x 1216 + {nx} boxes x 12), {W}x{H} pixel-centre primary rays + {ns} cosine secondary rays",
                   "l2": "inputs larger than L2: BVH (640 MB nodes, 480 MB triangles) and ray buffers exceed the 126 MB L2",
                   "nodes": "4-wide compressed (64 B, 8-bit child boxes)" if wide else "binary (64 B, fp32 child boxes)"},
        "bvh_build_ms": round(build_ms, 4), "bvh_build_ms_all": [round(b, 4) for b in builds], "bvh_faces": int(F), "bvh_max_depth": bstats["max_depth"],
        "bvh_refit_ms": round(refit_ms, 4), "bvh_sah_cost": round(bstats["sah_cost"], 2),
        "primary_mrays_per_s": round(n / (ms_pri * 1e-3) / 1e6, 2), "secondary_mrays_per_s": round(ns / (ms_sec * 1e-3) / 1e6, 2),
        "primary_hit_fraction": round(float((hits_h["prim"] >= 0).mean()), 4), "hits_sha256": hits_sha,
        "gpu_launches": int(launches), "clocks": clk,
        "roofline": {"bound": "hbm", "kernel": "k_trace_rays_persistent (secondary ray set, incoherent)", "achieved": round(ach, 2), "peak": peak,
                     "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback (B200_PROFILING.md)", "unit": "GB/s", "frac": round(ach / peak, 5),
                     "traffic": round(cap["dram_bytes"]) if cap else None,
                     "traffic_ratio": round(cap["dram_bytes"] / (bpr_sec * ns), 3) if cap else None,
                     "traffic_of": f"dram bytes of the secondary-ray launch, ncu --set full, {cap['file']}; traffic_ratio = measured DRAM bytes / algorithmic bytes "
                                   "(below 1: most node fetches are served by L1/L2; the kernel's limiter is the L1 data pipe, l1_wavefront_pct)" if cap else None,
                     "l1_wavefront_pct": cap["l1_wavefront_pct"] if cap else None,
                     "bytes_per_ray": round(bpr_sec, 1), "n_int_per_ray": round(c_sec["node_visits"] / ns, 2),
                     "n_leaf_per_ray": round(c_sec["tri_tests"] / ns, 2), "rays_per_launch": ns, "avg_launch_us": round(ms_sec * 1e3, 1),
                     "primary": {"bytes_per_ray": round(bpr_pri, 1), "n_int_per_ray": round(c_pri["node_visits"] / n, 2),
                                 "n_leaf_per_ray": round(c_pri["tri_tests"] / n, 2), "achieved": round(bpr_pri * n / (ms_pri * 1e-3) / 1e9, 2)},
                     "bvh_build": {"algorithmic_bytes_per_triangle": 408, "achieved": round(build_gbs, 2), "frac": round(build_gbs / peak, 5)}},
    }
    if e2e:
        line["e2e"] = e2e
    dscene.close()
    ctx.close()
    return line


def cpu_baseline(args, fixture, W, H, depth, env, seed):
    """The oracle (a port: the reference has no CPU path) on a bounded sample of the same workload."""
    from oracle import gltf_scene, pyoracle as orc
    scene = gltf_scene.load_scene(ROOT / "tests" / "golden" / "scenes" / f"{fixture}.npz")
    wv, wn = orc.transform(scene)
    nodes, _, _ = orc.build_reference_bvh(wv, scene["indices"])
    # centre crop, 4 spp; grow until ~10 s of CPU work
    cw, ch, spp = max(32, W // 8), max(32, H // 8), 4
    x0, y0 = (W - cw) // 2, (H - ch) // 2
    t_used, n_samples, n_rays = 0.0, 0, 0
    while t_used < args.cpu_seconds:
        t0 = time.perf_counter()
        _, cnt = orc.render(scene, W, H, spp, max_depth=depth, seed=seed, env=env, region=(x0, y0, x0 + cw, y0 + ch),
                            prepared=(wv, wn, nodes), counters=True, sample_offset=n_samples // (cw * ch))
        t_used += time.perf_counter() - t0
        n_samples += cw * ch * spp
        n_rays += int(cnt[0])
    return {"value": round(n_samples / t_used / 1e6, 4), "unit": "Msamples/s", "cores": orc.num_threads(), "kind": "port",
            "mrays_per_s": round(n_rays / t_used / 1e6, 3),
            "sample": f"centre {cw}x{ch} crop of the {W}x{H} frame, {n_samples // (cw * ch)} spp, {t_used:.1f} s of CPU work"}


def run_reference(args):
    """Reference arm.  The reference has NO CPU implementation of this path; its own code is CUDA.
    When oracle/_ref/tpt_ref_headless (the UNMODIFIED reference sources built headless for sm_100 by
    oracle/ref_build.py) is present it is run on GPU 0 for the same workload; otherwise the CPU oracle port
    is timed on all host cores on a bounded sample."""
    rank, world, local = dist_env()
    if rank != 0:
        return
    if args.workload == "soup":
        print(json.dumps(run_reference_soup(args)))
        return
    line = reference_frames(args, args.workload, args.spp, args.steps, args.warmup)
    # the other BASELINE configs with the same binary, so that the driver's record holds both arms of each
    if args.workload == "ball" and not args.spp and not args.no_also:
        also = {}
        for name in ("box", "tir", "light"):
            try:
                o = reference_frames(args, name, 0, max(3, args.steps), args.warmup)
                also[name] = {k: o[k] for k in ("ms_per_step", "value", "unit", "rays_per_sample", "mrays_per_s", "reference_kind") if k in o}
                also[name]["workload"] = o["config"]["workload"]
            except Exception as e:
                also[name] = {"error": str(e)[:300]}
        if not args.no_also_soup:
            try:
                o = run_reference_soup(args)
                also["soup"] = {k: o[k] for k in ("unavailable", "value", "unit", "ms_per_step", "bvh_build_ms", "primary_mrays_per_s", "secondary_mrays_per_s",
                                                  "reference_kind") if k in o}
            except Exception as e:
                also["soup"] = {"error": str(e)[:300]}
        line["also"] = also
    print(json.dumps(line))


def reference_frames(args, workload, spp_override, steps, warmup):
    fixture, W, H, spp, depth, envkind, seed = WORKLOADS[workload]
    if spp_override:
        spp = spp_override
    cfg = {"workload": f"{fixture}.gltf {W}x{H} {spp}spp depth{depth} env={envkind or 'black'} seed={seed}"}
    exe = ROOT / "oracle" / "_ref" / "tpt_ref_headless"
    ref_scene = ROOT / "oracle" / "_ref" / "input" / f"{fixture}.gltf"
    if exe.exists() and ref_scene.exists() and not args.ref_cpu:
        import torch
        if torch.cuda.is_available():
            out = run_reference_cuda(args, exe, ref_scene, W, H, spp, depth, envkind, cfg, steps, warmup)
            if out is not None:
                return out
    env = synthetic_sky() if envkind == "sky" else None
    res = []
    for _ in range(warmup + steps):
        res.append(cpu_baseline(args, fixture, W, H, depth, env, seed))
    res = res[warmup:]
    v = float(np.mean([r["value"] for r in res]))
    cb = dict(res[-1])
    cb["value"] = round(v, 4)
    return {"impl": "reference", "metric": "Msamples/s (whole frames: transform + BVH build + render + resolve)",
            "value": round(v, 4), "unit": "Msamples/s", "n_gpus": 1, "steps": steps, "warmup": warmup,
            "ms_per_step": round(1e3 * W * H * spp / (v * 1e6), 1), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg, "cpu_baseline": cb,
            "reference_kind": "cpu oracle port (reference CUDA build not present)",
            "e2e": {"value": round(v, 4), "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def _world_vertices_numpy(scene):
    """object -> world vertices with the reference's operation order in float32 (bit-identical to its transform kernel)."""
    idx = np.asarray(scene["indices"]).reshape(-1, 3)
    face_obj = np.searchsorted(scene["objects"]["begin"], np.arange(idx.shape[0]), side="right") - 1
    vobj = np.zeros(scene["vertices"].shape[0], np.int64)
    for k in range(3):
        vobj[idx[:, k]] = face_obj
    m = np.asarray(scene["vert_trans"], np.float32)[vobj]
    v = np.asarray(scene["vertices"], np.float32)
    out = np.empty_like(v)
    for i in range(3):
        out[:, i] = ((m[:, i] * v[:, 0] + m[:, 4 + i] * v[:, 1]) + m[:, 8 + i] * v[:, 2]) + m[:, 12 + i]
    return out


def _reference_morton_cells(c):
    """floatTo21Int of the reference (src/bvh.cu:23-46), vectorised: the 21-bit cell of one coordinate."""
    ix = np.ascontiguousarray(c, np.float32).view(np.int32).astype(np.int64)
    exponent = ((ix >> 23) & 0xff) - 127
    mant = (ix & 0x00ffffff) | 0x00800000
    sign = np.where(ix < 0, -1, 1)
    sh = np.clip(np.abs(exponent), 0, 31)
    value = np.where(exponent >= 8, 2**31 - 1, np.where(exponent >= 0, mant << sh, mant >> sh))
    value = (value * sign + (2**31 - 1)) & 0xffffffff
    return (value & 0xfffff800) >> 11


def run_reference_soup(args):
    """configs[3] through the reference's own BVH::construct and traverseBVH (oracle/_ref/tpt_ref_headless --soup),
    on the same world-space triangles and the same two ray files as the product arm."""
    import tempfile
    from oracle import pyoracle as orc
    from tinypathtracer_b200.soup import make_soup, secondary_rays
    from tinypathtracer_b200.scene_io import load_scene_npz
    exe = ROOT / "oracle" / "_ref" / "tpt_ref_headless"
    if not exe.exists():
        return {"impl": "reference", "unavailable": "oracle/_ref/tpt_ref_headless is not built (needs /root/reference at build time)"}
    nb, nx = (8192, 3211) if not args.soup_small else (64, 25)
    scene = make_soup(load_scene_npz(ROOT / "tests" / "golden" / "scenes" / "ball.npz"), nb, nx)
    F = scene["indices"].size // 3
    W, H = 3840, 2160
    wv = _world_vertices_numpy(scene)
    # the reference's hierarchy breaks (and its refit can spin forever) on runs of >= 3 equal Morton keys: check first
    tri = wv[np.asarray(scene["indices"]).reshape(-1, 3)]
    ctr = (tri.min(1) + tri.max(1)) * np.float32(0.5)
    cells = (_reference_morton_cells(ctr[:, 0]) | (_reference_morton_cells(ctr[:, 1]) << 21) | (_reference_morton_cells(ctr[:, 2]) << 42))
    _, counts = np.unique(cells, return_counts=True)
    if counts.max() >= 3:
        return {"impl": "reference", "unavailable": f"soup has a run of {int(counts.max())} equal reference Morton keys: its builder would hang"}
    rays = orc.primary_rays(scene, W, H).astype(np.float32).reshape(-1, 6)
    tmp = Path(tempfile.mkdtemp(prefix="tpt_soup_"))

    def write(path, rb):
        with open(path, "wb") as f:
            np.array([wv.shape[0], F, rays.shape[0], rb.shape[0]], np.uint64).tofile(f)
            wv.tofile(f)
            np.asarray(scene["indices"], np.uint32).tofile(f)
            rays.tofile(f)
            rb.tofile(f)

    def run(path, out, builds, reps):
        cmd = [str(exe), "--soup", str(path), "--soup-builds", str(builds), "--soup-reps", str(reps)] + (["--out", str(out)] if out else [])
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=args.ref_timeout)
        return json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])

    try:
        # pass 1: primary rays only -> the reference's own hits -> the seeded secondary ray file
        write(tmp / "a.bin", np.zeros((0, 6), np.float32))
        run(tmp / "a.bin", tmp / "a.hits", 1, 1)
        hs = np.fromfile(tmp / "a.hits", np.dtype([("prim", "<i4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4")]))
        rs = rays.view(np.dtype([("origin", "<f4", (3,)), ("dir", "<f4", (3,))])).reshape(-1)
        sec = secondary_rays(rs, hs, wv, scene["indices"]).view(np.float32).reshape(-1, 6)
        write(tmp / "b.bin", sec)
        d = run(tmp / "b.bin", None, max(1, min(args.steps, 3)), max(1, min(args.steps, 3)))
    except Exception as e:
        return {"impl": "reference", "unavailable": f"reference soup run failed: {e}"}
    finally:
        import shutil
        shutil.rmtree(tmp, ignore_errors=True)
    n, ns = d["primary_rays"], d["secondary_rays"]
    ms = d["primary_ms"] + d["secondary_ms"]
    v = (n + ns) / (ms * 1e-3) / 1e6
    return ({"impl": "reference", "metric": "Mrays/s (tpt_trace_rays, primary + secondary ray sets; BVH build ms beside it)", "value": round(v, 3),
                      "unit": "Mrays/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
                      "scaling": "replicas only", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": f"synthetic soup {F} triangles, {W}x{H} pixel-centre primary rays + {ns} cosine secondary rays"},
                      "bvh_build_ms": d["bvh_build_ms"], "primary_mrays_per_s": round(n / d["primary_ms"] / 1e3, 3),
                      "secondary_mrays_per_s": round(ns / d["secondary_ms"] / 1e3, 3), "primary_hit_fraction": round(float((hs["prim"] >= 0).mean()), 4),
                      "reference_kind": f"reference BVH::construct + traverseBVH, headless, {d['build']}, on GPU 0",
                      "cpu_baseline": {"value": round(v, 3), "unit": "Mrays/s", "cores": 0, "kind": "reference", "sample": "whole workload on the GPU (the reference has no CPU path)"},
                      "e2e": {"value": round(v, 3), "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})


def run_reference_cuda(args, exe, ref_scene, W, H, spp, depth, envkind, cfg, steps, warmup):
    """Runs the headless build of the reference's own CUDA sources; it prints one JSON object with frame times."""
    # the reference cannot run without an env texture (uninitialised handle): black env = a black PNG
    env_png = ROOT / "oracle" / "_ref" / ("sky.png" if envkind == "sky" else "black.png")
    if depth == 32:
        exe = exe.with_name("tpt_ref_headless_d32")     # DEPTH_TRACE is a compile-time constant of the reference
    base = ["--scene", str(ref_scene), "--width", str(W), "--height", str(H), "--spp", str(spp), "--depth", str(depth)]
    if env_png.exists():
        base += ["--env", str(env_png)]
    try:
        r = subprocess.run([str(exe), *base, "--frames", str(steps), "--warmup", str(warmup)], capture_output=True, text=True, timeout=args.ref_timeout)
        data = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    except Exception as e:   # the binary is a build of foreign code: fall back to the port, loudly
        sys.stderr.write(f"reference CUDA build failed to run ({e}); falling back to the CPU oracle port\n")
        return None
    ms = float(data["ms_per_frame"])
    v = W * H * spp / (ms * 1e-3) / 1e6
    out = {"impl": "reference", "metric": "Msamples/s (whole frames: transform + BVH build + render + resolve)", "value": round(v, 3),
           "unit": "Msamples/s", "n_gpus": 1, "steps": steps, "warmup": warmup, "ms_per_step": round(ms, 3),
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
           "reference_kind": f"reference CUDA sources, headless, -O3 sm_100 ({data.get('build', '')}) on GPU 0; depth is compile-time {data.get('depth')}",
           "cpu_baseline": {"value": round(v, 3), "unit": "Msamples/s", "cores": 0, "kind": "reference",
                            "sample": "whole workload on the GPU (the reference has no CPU path)"},
           "e2e": {"value": round(v, 3), "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    # rays per sample from the COUNTING build of the same sources (one atomicAdd per traverseBVH call; one untimed frame), so that
    # north_star's Mrays/s wording can be answered: Mrays/s = rays per sample x Msamples/s of the timed (uncounted) build
    cexe = exe.with_name("tpt_ref_headless_count_d32" if depth == 32 else "tpt_ref_headless_count")
    if cexe.exists():
        try:
            r = subprocess.run([str(cexe), *base, "--frames", "1", "--warmup", "0"], capture_output=True, text=True, timeout=args.ref_timeout)
            c = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
            rps = float(c["rays_all_frames"]) / (W * H * spp)
            out["rays_per_sample"] = round(rps, 4)
            out["mrays_per_s"] = round(rps * v, 2)
            out["rays_note"] = "traverseBVH calls per camera sample counted by oracle/_ref/" + cexe.name + " (one frame); mrays_per_s = rays_per_sample x value"
        except Exception as e:
            out["rays_note"] = f"counting build failed: {e}"
    return out


def _json_only_stdout():
    """The contract is ONE JSON line on stdout, but native libraries write there too (NCCL prints its version banner on fd 1 at
    any NCCL_DEBUG level above NONE).  File descriptor 1 is pointed at stderr for the whole run; Python's sys.stdout — where the
    JSON line is printed — keeps the original stdout."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)


def main():
    _json_only_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="ball", choices=sorted(WORKLOADS) + ["soup"])
    ap.add_argument("--soup-small", action="store_true", help="soup workload at 1/128 size (smoke test)")
    ap.add_argument("--spp", type=int, default=0, help="override the workload's spp (multiple of 8)")
    ap.add_argument("--paths-in-flight", type=int, default=0)
    ap.add_argument("--frames-in-flight", type=int, default=0,
                    help="independent frames kept in flight (each on its own stream, scene, pools and accumulator); 0 = auto: 2 or 4, whichever a few untimed frames show to be faster")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="how the ranks' accumulators meet at world > 1: the library's fused peer-memory kernel, or torch.distributed all-reduce")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the other BASELINE configs (box, tir, light, NEE, soup) after the headline")
    ap.add_argument("--no-also-soup", action="store_true", help="skip only the 10M-triangle soup in the `also` section")
    ap.add_argument("--ref-cpu", action="store_true", help="reference arm: force the CPU oracle port")
    ap.add_argument("--ref-timeout", type=float, default=900.0)
    args = ap.parse_args()
    if args.workload == "soup" and args.impl != "reference":
        run_soup(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
