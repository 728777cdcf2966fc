#!/usr/bin/env python
"""bench.py — rays/s and ms/frame of the voxel ray-march hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nproc-per-node N ... bench.py --gpus N ...     (one rank per GPU, NCCL)

A "step" is one frame: the default mod's scene at 1920x1080, samples=4, max_bounces=4, rough PBR + glass + mist
materials (BASELINE.json configs[2], the configuration the metric "1080p 4spp" is quoted on), cycling through
four camera poses.  One JSON line is printed by rank 0; beside the headline it carries a `configs` object with the two
other named shapes (C4: 512^3 terrain at 3840x2160 16 spp, C5: 9.9 M-voxel model at 1080p 8 spp with culling on), each
split over the N GPUs like the headline, and the CRC32 of every assembled frame with `identical_to_1gpu`.
See DESIGN.md "Measurement".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "rays_per_sec_1080p_4spp"
UNIT = "rays/s"

# C3: shipped config.cfg values except where BASELINE.json's config names others
C3_SETTINGS = dict(width=1920, height=1080, samples=4, static=True, shutter=0.25, falloff=0.25, dof=0.5, dist_min=0,
                   dist_max=192, max_light=1, max_bounces=4, lod_bounces=0.5, lod_samples=0, lod_random=0.25,
                   lod_edge=0, chunk_size=16, chunk_lod=2, fov=90, culling=False, threads=1)
# (position, euler degrees roll/yaw/pitch): default player camera + three poses framing the material cubes
POSES = [((-12.0, 4.0, 4.0), (0, 0, 0)), ((-30.0, 20.0, -10.0), (0, 20, 0)), ((-4.0, 18.0, -6.0), (0, 0, 0)),
         ((-36.0, 19.0, -60.0), (0, 0, 0))]
REFERENCE_STRIDE = 96          # --impl reference: every 96th pixel in x and y of the 1080p frame per step (880 rays)


def load_workload():
    from python_raytracer_b200.mathutil import euler_to_quaternion
    from python_raytracer_b200.world import chunk_presence, lod_table
    d = np.load(os.path.join(ROOT, "tests", "golden", "default_world.npz"))
    origin, cells, mats = d["origin"], d["cells"], d["materials"]
    s = C3_SETTINGS
    present = chunk_presence(cells, s["chunk_size"])
    poses = []
    for pos, eul in POSES:
        poses.append(dict(pos=pos, rot=euler_to_quaternion(*eul),
                          chunk_res=lod_table(origin, present, pos, s["chunk_size"], s["chunk_lod"], s["dist_max"])))
    return origin, cells, mats, poses


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed regions (B200_PROFILING.md).  NVML is polled from a thread
    every ~2 ms (an nvidia-smi -lms loop cannot deliver three samples inside an 8 ms timed region); nvidia-smi is the
    fallback.  Started before the warm-up; only samples that fall inside a marked region are reported."""
    REASONS = (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40))

    def __init__(self, index):
        self.samples, self.regions, self.stop_flag, self.how = [], [], False, None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = None
            bus = getattr(torch.cuda.get_device_properties(index), "pci_bus_id", None)    # CUDA_VISIBLE_DEVICES may renumber
            if bus is not None:
                for k in range(pynvml.nvmlDeviceGetCount()):
                    h = pynvml.nvmlDeviceGetHandleByIndex(k)
                    if int(pynvml.nvmlDeviceGetPciInfo(h).bus) == int(bus):
                        self.h = h
            if self.h is None:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            pynvml.nvmlDeviceGetClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.how = "nvml"
            self.t = threading.Thread(target=self._poll_nvml, daemon=True)
            self.t.start()
        except Exception:
            self.how = None
        if self.how is None:
            try:
                q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_power_cap,"
                     "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.hw_thermal_slowdown")
                self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                              "-lms", "10"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
                self.how = "nvidia-smi -lms 10"
                self.t = threading.Thread(target=self._pump_smi, daemon=True)
                self.t.start()
            except Exception:
                self.how = None

    def _poll_nvml(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    bits = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    bits = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.samples.append((time.perf_counter(), mhz, self.max_mhz, [n for n, b in self.REASONS if bits & b]))
            except Exception:
                pass
            time.sleep(0.002)

    def _pump_smi(self):
        for line in self.proc.stdout:
            f = [v.strip() for v in line.split(",")]
            try:
                self.samples.append((time.perf_counter(), float(f[0]), float(f[1]),
                                     [n for (n, _), v in zip(self.REASONS, f[2:6]) if v.lower().startswith("active")]))
            except Exception:
                continue

    def region(self, t0, t1):
        self.regions.append((t0, t1))

    def stop(self):
        self.stop_flag = True
        if self.how is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["no NVML / nvidia-smi"]}
        if self.how != "nvml":
            self.proc.terminate()
        time.sleep(0.01)
        inside = [s for s in self.samples if any(a <= s[0] <= b for a, b in self.regions)]
        sm = [s[1] for s in inside]
        reasons = sorted({r for s in inside for r in s[3]})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_min_mhz": min(sm) if sm else None,
                "sm_max_mhz": max(s[2] for s in self.samples) if self.samples else None, "samples": len(sm),
                "samples_total": len(self.samples), "reasons": reasons, "how": self.how + ", samples inside the timed regions only"}


# ---------------------------------------------------------------------------------------------------------------------
# CPU arms

def _strided_pixels(w, h, stride):
    xs, ys = np.meshgrid(np.arange(stride // 2, w, stride, dtype=np.int32), np.arange(stride // 2, h, stride, dtype=np.int32), indexing="ij")
    return np.stack([xs.ravel(), ys.ravel()], axis=1)


def oracle_rays_per_sec(sd, origin, cells, mats, poses, stride, steps, warmup, threads):
    """oracle/vt_oracle.c (C port of the reference trace loop, pinned bit-exact to it) on the host cores, on a strided
    pixel sample of the same frames.  Returns (rays/s, rays per step, seconds)."""
    from oracle import cport
    pixels = _strided_pixels(sd["width"], sd["height"], stride)
    lens = sd["fov"] * np.pi / 8
    total_rays, total_t = 0, 0.0
    for it in range(warmup + steps):
        p = poses[it % len(poses)]
        world = cport.World(origin, cells, p["chunk_res"], mats)
        t0 = time.perf_counter()
        o = cport.render(sd, p["pos"], p["rot"], lens, world, pixels=pixels, rng=cport.RNG_PHILOX, seed=0, threads=threads)
        dt = time.perf_counter() - t0
        if it >= warmup:
            total_rays += int(o["counters"][0]); total_t += dt
    return total_rays / total_t, total_rays / max(1, steps), total_t


def port_baseline(sd, origin, cells, mats, poses, stride, steps, what):
    cores = os.cpu_count() or 1
    v, rays_step, secs = oracle_rays_per_sec(sd, origin, cells, mats, poses, stride, steps, 1, cores)
    return {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "seconds": round(secs, 2),
            "sample": "%d frames, %s of %s (%d rays each), C port of the reference trace loop (oracle/vt_oracle.c), %d threads" % (
                steps, "every pixel" if stride == 1 else "every %s pixel in x and y" % _ordinal(stride), what, rays_step, cores)}


def _ordinal(n):
    return "%d%s" % (n, "th" if 10 <= n % 100 <= 20 else {1: "st", 2: "nd", 3: "rd"}.get(n % 10, "th"))


def reference_available():
    from oracle import refharness as rh
    return rh.available() or rh.available_bytecode()


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on the host cores.  When the reference is
    runnable here -- from /root/reference (build container) or from the bytecode build oracle/_ref/ that
    oracle/build_ref.py makes and that travels to the GPU box -- that is the UNMODIFIED Python reference (Camera.tile
    through a multiprocessing pool, like init.py:162, 193-205), on the same C3 workload, each step a strided pixel sample
    of one frame.  Otherwise its C port (the oracle) is timed.  Rank 0 only."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    frame_rays = C3_SETTINGS["width"] * C3_SETTINGS["height"] * C3_SETTINGS["samples"]
    if reference_available():
        from oracle import refbench
        r = refbench.C3Reference(C3_SETTINGS, POSES, REFERENCE_STRIDE, cores)
        for i in range(args.warmup):
            r.step(i)
        secs = sum(r.step(args.warmup + i) for i in range(args.steps))
        rays_step, kind = r.rays_per_step, "reference"
        r.close()
        v = rays_step * args.steps / secs
        sample = ("each step = every %s pixel in x and y of one 1920x1080x4spp frame (%d rays) through the reference's own "
                  "Camera.tile (unmodified Python, %s), multiprocessing pool of %d workers, pixel lists (x ^ y) %% workers; scene "
                  "build %.1f s excluded" % (_ordinal(REFERENCE_STRIDE), rays_step, r.source, cores, r.build_seconds))
    else:
        origin, cells, mats, poses = load_workload()
        v, rays_step, secs = oracle_rays_per_sec(C3_SETTINGS, origin, cells, mats, poses, 4, args.steps, args.warmup, cores)
        kind = "port"
        sample = "each step = every 4th pixel in x and y of one 1920x1080x4spp frame (%d rays), C port (oracle/vt_oracle.c), %d threads" % (rays_step, cores)
    line = {"metric": METRIC, "value": v, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * secs / args.steps, "ms_per_full_frame_est": 1000.0 * frame_rays / v,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args.gpus, args.gpus > 1),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(n_gpus, peer=False, **extra):
    c = {"workload": "default mod scene (castle + 6 material cubes + player), 1920x1080, samples=4, max_bounces=4, "
                     "dof=0.5, lod_random=0.25, lod_edge=0, shipped rough/glass/mist materials, 4 camera poses cycled "
                     "(BASELINE.json configs[2])",
         "rays_per_frame": C3_SETTINGS["width"] * C3_SETTINGS["height"] * C3_SETTINGS["samples"],
         "parallelism": "image split in interleaved 8-row stripes over %d GPU(s), scene replicated, %s" % (
             n_gpus, "pixels stored into rank 0's frame over NVLink peer memory by the render kernel + arrival flags (no collective)"
             if peer else ("NCCL gather to rank 0" if n_gpus > 1 else "single launch")),
         "l2": "flushed between frames (256 MiB memset); per-frame CUDA events"}
    c.update(extra)
    return c


# ---------------------------------------------------------------------------------------------------------------------
# GPU arm

class FramePath:
    """One rank's part of a frame + its assembly on rank 0, on either multi-GPU path (or the single-GPU one)."""

    def __init__(self, cam, rank, world, use_peer, visited=False):
        from python_raytracer_b200.parallel import FrameRenderer, PeerFrameRenderer
        self.cam, self.rank, self.world, self.use_peer, self.visited = cam, rank, world, use_peer, visited
        if use_peer:
            cam._prepare()
            self.fr = PeerFrameRenderer(cam, rank, world, visited_words=cam._visited.numel() if visited else 0)
        else:
            self.fr = FrameRenderer(cam, rank, world)
        self._b = None

    def render(self, e1=None):
        """Enqueue one frame (render + assembly); nothing is copied to the host."""
        if self.use_peer:
            if self._b is not None:
                self.fr.release(self._b)
            self._b = self.fr.render_frame(prepare=False, visited=self.visited)
            if e1 is not None:
                e1.record()
        else:
            self.fr.render(prepare=False, want=("rgba8", "visited") if self.visited else ("rgba8",))
            if e1 is not None:
                e1.record()
            self.fr.gather()

    def frame(self):
        """Rank 0: the assembled frame of the last render() as a device tensor."""
        if self.rank != 0:
            return None
        return self.fr.frame_views[self._b] if self.use_peer else self.fr.frame

    def union(self):
        """Visited-chunk bits of the last frame, OR-ed over the ranks, on every rank."""
        if self.use_peer:
            return self.fr.visited_union()
        return self.fr.reduce_visited(self.cam._visited)

    def close(self):
        if self.use_peer:
            if self._b is not None:
                self.fr.release(self._b)
            self.fr.close()


def crc_of(t):
    return "%08x" % (zlib.crc32(t.cpu().numpy().tobytes()) & 0xffffffff)


def alg_bytes(c, n_px):
    """B_alg = 1 * (in-chunk samples + neighbour probes) + 28 * hits + 2 * chunk re-evaluations + 4 * pixels"""
    return 1.0 * (float(c[1]) + float(c[4])) + 28.0 * float(c[3]) + 2.0 * float(c[5]) + 4.0 * n_px


def timed_frames(torch, dist, dev, world, path, set_pose, warmup, steps, flush, sampler, launch_count=None):
    """warmup untimed + `steps` timed frames, per-frame CUDA events, L2 flushed between frames.
    -> (ms per step [max over ranks of the summed render + assembly time], per-rank kernel ms per launch, this rank's
    sorted per-frame ms)"""
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(steps)]
    for i in range(warmup):
        set_pose(i)
        path.render()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    l0 = launch_count() if launch_count else 0
    for i in range(steps):
        set_pose(warmup + i)
        flush.zero_()
        e0, e1, e2 = ev[i]
        e0.record()
        path.render(e1)
        e2.record()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize(dev)
    if sampler:
        sampler.region(t0, time.perf_counter())
    t_total = sum(a.elapsed_time(c) for a, _, c in ev)
    t_kernel = sum(a.elapsed_time(b) for a, b, _ in ev)
    tt = torch.tensor([t_total, t_kernel], dtype=torch.float64, device=dev)
    per_rank = [tt.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, tt)
    ms_step = max(float(t[0]) for t in per_rank) / steps
    rank_kernel = [float(t[1]) / steps for t in per_rank]
    if launch_count:
        timed_frames.launches = launch_count() - l0
    return ms_step, rank_kernel, sorted(a.elapsed_time(c) for a, _, c in ev)


def rank_stats(v):
    s = sorted(v)
    return {"min": s[0], "median": s[len(s) // 2], "max": s[-1], "all": [round(x, 5) for x in v]}


def frame_checks(torch, dist, dev, rank, world, cam, path, set_pose, n_poses, alone_poses=None):
    """Outside the timed regions: CRC32 of rank 0's assembled frame per pose, the same frame rendered by rank 0 ALONE
    (N > 1; for the poses in `alone_poses`, default all) and the kernel's counters summed over the ranks."""
    crcs, alone, counters, n_px = [], [], np.zeros(8, np.int64), 0
    for k in range(n_poses):
        set_pose(k)
        path.render()
        torch.cuda.synchronize(dev)
        if rank == 0:
            crcs.append(crc_of(path.frame()))
        if world > 1:
            dist.barrier()
        out = cam.render(stripes=(8, world, rank), want=("rgba8", "counters"), prepare=False)
        c = out["counters"].clone()
        if world > 1:
            dist.all_reduce(c)
        counters += c.cpu().numpy()
        n_px += int(cam.settings.width) * int(cam.settings.height)
        if rank == 0:
            if world > 1 and (alone_poses is None or k in alone_poses):
                alone.append(crc_of(cam.render(want=("rgba8",), prepare=False)["rgba8"]))
            else:
                alone.append(crcs[-1] if world == 1 else None)
        if world > 1:
            dist.barrier()
    return crcs, alone, counters, n_px


def same_frames(crcs, alone):
    pairs = [(a, b) for a, b in zip(crcs, alone) if b is not None]
    return bool(pairs) and all(a == b for a, b in pairs)


def roofline_of(counters, n_px, launches, kernel_ms, peak, peaks_ok, extra=None):
    b = alg_bytes(counters, n_px) / launches
    ach = b / (kernel_ms * 1e-3) / 1e9
    r = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
         "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks_ok else "fallback 6650", "kernel": "render_persistent",
         "kernel_ms_per_launch": kernel_ms, "alg_bytes_per_launch": b,
         "per_ray": {k: float(counters[j]) / max(1, counters[0]) for j, k in
                     enumerate(["rays", "in_chunk_samples", "void_samples", "hits", "neighbour_probes", "chunk_reevals", "iterations"])}}
    if extra:
        r.update(extra)
    return r


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from python_raytracer_b200 import GpuCamera, Settings
    from python_raytracer_b200 import build as vt_build

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the GPU path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    sampler = ClockSampler(local_rank) if rank == 0 else None          # before the warm-up
    origin, cells, mats, poses = load_workload()
    s = Settings.from_object(C3_SETTINGS)
    cam = GpuCamera(s, device=local_rank, background_function=BUILTIN_BG, background=BUILTIN_BG)
    cam.set_world_dense(origin, cells, poses[0]["chunk_res"], mats)
    # N > 1: the render kernel stores finished pixels straight into rank 0's frame over NVLink peer memory and raises
    # an arrival flag (PeerFrameRenderer); N == 1: the band is the frame.  --gather nccl keeps the NCCL gather (A/B).
    use_peer = world > 1 and args.gather == "peer"
    path = FramePath(cam, rank, world, use_peer)
    fr = path.fr
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rays_frame = C3_SETTINGS["width"] * C3_SETTINGS["height"] * C3_SETTINGS["samples"]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def set_pose(i):
        p = poses[i % len(poses)]
        cam.pos, cam.rot = p["pos"], p["rot"]
        cam.set_chunk_res(p["chunk_res"])
        cam._prepare()

    # ---- kernel-resident timing: inputs in HBM, per-frame events around render (+ assembly) ----------------------
    ms_step, rank_kernel, per_frame = timed_frames(torch, dist, dev, world, path, set_pose, args.warmup, args.steps, flush, sampler,
                                                   launch_count=lambda: cam.launch_count)
    launches = timed_frames.launches
    t_kernel_ms = max(rank_kernel)
    value = rays_frame / (ms_step * 1e-3)
    ms_median, ms_p95 = per_frame[len(per_frame) // 2], per_frame[min(len(per_frame) - 1, int(0.95 * len(per_frame)))]

    # ---- end to end through the public API with host buffers ------------------------------------------------
    # Every step: camera + LOD table + materials host -> device, render, assembly, frame device -> host into pinned
    # memory.  Frames are pipelined two deep (the read-out of frame i overlaps the render of frame i + 1), all copies
    # are complete when the timed region ends.
    if use_peer and path._b is not None:
        fr.release(path._b)
        path._b = None
    hosts = [torch.empty((C3_SETTINGS["height"], C3_SETTINGS["width"], 4), dtype=torch.uint8).pin_memory() for _ in range(2)] if rank == 0 else [None, None]
    host = hosts[0]
    for i in range(2):
        p = poses[i % len(poses)]
        fr.wait(fr.submit_frame(p["pos"], p["rot"], p["chunk_res"], hosts[i % 2]))
    barrier()
    t0 = time.perf_counter()
    prev = None
    for i in range(args.steps):
        p = poses[(args.warmup + i) % len(poses)]
        tk = fr.submit_frame(p["pos"], p["rot"], p["chunk_res"], hosts[i % 2])
        fr.wait(prev)
        prev = tk
    fr.wait(prev)
    barrier()
    e2e_s = time.perf_counter() - t0
    if sampler:
        sampler.region(t0, t0 + e2e_s)
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = rays_frame * args.steps / float(te[0])
    h2d = int(mats.shape[0] * 64 + poses[0]["chunk_res"].size + 8 * 8)
    d2h = int(host.numel()) if host is not None else 0

    # ---- frame checksums, N-GPU == 1-GPU, counters (outside the timed regions) ---------------------------------------
    crcs, alone, counters, n_px = frame_checks(torch, dist, dev, rank, world, cam, path, set_pose, len(poses))
    path.close()
    cam.close()
    line = None
    if rank == 0:
        # ---- roofline of the dominant kernel (render_persistent): algorithmic bytes from its own counters -----------
        traffic, traffic_src, ncu_side, stale = None, None, None, None
        for name in ("r2_traffic.json", "r1_traffic.json"):
            try:          # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
                tj = json.load(open(os.path.join(ROOT, "profiles", name)))
            except Exception:
                continue
            traffic, traffic_src = float(tj["traffic_bytes_per_launch_mean"]), tj["source"]
            stale = tj.get("kernel_sha256") != vt_build.kernel_source_sha256()      # the capture was taken from other sources
            ncu_side = {k: [round(float(v), 2) for v in tj[k]] for k in
                        ("issue_active_pct", "lanes_per_instruction", "l1_hit_pct", "l2_hit_pct", "fp64_pipe_pct") if k in tj}
            break
        roofline = roofline_of(counters, n_px, len(poses), t_kernel_ms, peak, bool(peaks), extra={
            "traffic": traffic, "traffic_source": traffic_src, "traffic_stale": stale, "ncu_per_pose": ncu_side,
            "binding_resource": "instruction issue on a divergent fp64 loop, not HBM: the scene is L1/L2 resident (DESIGN.md §4); "
                                "issue-slot utilisation and lanes per instruction are in ncu_per_pose",
            "issue_frac": (float(np.mean(ncu_side["issue_active_pct"])) / 100.0) if ncu_side and "issue_active_pct" in ncu_side else None})
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "ms_median": ms_median, "ms_p95": ms_p95, "frames_per_sec": 1000.0 / ms_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(world, use_peer),
                "gpu_launches": int(launches), "rank_kernel_ms": rank_stats(rank_kernel),
                "value_is": ("pipelined throughput: frames are independent and ranks > 0 may run up to two frame buffers ahead of rank 0; "
                             "not per-frame latency") if world > 1 else "one launch per frame, per-frame CUDA events",
                "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": 1000.0 * float(te[0]) / args.steps,
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "roofline": roofline, "frame_crc": {"c3": crcs}, "frame_crc_1gpu": {"c3": alone},
                "identical_to_1gpu": same_frames(crcs, alone)}
    # ---- the other named shapes -------------------------------------------------------------------------------------
    if not args.no_configs:
        cfg = {}
        for name, fn in (("c4", config_c4), ("c5", config_c5)):
            try:
                cfg[name] = fn(args, torch, dist, dev, rank, world, local_rank, use_peer, flush, sampler, peak, bool(peaks))
            except Exception as e:      # a failure here must not take the headline line down with it (single GPU)
                if world > 1:
                    raise
                cfg[name] = {"error": "%s: %s" % (type(e).__name__, e)}
        if rank == 0:
            line["configs"] = cfg
            for name in cfg:
                if cfg[name] and "frame_crc" in cfg[name]:
                    line["frame_crc"][name] = cfg[name]["frame_crc"]
                    line["frame_crc_1gpu"][name] = cfg[name]["frame_crc_1gpu"]
            line["identical_to_1gpu"] = all(same_frames(line["frame_crc"][k], line["frame_crc_1gpu"][k]) for k in line["frame_crc"])
    if rank == 0:
        line["clocks"] = sampler.stop() if sampler else None
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = port_baseline(C3_SETTINGS, origin, cells, mats, poses, 1, 8, "1920x1080x4spp, two frames per pose")
            line["cpu_baselines"] = [line["cpu_baseline"]]
            if reference_available():
                try:
                    from oracle import refbench
                    r = refbench.time_reference(5.0)
                    line["cpu_baselines"].append({
                        "value": r["rays_per_sec"], "unit": UNIT, "cores": r["procs"], "kind": "reference", "seconds": round(r["seconds"], 2),
                        "sample": "%d frames of C1 (the default mod exactly as shipped: 64x48, 1 spp, %d rays per frame) through the reference's "
                                  "own Camera.tile (unmodified Python, %s) on a persistent multiprocessing pool of %d workers; scene build "
                                  "%.1f s excluded" % (r["frames"], r["rays_per_frame"], r["source"], r["procs"], r["build_seconds"])})
                    c3 = refbench.C3Reference(C3_SETTINGS, POSES, REFERENCE_STRIDE)
                    c3.step(0)
                    secs = sum(c3.step(k) for k in range(8))
                    line["cpu_baselines"].append({
                        "value": c3.rays_per_step * 8 / secs, "unit": UNIT, "cores": c3.procs, "kind": "reference", "seconds": round(secs, 2),
                        "sample": "8 steps of the bench workload (C3), every %s pixel in x and y of the 1080p frame (%d rays each), same "
                                  "reference code and pool" % (_ordinal(REFERENCE_STRIDE), c3.rays_per_step)})
                    c3.close()
                except Exception as e:
                    line["cpu_baselines"].append({"kind": "reference", "error": "%s: %s" % (type(e).__name__, e)})
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def config_c4(args, torch, dist, dev, rank, world, local_rank, use_peer, flush, sampler, peak, peaks_ok):
    """BASELINE.json configs[3]: synthetic 512^3 bricked terrain at 3840x2160, 16 spp, tile-split across the GPUs."""
    from python_raytracer_b200 import GpuCamera, Settings
    from python_raytracer_b200.scenes import C4_SETTINGS, c4_workload
    sd = dict(C4_SETTINGS); sd["threads"] = 1
    origin, cells, mats, res, poses = c4_workload()
    cam = GpuCamera(Settings.from_object(sd), device=local_rank, background_function=BUILTIN_BG, background=BUILTIN_BG)
    cam.set_world_dense(origin, torch.from_numpy(cells), res, mats)
    path = FramePath(cam, rank, world, use_peer)

    def set_pose(i):
        p = poses[i % len(poses)]
        cam.pos, cam.rot = p["pos"], p["rot"]
        cam._prepare()

    steps = 8
    ms_step, rank_kernel, per_frame = timed_frames(torch, dist, dev, world, path, set_pose, 3, steps, flush, sampler)
    # rank 0 alone renders one of the four poses (60 ms a frame) for the N-GPU == 1-GPU check
    crcs, alone, counters, n_px = frame_checks(torch, dist, dev, rank, world, cam, path, set_pose, len(poses), alone_poses=(0,))
    stats = cam.world_stats()
    path.close()
    cam.close()
    if rank != 0:
        return None
    rays = sd["width"] * sd["height"] * sd["samples"]
    out = {"workload": "512^3 terrain (value-noise height field, 4 stone bands + 1 %% emissive surface cells, %d bricks = %.0f MB pool), "
                       "3840x2160, 16 spp, max_bounces 4, dist_max 768, chunk_lod 2, camera (0, 300, 0) 30 degrees down, 4 yaw angles "
                       "(BASELINE.json configs[3])" % (stats["bricks"], stats["bricks"] * 512 / 1e6),
           "rays_per_frame": rays, "steps": steps, "warmup": 3, "ms_per_step": ms_step, "value": rays / (ms_step * 1e-3), "unit": UNIT,
           "ms_median": per_frame[len(per_frame) // 2], "rank_kernel_ms": rank_stats(rank_kernel),
           "roofline": roofline_of(counters, n_px, len(poses), max(rank_kernel), peak, peaks_ok),
           "frame_crc": crcs, "frame_crc_1gpu": alone, "identical_to_1gpu": same_frames(crcs, alone)}
    if world == 1 and not args.no_cpu_baseline:
        pp = [dict(pos=p["pos"], rot=p["rot"], chunk_res=res) for p in poses]
        out["cpu_baseline"] = port_baseline(sd, origin, cells, mats, pp, 4, 4, "3840x2160x16spp, one frame per pose")
    return out


def config_c5(args, torch, dist, dev, rank, world, local_rank, use_peer, flush, sampler, peak, peaks_ok):
    """BASELINE.json configs[4]: large Goxel-text model (~10 M voxels) with chunk culling on, 1080p 8 spp.  The model goes
    through the device world builder; 8 warm-up frames feed the traversed chunks back into the resident set
    (init.py:446-452) -- at N > 1 the union of the ranks' visited bits is assembled over NVLink by the render kernels --
    then frames are timed with the visited output on."""
    from python_raytracer_b200 import GpuCamera, Settings
    from python_raytracer_b200.builder import SceneObject, WorldBuilder
    from python_raytracer_b200.mathutil import euler_to_quaternion
    from python_raytracer_b200.scenes import C5_EULER, C5_MATERIALS, C5_POS, C5_SETTINGS, C5_SIZE, goxel_model_grid
    sd = dict(C5_SETTINGS); sd["threads"] = 1
    grid = goxel_model_grid(C5_SIZE)
    cam = GpuCamera(Settings.from_object(sd), device=local_rank, background_function=BUILTIN_BG, background=BUILTIN_BG)
    rot = euler_to_quaternion(*C5_EULER)
    cam.pos, cam.rot = C5_POS, rot
    wb = WorldBuilder(cam)
    cells = wb.build([SceneObject(sprite=grid, pos=(0.0, 0.0, 0.0))], C5_MATERIALS)
    res_all = wb.update_lod(visible=None)
    path = FramePath(cam, rank, world, use_peer, visited=True)
    cam._prepare()
    res = res_all
    for _ in range(8):                                                     # culling warm-up
        cam._visited.zero_()
        path.render()
        vis = cam.visited_slots(path.union()).copy()
        torch.cuda.synchronize(dev)
        res = wb.update_lod(visible=vis)
        cam._prepare()

    def set_pose(i):
        cam._visited.zero_()

    steps = 16
    ms_step, rank_kernel, per_frame = timed_frames(torch, dist, dev, world, path, set_pose, 3, steps, flush, sampler)
    crcs, alone, counters, n_px = frame_checks(torch, dist, dev, rank, world, cam, path, set_pose, 1)
    stats = cam.world_stats()
    path.close()
    if rank != 0:
        cam.close()
        return None
    rays = sd["width"] * sd["height"] * sd["samples"]
    out = {"workload": "synthetic Goxel model, %d voxels in a 512x256x512 sprite (%d bricks), built on the device, 1920x1080, 8 spp, "
                       "max_bounces 4, dist_max 384, culling on: %d of %d chunks resident after an 8-frame warm-up, visited-chunk output "
                       "on in the timed frames (BASELINE.json configs[4])" % (int((grid > 0).sum()), stats["bricks"], int((res > 0).sum()), int((res_all > 0).sum())),
           "rays_per_frame": rays, "steps": steps, "warmup": 3, "ms_per_step": ms_step, "value": rays / (ms_step * 1e-3), "unit": UNIT,
           "ms_median": per_frame[len(per_frame) // 2], "rank_kernel_ms": rank_stats(rank_kernel),
           "resident_chunks": int((res > 0).sum()), "resident_chunks_crc": "%08x" % (zlib.crc32(res.tobytes()) & 0xffffffff),
           "roofline": roofline_of(counters, n_px, 1, max(rank_kernel), peak, peaks_ok),
           "frame_crc": crcs, "frame_crc_1gpu": alone, "identical_to_1gpu": same_frames(crcs, alone)}
    if world == 1 and not args.no_cpu_baseline:
        pp = [dict(pos=C5_POS, rot=rot, chunk_res=res)]
        out["cpu_baseline"] = port_baseline(sd, wb.origin.astype(np.int32), cells.cpu().numpy(), C5_MATERIALS, pp, 2, 4, "1920x1080x8spp")
    cam.close()
    return out


def BUILTIN_BG(ray, settings):   # identity token for "data.background is lib.material_background"
    raise AssertionError("the GPU path compiles the built-in background; this is never called")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="headline (C3) only: skip the C4 / C5 blocks")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"], help="N > 1: how bands reach rank 0")
    args = ap.parse_args()
    args.warmup = max(3, args.warmup) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        if world != args.gpus and world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus %d must be launched with torchrun (one rank per GPU)" % args.gpus)
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
