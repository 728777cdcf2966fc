#!/usr/bin/env python
"""bench.py — rays/s and ms/frame of the voxel ray-march hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nproc-per-node N ... bench.py --gpus N ...     (one rank per GPU, NCCL)

A "step" is one frame: the default mod's scene at 1920x1080, samples=4, max_bounces=4, rough PBR + glass + mist
materials (BASELINE.json configs[2], the configuration the metric "1080p 4spp" is quoted on), cycling through
four camera poses.  One JSON line is printed by rank 0.  See DESIGN.md "Measurement".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

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
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.lines, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [v.strip() for v in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def oracle_rays_per_sec(origin, cells, mats, poses, stride, steps, warmup, threads):
    """The CPU arm: oracle/vt_oracle.c (C port of the reference trace loop) on the host cores, on a strided pixel
    sample of the same frames.  Returns (rays/s, rays per step, seconds)."""
    from oracle import cport
    s = C3_SETTINGS
    xs, ys = np.meshgrid(np.arange(stride // 2, s["width"], stride, dtype=np.int32),
                         np.arange(stride // 2, s["height"], stride, dtype=np.int32), indexing="ij")
    pixels = np.stack([xs.ravel(), ys.ravel()], axis=1)
    lens = s["fov"] * np.pi / 8
    total_rays, total_t = 0, 0.0
    for it in range(warmup + steps):
        p = poses[it % len(poses)]
        world = cport.World(origin, cells, p["chunk_res"], mats)
        t0 = time.perf_counter()
        o = cport.render(s, p["pos"], p["rot"], lens, world, pixels=pixels, rng=cport.RNG_PHILOX, seed=0, threads=threads)
        dt = time.perf_counter() - t0
        if it >= warmup:
            total_rays += int(o["counters"][0]); total_t += dt
    return total_rays / total_t, total_rays / max(1, steps), total_t


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path.  The Python reference itself cannot
    travel to the GPU box (it is not in this repo); its C port (the oracle, pinned bit-exact to it) is timed with
    all host threads.  Rank 0 only."""
    if rank != 0:
        return
    origin, cells, mats, poses = load_workload()
    cores = os.cpu_count() or 1
    stride = 4
    v, rays_step, secs = oracle_rays_per_sec(origin, cells, mats, poses, stride, args.steps, args.warmup, cores)
    frame_rays = C3_SETTINGS["width"] * C3_SETTINGS["height"] * C3_SETTINGS["samples"]
    line = {"metric": METRIC, "value": v, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * secs / args.steps, "ms_per_full_frame_est": 1000.0 * frame_rays / v,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(1, False, sample="every %dth pixel in x and y" % stride),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "each step = every %dth pixel in x and y of one 1920x1080x4spp frame (%d rays)" % (stride, rays_step)},
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


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from python_raytracer_b200 import GpuCamera, Settings
    from python_raytracer_b200.parallel import FrameRenderer, PeerFrameRenderer

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the GPU path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    origin, cells, mats, poses = load_workload()
    s = Settings.from_object(C3_SETTINGS)
    cam = GpuCamera(s, device=local_rank, background_function=BUILTIN_BG, background=BUILTIN_BG)
    cam.set_world_dense(origin, cells, poses[0]["chunk_res"], mats)
    # N > 1: the render kernel stores finished pixels straight into rank 0's frame over NVLink peer memory and raises
    # an arrival flag (PeerFrameRenderer); N == 1: the band is the frame.  --gather nccl keeps the NCCL gather (A/B).
    use_peer = world > 1 and args.gather == "peer"
    fr = PeerFrameRenderer(cam, rank, world) if use_peer else FrameRenderer(cam, rank, world)

    def render_and_assemble(e1=None):
        if use_peer:
            b = fr.render_frame(prepare=False)
            if e1 is not None:
                e1.record()
            fr.release(b)
        else:
            fr.render(prepare=False)
            if e1 is not None:
                e1.record()
            fr.gather()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rays_frame = C3_SETTINGS["width"] * C3_SETTINGS["height"] * C3_SETTINGS["samples"]

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

    # ---- kernel-resident timing: inputs in HBM, per-frame events around render (+ gather) ----------------
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    for i in range(args.warmup):
        set_pose(i)
        render_and_assemble()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    l0 = cam.launch_count
    for i in range(args.steps):
        set_pose(args.warmup + i)
        flush.zero_()
        e0, e1, e2 = ev[i]
        e0.record()
        render_and_assemble(e1)
        e2.record()
    barrier()
    launches = cam.launch_count - l0
    clocks = sampler.stop() if sampler else None
    t_total = sum(a.elapsed_time(c) for a, _, c in ev)          # ms, render + gather
    t_kernel = sum(a.elapsed_time(b) for a, b, _ in ev)         # ms, this rank's render kernel only
    tt = torch.tensor([t_total, t_kernel], dtype=torch.float64, device=dev)
    per_rank = [tt.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, tt)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    rank_kernel_ms = sorted(float(t[1]) / args.steps for t in per_rank)
    t_total, t_kernel = float(tt[0]), float(tt[1])
    ms_step = t_total / args.steps
    value = rays_frame / (ms_step * 1e-3)
    per_frame = sorted(a.elapsed_time(c) for a, _, c in ev)             # this rank's frames (rank 0 is the one that prints)
    ms_median, ms_p95 = per_frame[len(per_frame) // 2], per_frame[min(len(per_frame) - 1, int(0.95 * len(per_frame)))]

    # ---- end to end through the public API with host buffers ------------------------------------------------
    # Every step: camera + LOD table + materials host -> device, render, assembly, frame device -> host into pinned
    # memory.  Frames are pipelined two deep (the read-out of frame i overlaps the render of frame i + 1), all copies
    # are complete when the timed region ends.
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
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = rays_frame * args.steps / float(te[0])
    h2d = int(mats.shape[0] * 64 + poses[0]["chunk_res"].size + 8 * 8)
    d2h = int(host.numel()) if host is not None else 0

    if use_peer:
        fr.close()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (render_persistent): algorithmic bytes from its own counters -----------
    alg_bytes, cnt_tot = 0.0, np.zeros(8, dtype=np.int64)
    for i in range(args.steps):
        set_pose(args.warmup + i)
        out = cam.render(stripes=(8, world, rank), want=("rgba8", "counters"), prepare=False)
        c = out["counters"].cpu().numpy()
        cnt_tot += c
        n_px = out["rgba8"].shape[0] * out["rgba8"].shape[1]
        # B_alg = 1*(in-chunk samples + neighbour probes) + 28*hits + 2*chunk re-evaluations + 4*pixels
        alg_bytes += 1.0 * (c[1] + c[4]) + 28.0 * c[3] + 2.0 * c[5] + 4.0 * n_px
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic, traffic_src, ncu_side = None, None, None
    try:          # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
        tj = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json")))
        traffic, traffic_src = float(tj["traffic_bytes_per_launch_mean"]), tj["source"]
        # what actually bounds this kernel (same capture; profiler numbers, reported beside the roofline, not timed here)
        ncu_side = {k: [round(float(v), 2) for v in tj[k]] for k in ("issue_active_pct", "lanes_per_instruction", "l1_hit_pct", "l2_hit_pct") if k in tj}
    except Exception:
        ncu_side = None
    achieved = alg_bytes / (t_kernel * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650",
                "kernel": "render_persistent", "kernel_ms_per_launch": t_kernel / args.steps,
                "alg_bytes_per_launch": alg_bytes / args.steps, "ncu_per_pose": ncu_side,
                "per_ray": {k: float(cnt_tot[j]) / max(1, cnt_tot[0]) for j, k in
                            enumerate(["rays", "in_chunk_samples", "void_samples", "hits", "neighbour_probes", "chunk_reevals", "iterations"])}}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "ms_median": ms_median, "ms_p95": ms_p95, "frames_per_sec": 1000.0 / ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(world, use_peer),
            "clocks": clocks, "gpu_launches": int(launches),
            "rank_kernel_ms": {"min": rank_kernel_ms[0], "median": rank_kernel_ms[len(rank_kernel_ms) // 2], "max": rank_kernel_ms[-1],
                               "all": [round(v, 5) for v in rank_kernel_ms]},
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": 1000.0 * float(te[0]) / args.steps,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "roofline": roofline}
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        v, rays_step, secs = oracle_rays_per_sec(origin, cells, mats, poses, 3, 4, 1, cores)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": "4 frames (one per pose), every 3rd pixel in x and y of 1920x1080x4spp (%d rays each), "
                                          "C port of the reference trace loop (oracle/vt_oracle.c), %d threads" % (rays_step, cores)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def BUILTIN_BG(ray, settings):   # identity token for "data.background is lib.material_background"
    raise AssertionError("the GPU path compiles the built-in background; this is never called")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
