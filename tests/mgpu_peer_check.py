"""Multi-GPU check of the peer-memory frame path (run under torchrun on a box with >= 2 GPUs; tests/test_gpu_multi.py
launches it when the box has them):
   torchrun --nproc-per-node N tests/mgpu_peer_check.py
Every rank renders its stripes straight into rank 0's frame; rank 0 compares 6 frames (two buffers, three rounds, the
camera moving) with its own single-GPU render of the same frames and with the NCCL-gather path, bit for bit."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from python_raytracer_b200 import GpuCamera, Settings
from python_raytracer_b200.parallel import FrameRenderer, PeerFrameRenderer


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    origin, cells, mats, poses = bench.load_workload()
    sd = dict(bench.C3_SETTINGS); sd.update(width=960, height=540)
    cam = GpuCamera(Settings.from_object(sd), device=local, background_function=bench.BUILTIN_BG, background=bench.BUILTIN_BG)
    cam.set_world_dense(origin, cells, poses[0]["chunk_res"], mats)
    peer = PeerFrameRenderer(cam, rank, world)
    nccl = FrameRenderer(cam, rank, world)
    host = torch.empty((540, 960, 4), dtype=torch.uint8).pin_memory()
    host2 = torch.empty((540, 960, 4), dtype=torch.uint8).pin_memory()
    bad = 0
    for f in range(6):
        p = poses[f % len(poses)]
        peer.frame_to_host(p["pos"], p["rot"], p["chunk_res"], host)
        nccl.frame_to_host(p["pos"], p["rot"], p["chunk_res"], host2)
        if rank == 0:
            cam.pos, cam.rot = p["pos"], p["rot"]
            single = cam.render(want=("rgba8",))["rgba8"].cpu()
            ok1, ok2 = torch.equal(single, host), torch.equal(single, host2)
            print("frame %d  peer==single %s  nccl==single %s" % (f, ok1, ok2), flush=True)
            bad += (not ok1) + (not ok2)
    # pipelined: frame i is read out on the copy stream while frame i + 1 renders (bench.py's e2e loop)
    n_pipe = 9
    want = []
    if rank == 0:
        for f in range(n_pipe):
            p = poses[f % len(poses)]
            cam.pos, cam.rot = p["pos"], p["rot"]
            cam.set_chunk_res(p["chunk_res"])
            want.append(cam.render(want=("rgba8",))["rgba8"].cpu())
    hosts = [host, host2]
    prev = None
    for f in range(n_pipe + 1):
        tk = None
        if f < n_pipe:
            p = poses[f % len(poses)]
            tk = peer.submit_frame(p["pos"], p["rot"], p["chunk_res"], hosts[f % 2])
        if f > 0:
            peer.wait(prev)
            if rank == 0:
                ok = torch.equal(want[f - 1], hosts[(f - 1) % 2])
                print("pipelined frame %d == single %s" % (f - 1, ok), flush=True)
                bad += not ok
        prev = tk
    st = peer.wait_status()
    peer.close()
    cam.close()
    bad += culling_check(rank, world, local)
    t = torch.tensor([bad + (st != 0)], device=dev)
    dist.all_reduce(t)
    dist.destroy_process_group()
    if rank == 0:
        print("PEER CHECK", "OK" if int(t) == 0 else "FAILED", flush=True)
    sys.exit(0 if int(t) == 0 else 1)


def culling_check(rank, world, local):
    """C5 (9.9 M-voxel model, 1080p 8 spp, culling on) split over the ranks through the fused path: every rank's visited
    bits reach rank 0's mask over NVLink with the arrival flag, every rank applies the same chunk update
    (init.py:446-452), and after the same warm-up the N-GPU frames, visited unions and resident-chunk sets are the ones a
    single GPU produces."""
    from python_raytracer_b200.builder import SceneObject, WorldBuilder
    from python_raytracer_b200.mathutil import euler_to_quaternion
    from python_raytracer_b200.scenes import C5_EULER, C5_MATERIALS, C5_POS, C5_SETTINGS, C5_SIZE, goxel_model_grid
    sd = dict(C5_SETTINGS); sd["threads"] = 1
    grid = goxel_model_grid(C5_SIZE)
    cams, builders = [], []
    for _ in range(2 if rank == 0 else 1):                 # rank 0 keeps a second, single-GPU camera as the yardstick
        cam = GpuCamera(Settings.from_object(sd), device=local, background_function=bench.BUILTIN_BG, background=bench.BUILTIN_BG)
        cam.pos, cam.rot = C5_POS, euler_to_quaternion(*C5_EULER)
        wb = WorldBuilder(cam)
        wb.build([SceneObject(sprite=grid, pos=(0.0, 0.0, 0.0))], C5_MATERIALS)
        wb.update_lod(visible=None)
        cams.append(cam); builders.append(wb)
    cam, wb = cams[0], builders[0]
    cam._prepare()
    peer = PeerFrameRenderer(cam, rank, world, visited_words=cam._visited.numel())
    host = torch.empty((sd["height"], sd["width"], 4), dtype=torch.uint8).pin_memory()
    bad = 0
    for f in range(5):
        b = peer.render_frame(prepare=True, visited=True)
        union = peer.visited_union()
        if rank == 0:
            host.copy_(peer.frame_views[b])
            peer.release(b)
        torch.cuda.synchronize()
        vis = cam.visited_slots(union).copy()
        res = wb.update_lod(visible=vis)
        if rank == 0:
            c1, w1 = cams[1], builders[1]
            out = c1.render(want=("rgba8", "visited"))
            torch.cuda.synchronize()
            vis1 = c1.visited_slots(out["visited"]).copy()
            res1 = w1.update_lod(visible=vis1)
            ok = torch.equal(out["rgba8"].cpu(), host) and np.array_equal(vis, vis1) and np.array_equal(res, res1)
            print("culling frame %d: frame, visited union (%d chunks) and resident set (%d chunks) == single GPU: %s" % (
                f, int(vis.sum()), int((res > 0).sum()), ok), flush=True)
            bad += not ok
    bad += peer.wait_status() != 0
    peer.close()
    for c in cams:
        c.close()
    return bad


if __name__ == "__main__":
    main()
