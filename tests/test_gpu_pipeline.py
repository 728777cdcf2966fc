"""GPU tests of the host-facing frame loop (FrameRenderer / PeerFrameRenderer at world size 1): the pipelined
submit_frame / wait pair returns the same frames as the blocking frame_to_host call, bit for bit, while the camera
moves and the per-chunk LOD table changes every frame."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def _setup(width=640, height=360):
    import bench
    from python_raytracer_b200 import GpuCamera, Settings
    origin, cells, mats, poses = bench.load_workload()
    sd = dict(bench.C3_SETTINGS); sd.update(width=width, height=height)
    cam = GpuCamera(Settings.from_object(sd), device=0, background_function=bench.BUILTIN_BG, background=bench.BUILTIN_BG)
    cam.set_world_dense(origin, cells, poses[0]["chunk_res"], mats)
    return cam, poses


@pytest.mark.parametrize("kind", ["gather", "peer"])
def test_pipelined_frames_equal_blocking_frames(kind):
    from python_raytracer_b200.parallel import FrameRenderer, PeerFrameRenderer
    cam, poses = _setup()
    fr = (PeerFrameRenderer if kind == "peer" else FrameRenderer)(cam, 0, 1)
    h, w = 360, 640
    n = 7
    blocking = []
    host = torch.empty((h, w, 4), dtype=torch.uint8).pin_memory()
    for f in range(n):
        p = poses[f % len(poses)]
        fr.frame_to_host(p["pos"], p["rot"], p["chunk_res"], host)
        blocking.append(host.clone())
    assert any(not torch.equal(blocking[0], b) for b in blocking[1:])      # the camera really moves
    hosts = [torch.empty((h, w, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]
    prev = None
    for f in range(n + 1):
        tk = None
        if f < n:
            p = poses[f % len(poses)]
            tk = fr.submit_frame(p["pos"], p["rot"], p["chunk_res"], hosts[f % 2])
        if f > 0:
            fr.wait(prev)
            assert torch.equal(hosts[(f - 1) % 2], blocking[f - 1]), "frame %d" % (f - 1)
        prev = tk
    if kind == "peer":
        assert fr.wait_status() == 0
        fr.close()
    cam.close()


def test_table_uploads_do_not_wait_for_the_frame_in_flight():
    """vt_set_materials / vt_set_chunk_table stage through pinned slots: many back-to-back updates with renders in
    between give the same frame as one update followed by one render."""
    cam, poses = _setup(320, 180)
    p = poses[3 % len(poses)]
    cam.pos, cam.rot = p["pos"], p["rot"]
    cam.set_chunk_res(p["chunk_res"])
    want = cam.render(want=("rgba8",))["rgba8"].clone()
    for k in range(20):                                 # more updates than staging slots, no sync in between
        q = poses[k % len(poses)]
        cam.pos, cam.rot = q["pos"], q["rot"]
        cam.set_chunk_res(q["chunk_res"])
        cam.render(want=("rgba8",))
    cam.pos, cam.rot = p["pos"], p["rot"]
    cam.set_chunk_res(p["chunk_res"])
    got = cam.render(want=("rgba8",))["rgba8"]
    assert torch.equal(want, got)
    cam.close()


def test_bench_line_contract():
    """bench.py end to end on one GPU (short run): one JSON line with every key the measurement contract names."""
    import json
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--steps", "4", "--warmup", "3"], capture_output=True,
                       text=True, timeout=600, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert key in line, key
    assert line["steps"] == 4 and line["n_gpus"] == 1 and line["gpu_launches"] >= 4
    assert line["value"] > 1e8 and line["e2e"]["value"] > 1e8 and line["e2e"]["d2h_bytes_per_step"] == 1920 * 1080 * 4
    rf = line["roofline"]
    assert rf["bound"] == "hbm" and 0 < rf["frac"] < 1 and abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9
    assert line["cpu_baseline"]["kind"] in ("port", "reference") and line["cpu_baseline"]["value"] > 0
