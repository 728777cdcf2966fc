import cProfile, pstats, sys, time
sys.path.insert(0, '/root/repo')
import torch, bench
from python_raytracer_b200 import GpuCamera, Settings
from python_raytracer_b200.parallel import FrameRenderer
origin, cells, mats, poses = bench.load_workload()
sd = dict(bench.C3_SETTINGS); sd.update(width=240, height=136)
cam = GpuCamera(Settings.from_object(sd), device=0, background_function=bench.BUILTIN_BG, background=bench.BUILTIN_BG)
cam.set_world_dense(origin, cells, poses[0]["chunk_res"], mats)
fr = FrameRenderer(cam, 0, 1)
hosts = [torch.empty((136, 240, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]
def loop(n):
    prev = None
    for i in range(n):
        p = poses[i % 4]
        tk = fr.submit_frame(p["pos"], p["rot"], p["chunk_res"], hosts[i % 2])
        fr.wait(prev); prev = tk
    fr.wait(prev)
loop(50)
t0 = time.perf_counter(); loop(2000); dt = time.perf_counter() - t0
print("host-bound frame loop: %.1f us per frame" % (1e6 * dt / 2000))
pr = cProfile.Profile(); pr.enable(); loop(2000); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
