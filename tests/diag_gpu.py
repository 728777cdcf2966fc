"""Diagnostic (not a test): per-case mismatch statistics GPU vs golden / oracle.  python -m tests.diag_gpu"""
import sys
import numpy as np
from tests.golden_util import case_names, load_case, stochastic
from tests.test_gpu_parity import gpu_case
from oracle import cport

COLS = "r g b energy bounces step hits fx fy fz fflag fstep vx vy vz".split()
for name in (sys.argv[1:] or case_names()):
    c = load_case(name)
    g = gpu_case(c)
    world = cport.World(c["origin"], c["cells"], c["chunk_res"], c["materials"])
    o = cport.render(c["settings"], c["cam_pos"], c["cam_rot"], float(c["cam_lens"]), world, pixels=c["pixels"],
                     rng=cport.RNG_PHILOX, seed=0)
    for tag, ref in (("gold", c["ray0"]), ("orac", o["ray0"])):
        if tag == "gold" and stochastic(c):
            continue
        d = np.abs(g["ray0"][:, :15] - ref)
        nz = {COLS[k]: (int((d[:, k] > 0).sum()), float(d[:, k].max())) for k in range(15) if (d[:, k] > 0).any()}
        print("%-18s %s n=%d mismatches: %s" % (name, tag, len(ref), nz))
    print("   counters gpu", g["counters"][:7].tolist(), "oracle", o["counters"][:6].tolist(),
          "visited eq", np.array_equal(g["visited"], o["visited"]), "mean maxdiff vs oracle", float(np.abs(g["mean"] - o["mean"]).max()))
