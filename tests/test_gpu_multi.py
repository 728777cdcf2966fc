"""N-GPU == 1-GPU, through pytest: launches tests/mgpu_peer_check.py under torchrun on every GPU of the box (skipped on a
one-GPU box).  The script checks, bit for bit on rank 0: the fused NVLink frame path and the NCCL gather against the
single-GPU frame (blocking and pipelined), and the C5 culling loop -- visited union over peer memory, chunk updates on
every rank -- against a single-GPU camera.  Its log is kept in gpurun_out/r2_multi_gpu_check.log."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
def test_n_gpu_frames_and_culling_equal_single_gpu():
    n = min(8, torch.cuda.device_count())
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "mgpu_peer_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "r2_multi_gpu_check.log"), "w") as f:
        f.write("world size %d\n" % n + res.stdout + "\n--- stderr ---\n" + res.stderr[-4000:])
    assert res.returncode == 0 and "PEER CHECK OK" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]
