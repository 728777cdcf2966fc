"""world_size-2 (and 3) gloo tests of the multi-GPU host logic on CPU: stripe partition, band gather, visited OR."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from python_raytracer_b200.parallel import StripeGather, stripe_rows

W, H = 50, 37


def pattern(y, x):
    return np.stack([(x * 7 + y) % 251, (y * 13 + x) % 241, (x ^ y) % 239, (x + y) % 233], axis=-1).astype(np.uint8)


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = StripeGather(W, H, rank, world, torch.device("cpu"))
        rows = stripe_rows(H, rank, world)
        ys, xs = np.meshgrid(rows, np.arange(W), indexing="ij")
        g.my_band().copy_(torch.from_numpy(pattern(ys, xs)))
        frame = g.gather()
        bits = torch.zeros(5, dtype=torch.int32)
        bits[rank % 5] = 1 << (3 * rank)
        bits[4] |= 1 << 30
        red = g.reduce_visited(bits)
        if rank == 0:
            ys, xs = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
            ok = np.array_equal(frame.numpy(), pattern(ys, xs))
            want = np.zeros(5, dtype=np.int64)
            for r in range(world):
                want[r % 5] |= 1 << (3 * r)
            want[4] |= 1 << 30
            ok = ok and np.array_equal(red.numpy().astype(np.int64), want)
            ret.put(bool(ok))
        else:
            assert frame is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_stripe_gather_gloo(world):
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert ret.get(timeout=5) is True


def test_stripe_partition_is_a_partition():
    for world in (1, 2, 3, 4, 8):
        allrows = np.concatenate([stripe_rows(1080, r, world) for r in range(world)])
        assert np.array_equal(np.sort(allrows), np.arange(1080))
        sizes = [len(stripe_rows(1080, r, world)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 8
