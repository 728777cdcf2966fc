"""Multi-GPU frame sharding: one process per GPU, scene replicated, image split in interleaved row stripes.

The reference's only parallelism is pixel data parallelism over worker processes (data.py:70-77, init.py:162,
193-205): each worker renders `settings.pixels[t]` and the main process composites.  Here rank g of N renders the
rows whose 8-row stripe index is g modulo N (same load-balancing idea as the reference's (x ^ y) % threads
interleave, but contiguous enough to gather), and the bands are gathered to rank 0 over NVLink with NCCL.
Pixels never need a reduction, so the N-GPU image is bit-identical to the 1-GPU image (RNG streams are keyed by
the global pixel index).  The visited-chunk mask is OR-reduced (as a max over u8 flags; NCCL has no bit-OR).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

STRIPE_ROWS = 8


def stripe_rows(height: int, rank: int, world: int, stripe: int = STRIPE_ROWS) -> np.ndarray:
    """Absolute y of the rows rank `rank` owns, increasing."""
    y = np.arange(height)
    return y[(y // stripe) % world == rank]


class StripeGather:
    """Band buffers and the collective that assembles them; device-agnostic (CUDA + NCCL, or CPU + gloo in tests)."""

    def __init__(self, width, height, rank, world, device, stripe=STRIPE_ROWS):
        self.w, self.h, self.rank, self.world, self.stripe, self.device = int(width), int(height), rank, world, stripe, device
        self.rows = [stripe_rows(self.h, r, world, stripe) for r in range(world)]
        self.max_rows = max(len(r) for r in self.rows)
        self.my_rows = len(self.rows[rank])
        self.frame = torch.zeros((self.h, self.w, 4), dtype=torch.uint8, device=device) if rank == 0 else None
        # one GPU: the band IS the frame; otherwise bands are padded to a common size for the gather
        self.band = self.frame if world == 1 else torch.zeros((self.max_rows, self.w, 4), dtype=torch.uint8, device=device)
        if world > 1 and rank == 0:
            self.recv = [torch.empty_like(self.band) for _ in range(world)]
            self.row_idx = [torch.from_numpy(r).to(device) for r in self.rows]

    def my_band(self):
        return self.band[:self.my_rows]

    def gather(self):
        """Bands -> full frame on rank 0 (returned there, None elsewhere)."""
        if self.world == 1:
            return self.frame
        dist.gather(self.band, self.recv if self.rank == 0 else None, dst=0)
        if self.rank == 0:
            for r in range(self.world):
                self.frame[self.row_idx[r]] = self.recv[r][:len(self.rows[r])]
            return self.frame
        return None

    def reduce_visited(self, visited_bits):
        """OR of the per-rank visited-chunk bitmasks (int32 words), on every rank."""
        if self.world == 1:
            return visited_bits
        bits = visited_bits.view(torch.uint8)
        shifts = torch.arange(8, device=bits.device, dtype=torch.uint8)
        flags = ((bits.unsqueeze(1) >> shifts) & 1).contiguous()
        dist.all_reduce(flags, op=dist.ReduceOp.MAX)
        packed = (flags.to(torch.int32) << shifts.to(torch.int32)).sum(dim=1).to(torch.uint8)
        return packed.view(torch.int32)


class FrameRenderer(StripeGather):
    """GpuCamera + StripeGather: what one rank does per frame."""

    def __init__(self, cam, rank: int, world: int, stripe: int = STRIPE_ROWS):
        super().__init__(cam.settings.width, cam.settings.height, rank, world, cam.device, stripe)
        self.cam = cam

    def render(self, prepare=True, want=("rgba8",)):
        """Render this rank's stripes into its band (one kernel launch)."""
        out = {"rgba8": self.my_band()}
        return self.cam.render(stripes=(self.stripe, self.world, self.rank), want=want, prepare=prepare, out=out)

    def frame_to_host(self, pos, rot, chunk_res, host_rgba8):
        """The call a user of the engine makes per frame, with HOST buffers: camera + per-chunk LOD table +
        materials go host -> device, the frame is rendered (one launch per GPU), gathered to rank 0 and copied
        device -> host into ``host_rgba8`` (pinned [H, W, 4] u8 on rank 0; ignored elsewhere)."""
        self.cam.pos, self.cam.rot = pos, rot
        if chunk_res is not None:
            self.cam.set_chunk_res(chunk_res)
        self.render(prepare=True)
        frame = self.gather()
        if self.rank == 0:
            host_rgba8.copy_(frame, non_blocking=True)
        torch.cuda.current_stream(self.cam.device).synchronize()
        return host_rgba8
