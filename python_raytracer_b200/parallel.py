"""Multi-GPU frame sharding: one process per GPU, scene replicated, image split in interleaved row stripes.

The reference's only parallelism is pixel data parallelism over worker processes (data.py:70-77, init.py:162,
193-205): each worker renders `settings.pixels[t]` and the main process composites.  Here rank g of N renders the
rows whose 8-row stripe index is g modulo N (same load-balancing idea as the reference's (x ^ y) % threads
interleave, but contiguous enough to gather), and the bands are gathered to rank 0 over NVLink with NCCL.
Pixels never need a reduction, so the N-GPU image is bit-identical to the 1-GPU image (RNG streams are keyed by
the global pixel index).  The visited-chunk mask is OR-reduced (as a max over u8 flags; NCCL has no bit-OR).
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.distributed as dist

STRIPE_ROWS = 8       # rows per interleaved stripe (2-16 measure the same; pass `stripe=` to the renderers to change it)


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

    # ---- pipelined variant (single GPU): the frame's device -> host copy overlaps the next frame's render ----------
    def submit_frame(self, pos, rot, chunk_res, host_rgba8):
        """Like frame_to_host but returns at once with a ticket; ``wait(ticket)`` blocks until ``host_rgba8`` holds
        the frame.  Two device frames alternate, the copy runs on its own stream."""
        if self.world > 1:              # the NCCL gather baseline stays blocking; PeerFrameRenderer is the pipelined path
            self.frame_to_host(pos, rot, chunk_res, host_rgba8)
            return None
        dev = self.cam.device
        if not hasattr(self, "_pipe"):
            self._pipe = dict(frames=[torch.zeros_like(self.frame) for _ in range(2)], copy=torch.cuda.Stream(dev),
                              rendered=[torch.cuda.Event() for _ in range(2)], copied=[None, None], n=0)
        p = self._pipe
        b = p["n"] % 2
        p["n"] += 1
        main = torch.cuda.current_stream(dev)
        self.cam.pos, self.cam.rot = pos, rot
        if chunk_res is not None:
            self.cam.set_chunk_res(chunk_res)
        if p["copied"][b] is not None:
            main.wait_event(p["copied"][b])                  # frame b was read out before it is rendered again
        self.cam.render(stripes=(self.stripe, 1, 0), want=("rgba8",), prepare=True, out={"rgba8": p["frames"][b]})
        p["rendered"][b].record(main)
        p["copy"].wait_event(p["rendered"][b])
        with torch.cuda.stream(p["copy"]):
            host_rgba8.copy_(p["frames"][b], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(p["copy"])
        p["copied"][b] = ev
        return ev

    @staticmethod
    def wait(ticket):
        if ticket is not None:
            ticket.synchronize()


# ------------------------------------------------------------------------------------------------------------
# Frame assembly over NVLink peer memory: the render kernel itself delivers the pixels (no collective per frame).

class _DevView:
    """Raw device pointer -> torch tensor view (no copy) through __cuda_array_interface__."""

    def __init__(self, ptr, shape, typestr="|u1"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class PeerFrameRenderer:
    """One rank of the multi-GPU frame path with the gather fused into the render kernel.

    Rank 0 owns ``buffers`` full frames and the arrival flags (cudaMalloc'ed by libvoxtrace, exported with CUDA IPC);
    every other rank opens them and renders its stripes with a *frame target*: finished pixels are stored straight
    into rank 0's frame over NVLink and the last warp of the launch raises the rank's arrival flag.  Rank 0 queues
    a flag wait on its stream where the NCCL gather used to be.  A frame buffer is handed back to the renderers
    with a credit flag in each renderer's own memory, so nobody overwrites a frame rank 0 is still reading.
    The reference's equivalent is the pool's result pipe + ``draw_tile`` (init.py:185-190, 205).
    """

    FLAGS_PER_BUFFER = 32
    VIS_RING = 8          # visited-union masks in rank 0's memory, one per frame modulo 8 (see render_frame)

    def __init__(self, cam, rank: int, world: int, stripe: int = STRIPE_ROWS, buffers: int = 2, visited_words: int = 0):
        """visited_words > 0 (u32 words of the camera's visited-chunk bitmask) enables the culling union: rank 0 then also
        owns VIS_RING masks of that size which the render kernels OR their visited bits into (render_frame(visited=True))."""
        import ctypes as C
        self.cam, self.rank, self.world, self.stripe, self.buffers = cam, rank, world, stripe, buffers
        self.w, self.h = int(cam.settings.width), int(cam.settings.height)
        self.frame_no = 0
        self._C = C
        lib, ctx = cam._lib, cam._ctx
        nbytes = self.w * self.h * 4
        self.visited_words = int(visited_words)
        self.vis_ring = None
        handles = {}
        if rank == 0:
            self.frames = []
            for b in range(buffers):
                p = C.c_void_p()
                cam._check(lib.vt_shared_alloc(ctx, nbytes, C.byref(p)))
                self.frames.append(p.value)
            p = C.c_void_p()
            cam._check(lib.vt_shared_alloc(ctx, 4 * self.FLAGS_PER_BUFFER * buffers, C.byref(p)))
            self.arrive = p.value
            shared = self.frames + [self.arrive]
            if self.visited_words:
                p = C.c_void_p()
                cam._check(lib.vt_shared_alloc(ctx, 4 * self.visited_words * self.VIS_RING, C.byref(p)))
                self.vis_ring = p.value
                shared.append(self.vis_ring)
                self.vis_views = [torch.as_tensor(_DevView(self.vis_ring + 4 * self.visited_words * k, (self.visited_words,), "<i4"),
                                                  device=cam.device) for k in range(self.VIS_RING)]
            for k, ptr in enumerate(shared):
                hb = C.create_string_buffer(64)
                cam._check(lib.vt_ipc_export(ctx, C.c_void_p(ptr), hb))
                handles[k] = hb.raw
            self.frame_views = [torch.as_tensor(_DevView(f, (self.h, self.w, 4)), device=cam.device) for f in self.frames]
        else:
            p = C.c_void_p()
            cam._check(lib.vt_shared_alloc(ctx, 4 * buffers, C.byref(p)))
            self.credit = p.value
            hb = C.create_string_buffer(64)
            cam._check(lib.vt_ipc_export(ctx, C.c_void_p(self.credit), hb))
            handles["credit"] = hb.raw
        gathered = [None] * world
        if world > 1:
            dist.all_gather_object(gathered, handles)
        else:
            gathered = [handles]
        self._opened = []
        if rank == 0:
            self.peer_credit = [None]
            for r in range(1, world):
                q = C.c_void_p()
                cam._check(lib.vt_ipc_open(ctx, gathered[r]["credit"], C.byref(q)))
                self.peer_credit.append(q.value)
                self._opened.append(q.value)
        else:
            self.frames = []
            for k in range(buffers + 1 + (1 if self.visited_words else 0)):
                q = C.c_void_p()
                cam._check(lib.vt_ipc_open(ctx, gathered[0][k], C.byref(q)))
                self._opened.append(q.value)
                if k < buffers:
                    self.frames.append(q.value)
                elif k == buffers:
                    self.arrive = q.value
                else:
                    self.vis_ring = q.value

    def _flag(self, b, r):
        return self.arrive + 4 * (b * self.FLAGS_PER_BUFFER + r)

    def render_frame(self, prepare=True, visited=False):
        """Renders this rank's stripes of the next frame into rank 0's frame buffer.  Returns the buffer index.

        visited=True (culling): the union of every rank's visited-chunk bits (the reference merges the workers'
        traversed lists every frame, init.py:189, 393) is assembled in rank 0's memory by the render kernels themselves.
        Rank 0 marks straight into mask f % VIS_RING, the other ranks mark a local mask and their launch's last warp ORs
        its non-zero words into the same mask over NVLink before the arrival flag goes up, so once rank 0 has seen the
        flags it holds the union (``visited_union``).  Mask (f + 4) % 8 is cleared by rank 0 ahead of time: ranks are
        never more than ``buffers`` frames apart (frame credits), so nobody writes it yet and nobody reads it any more."""
        C = self._C
        cam, lib, ctx = self.cam, self.cam._lib, self.cam._ctx
        f = self.frame_no
        b, k = f % self.buffers, f // self.buffers
        if prepare:
            cam._prepare()
        if visited and not self.visited_words:
            raise RuntimeError("PeerFrameRenderer was built without visited_words")
        if visited and cam._visited.numel() != self.visited_words:
            raise RuntimeError("the camera's chunk grid changed: %d visited words, the shared masks hold %d" % (cam._visited.numel(), self.visited_words))
        if self.rank > 0 and k > 0:            # buffer b must have been released k times by rank 0
            cam._check(lib.vt_stream_wait_flags(ctx, C.c_void_p(self.credit + 4 * b), 1, k, 30.0))
        cam._check(lib.vt_set_frame_target(ctx, C.c_void_p(self.frames[b]), C.c_void_p(self._flag(b, self.rank)), f + 1))
        slot = f % self.VIS_RING
        vis_out = None
        if visited:
            if self.rank == 0:
                self.vis_views[(f + self.VIS_RING // 2) % self.VIS_RING].zero_()
                vis_out = self.vis_views[slot]
            else:
                cam._check(lib.vt_set_visited_peer(ctx, C.c_void_p(self.vis_ring + 4 * self.visited_words * slot), self.visited_words))
        try:
            cam.render(stripes=(self.stripe, self.world, self.rank), want=("visited",) if visited else (), prepare=False,
                       visited_out=vis_out)
        finally:
            cam._check(lib.vt_set_frame_target(ctx, None, None, 0))
            if visited and self.rank > 0:
                cam._check(lib.vt_set_visited_peer(ctx, None, 0))
        if self.rank == 0 and self.world > 1:
            cam._check(lib.vt_stream_wait_flags(ctx, C.c_void_p(self._flag(b, 1)), self.world - 1, f + 1, 30.0))
        self._last_visited_slot = slot if visited else None
        self.frame_no += 1
        return b

    def visited_union(self):
        """The union of all ranks' visited bits of the last render_frame(visited=True), as an int32 word tensor on
        EVERY rank (rank 0 holds it once the arrival flags are in; the others get it by one small broadcast -- the
        chunk update that consumes it is host logic on every rank, init.py:446-452)."""
        if getattr(self, "_last_visited_slot", None) is None:
            raise RuntimeError("visited_union: the last frame was not rendered with visited=True")
        if self.rank == 0:
            t = self.vis_views[self._last_visited_slot].clone()
        else:
            t = torch.empty(self.visited_words, dtype=torch.int32, device=self.cam.device)
        if self.world > 1:
            dist.broadcast(t, src=0)
        return t

    def release(self, b):
        """Rank 0: frame buffer b has been consumed (stream-ordered); the renderers may overwrite it."""
        if self.rank != 0 or self.world == 1:
            return
        C = self._C
        k = (self.frame_no - 1) // self.buffers + 1
        arr = (C.c_void_p * (self.world - 1))(*[p + 4 * b for p in self.peer_credit[1:]])
        self.cam._check(self.cam._lib.vt_stream_signal_flags(self.cam._ctx, arr, self.world - 1, k))

    def frame_to_host(self, pos, rot, chunk_res, host_rgba8):
        """Same contract as FrameRenderer.frame_to_host, without the per-frame collective."""
        self.cam.pos, self.cam.rot = pos, rot
        if chunk_res is not None:
            self.cam.set_chunk_res(chunk_res)
        b = self.render_frame(prepare=True)
        if self.rank == 0:
            host_rgba8.copy_(self.frame_views[b], non_blocking=True)
            self.release(b)
            torch.cuda.current_stream(self.cam.device).synchronize()
            self._check_waits()
        return host_rgba8

    # ---- pipelined variant: rank 0 reads frame b out on a copy stream while everybody renders frame b + 1 --------
    def submit_frame(self, pos, rot, chunk_res, host_rgba8):
        self.cam.pos, self.cam.rot = pos, rot
        if chunk_res is not None:
            self.cam.set_chunk_res(chunk_res)
        dev = self.cam.device
        if self.rank == 0 and not hasattr(self, "_pipe"):
            self._pipe = dict(copy=torch.cuda.Stream(dev), copied=[None] * self.buffers)
        if self.rank == 0:
            b_next = self.frame_no % self.buffers
            if self._pipe["copied"][b_next] is not None:     # rank 0 renders into it too: the read-out must be over
                torch.cuda.current_stream(dev).wait_event(self._pipe["copied"][b_next])
        b = self.render_frame(prepare=True)
        if self.rank != 0:
            return None
        main, cp = torch.cuda.current_stream(dev), self._pipe["copy"]
        ev_r = torch.cuda.Event()
        ev_r.record(main)
        cp.wait_event(ev_r)
        C = self._C
        with torch.cuda.stream(cp):
            host_rgba8.copy_(self.frame_views[b], non_blocking=True)
            if self.world > 1:                               # hand the buffer back once it has been read out
                k = (self.frame_no - 1) // self.buffers + 1
                arr = (C.c_void_p * (self.world - 1))(*[p + 4 * b for p in self.peer_credit[1:]])
                self.cam._check(self.cam._lib.vt_stream_signal_flags_on(self.cam._ctx, C.c_void_p(cp.cuda_stream), arr, self.world - 1, k))
            ev = torch.cuda.Event()
            ev.record(cp)
        self._pipe["copied"][b] = ev
        return ev

    def wait(self, ticket):
        if ticket is not None:
            ticket.synchronize()
            self._check_waits()

    def _check_waits(self):
        """A flag wait that gave up (a peer that never arrived) must not hand a torn frame to the caller silently."""
        C = self._C
        st = C.c_int(0)
        self.cam._check(self.cam._lib.vt_wait_status_peek(self.cam._ctx, C.byref(st)))
        if st.value:
            from .abi import VoxtraceError
            raise VoxtraceError("rank %d: a flag wait timed out (flag index %d): a peer did not deliver its part of the frame" % (self.rank, st.value - 1))

    def wait_status(self) -> int:
        C = self._C
        st = C.c_int(0)
        self.cam._check(self.cam._lib.vt_wait_status(self.cam._ctx, C.byref(st)))
        return st.value

    def close(self):
        lib, ctx, C = self.cam._lib, self.cam._ctx, self._C
        torch.cuda.synchronize(self.cam.device)
        if self.world > 1:
            dist.barrier()
        for p in self._opened:
            lib.vt_ipc_close(ctx, C.c_void_p(p))
        if self.world > 1:
            dist.barrier()
        if self.rank == 0:
            self.frame_views = None
            self.vis_views = None
            for p in self.frames + [self.arrive] + ([self.vis_ring] if self.vis_ring else []):
                lib.vt_shared_free(ctx, C.c_void_p(p))
        else:
            lib.vt_shared_free(ctx, C.c_void_p(self.credit))
