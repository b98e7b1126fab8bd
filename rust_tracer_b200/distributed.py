"""Multi-GPU plumbing: one process per GPU (torch.distributed), interleaved row tiles, and
getting the quantised RGB8 rows to rank 0 -- the only exchange step of the path.
Rendering itself has no collective: pixels are independent (reference src/render.rs:31-38
writes one cell per iteration).

PeerFrame: the frame lives on rank 0's GPU, every rank's resolve kernel stores its rows there
over NVLink, completion travels through signal words in the same memory.  gather_rows is the
fallback where peer mapping is not available, and what the CPU (gloo) tests exercise.

NCCL has no native gather to ragged sizes, so the gather is point-to-point: every rank
!= 0 sends its tile buffer, rank 0 receives each into place (batched isend/irecv, which
NCCL runs as one grouped send/recv over NVLink).  On CPU (gloo) the same code runs the
world_size-2 tests.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import abi
from ._lib import check, lib
from .api import tile_rows


class _RawCudaBuffer:
    """Lets torch view a raw device allocation (rank 0 reads the finished frame as a tensor)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


# Signal words of a PeerFrame (include/rt_b200.h: RT_FRAME_SIGNAL_WORDS = 256 words behind the pixels):
# word r = rank r has finished frame #value; word 64 + r = rank r is about to start frame #value (align());
# word RT_FRAME_WORD_RELEASED = rank 0 has consumed frame #value.
_WORD_READY0 = 64
_WORD_READY_MAX_RANKS = 64


class PeerFrame:
    """The whole RGB8 frame on rank 0's GPU, mapped into every rank (include/rt_b200.h:
    rt_frame_*): each rank renders its row tiles straight into it with RT_FLAG_FRAME_ROWS --
    the stores of the resolve kernel travel over NVLink -- and tells rank 0 it is done through
    a signal word behind the pixels, in stream order.  No host barrier, no collective:

        frame.begin()                       # peers: wait until rank 0 has released the previous frame
        frame.align()                       # optional (measurements): a device-side start line for all ranks
        scene.render_device(..., d_rgb_u8=frame.ptr, flags=frame.flags, rank=, world=)
        full = frame.collect()              # every rank signals; rank 0 waits for all and gets the frame
        ... rank 0 consumes `full` ...
        frame.release()                     # rank 0: the peers may overwrite the pixels

    Replaces the NCCL send/recv gather below; `PeerFrame.create` returns None where the
    mapping is refused (no peer access, IPC not permitted), and callers fall back to it."""

    def __init__(self, ptr, height, width, rank, world, local_device, imported, timeout_ms=20000):
        self.ptr, self.height, self.width = ptr, height, width
        self.rank, self.world, self.local_device, self.imported = rank, world, local_device, imported
        self.nbytes = height * width * 3
        self.flags = abi.RT_FLAG_FRAME_ROWS
        self.epoch = 0          # number of the frame being rendered / last collected
        self.timeout_ms = timeout_ms

    @classmethod
    def create(cls, height, width, rank, world, local_device):
        if world > _WORD_READY0:  # the "finished" words must stay below the "ready" words
            return None
        nbytes = height * width * 3
        dev = torch.device("cuda", local_device)
        handle = torch.zeros(abi.RT_FRAME_HANDLE_BYTES, dtype=torch.uint8, device=dev)
        ptr = C.c_void_p()
        ok = torch.ones(1, dtype=torch.int32, device=dev)
        if rank == 0:
            hbuf = (C.c_uint8 * abi.RT_FRAME_HANDLE_BYTES)()
            try:
                check(lib().rt_frame_alloc(local_device, nbytes, C.byref(ptr)))
                check(lib().rt_frame_export(local_device, ptr, hbuf))
                handle.copy_(torch.tensor(list(hbuf), dtype=torch.uint8))
            except RuntimeError:
                ok.zero_()
        if world > 1:
            dist.broadcast(handle, 0)
        if rank != 0:
            hbuf = (C.c_uint8 * abi.RT_FRAME_HANDLE_BYTES)(*handle.cpu().tolist())
            try:
                check(lib().rt_frame_import(local_device, hbuf, C.byref(ptr)))
            except RuntimeError:
                ok.zero_()
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)  # all ranks or none
        frame = cls(ptr.value, height, width, rank, world, local_device, rank != 0) if ptr.value else None
        if int(ok.item()) == 0:
            if frame is not None:
                frame.close()
            return None
        return frame

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.local_device).cuda_stream)

    def begin(self):
        """Before rendering the next frame into the buffer.  A peer waits (on the device) until rank 0
        has released the previous one; rank 0 itself never waits here."""
        self.epoch += 1
        if self.rank != 0 and self.epoch > 1:
            check(lib().rt_frame_wait(self.local_device, C.c_void_p(self.ptr), self.nbytes, abi.RT_FRAME_WORD_RELEASED, 1,
                                      self.epoch - 1, self.timeout_ms, self._stream()))

    def align(self):
        """Every rank, after begin(): publish "I am about to start frame #epoch" and wait (on the device;
        the call returns when the wait is over) until every rank has.  A start line for measurements:
        after it the ranks' render calls begin within a launch latency of each other, so the longest
        (start -> done) span over the ranks IS the frame time.  Not needed for correctness."""
        if self.world > _WORD_READY_MAX_RANKS:
            raise RuntimeError("PeerFrame.align() supports at most %d ranks" % _WORD_READY_MAX_RANKS)
        check(lib().rt_frame_signal(self.local_device, C.c_void_p(self.ptr), self.nbytes, _WORD_READY0 + self.rank, self.epoch,
                                    self._stream()))
        check(lib().rt_frame_wait(self.local_device, C.c_void_p(self.ptr), self.nbytes, _WORD_READY0, self.world, self.epoch,
                                  self.timeout_ms, self._stream()))

    def collect(self):
        """After this rank's render call: publish "my rows of frame #epoch are in place".  Rank 0 then
        waits (on the device) for every rank's word and returns the frame as a [height, width, 3]
        uint8 tensor (a view, no copy); the other ranks return None without waiting for anyone."""
        if self.epoch == 0:
            raise RuntimeError("PeerFrame.collect() without begin()")
        check(lib().rt_frame_signal(self.local_device, C.c_void_p(self.ptr), self.nbytes, self.rank, self.epoch, self._stream()))
        if self.rank != 0:
            return None
        check(lib().rt_frame_wait(self.local_device, C.c_void_p(self.ptr), self.nbytes, 0, self.world, self.epoch,
                                  self.timeout_ms, self._stream()))
        return torch.as_tensor(_RawCudaBuffer(self.ptr, (self.height, self.width, 3)), device=torch.device("cuda", self.local_device))

    def release(self):
        """Rank 0, once it has consumed the collected frame (ordered after the work already on the
        current stream, e.g. a copy to the host): the peers may start writing the next one."""
        if self.rank == 0 and self.epoch > 0:
            check(lib().rt_frame_signal(self.local_device, C.c_void_p(self.ptr), self.nbytes, abi.RT_FRAME_WORD_RELEASED,
                                        self.epoch, self._stream()))

    def close(self):
        """Peers unmap, the owner frees; the caller makes sure (a barrier) that the peers are done first."""
        if self.ptr:
            lib().rt_frame_release(self.local_device, C.c_void_p(self.ptr), 1 if self.imported else 0)
            self.ptr = None


_row_index_cache = {}


def _row_index(height, rank, world, rows_per_tile, device):
    """Device tensor of the rows rank `rank` renders (cached: the gather runs every frame)."""
    key = (height, rank, world, rows_per_tile, str(device))
    idx = _row_index_cache.get(key)
    if idx is None:
        idx = torch.from_numpy(tile_rows(height, rank, world, rows_per_tile)).to(device, torch.long)
        _row_index_cache[key] = idx
    return idx


def assemble_rows(tiles, height, width, world, rows_per_tile):
    """tiles[r]: [rows_r, width, 3] uint8 tensor of rank r -> [height, width, 3] image."""
    full = torch.empty((height, width, 3), dtype=tiles[0].dtype, device=tiles[0].device)
    for r in range(world):
        full[_row_index(height, r, world, rows_per_tile, full.device)] = tiles[r]
    return full


def gather_rows(tile, height, width, rank, world, rows_per_tile, device=None):
    """Gather every rank's rendered rows to rank 0; returns the full image there, None elsewhere.

    tile: this rank's [rows, width, 3] uint8 tensor (device tensor for NCCL, CPU for gloo)."""
    if world == 1:
        return tile
    device = device if device is not None else tile.device
    if rank == 0:
        bufs = [tile]
        ops = []
        for r in range(1, world):
            n = len(tile_rows(height, r, world, rows_per_tile))
            b = torch.empty((n, width, 3), dtype=tile.dtype, device=device)
            bufs.append(b)
            if n:
                ops.append(dist.P2POp(dist.irecv, b, r))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        return assemble_rows(bufs, height, width, world, rows_per_tile)
    if tile.numel():
        for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, tile.contiguous(), 0)]):
            req.wait()
    return None
