"""CPU, gloo, world_size 2: the N > 1 host path -- interleaved row tiles per rank and the
gather of finished tiles to rank 0 (rust_tracer_b200/distributed.py).  The per-rank pixels
come from the oracle here (test infrastructure); on the GPU box the same code gathers the
tiles librt_b200 rendered, over NCCL."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, w, h, depth, rows_per_tile, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import orc
    import rust_tracer_b200 as rt
    from rust_tracer_b200 import abi, distributed

    rows = rt.tile_rows(h, rank, world, rows_per_tile)
    sa = orc.default_scene()
    _, u8, _ = orc.render(sa.desc, abi.default_camera(w, h), depth, rows=rows, threads=2, want_f32=False)
    tile = torch.from_numpy(u8)
    full = distributed.gather_rows(tile, h, w, rank, world, rows_per_tile, device="cpu")
    if rank == 0:
        np.save(out_path, full.numpy())
    else:
        assert full is None
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_row_tiles_gather_to_rank0(tmp_path):
    import orc
    from rust_tracer_b200 import abi

    w, h, depth, rpt = 64, 44, 4, 8  # 44 rows: the last tile is ragged
    out = str(tmp_path / "full.npy")
    mp.spawn(_worker, args=(2, _free_port(), w, h, depth, rpt, out), nprocs=2, join=True)
    sa = orc.default_scene()
    _, ref, _ = orc.render(sa.desc, abi.default_camera(w, h), depth, want_f32=False)
    assert (np.load(out) == ref).all()


def test_gather_layout_single_process():
    from rust_tracer_b200 import distributed
    import rust_tracer_b200 as rt

    h, w, world, rpt = 21, 5, 3, 4
    full = np.arange(h * w * 3, dtype=np.uint8).reshape(h, w, 3)
    tiles = [torch.from_numpy(full[rt.tile_rows(h, r, world, rpt)]) for r in range(world)]
    out = distributed.assemble_rows(tiles, h, w, world, rpt)
    assert (out.numpy() == full).all()


def test_peer_frame_protocol_against_a_recording_library(monkeypatch):
    """PeerFrame's signal-word protocol (begin / align / collect / release) with the C library replaced
    by a recorder that keeps the 256 words of one shared frame: two ranks step through three frames
    in an interleaving a real run can produce; every wait must already be satisfied when it is
    issued in that order (no deadlock), and the words used are the documented ones."""
    from rust_tracer_b200 import abi, distributed as D

    words = [0] * abi.RT_FRAME_SIGNAL_WORDS
    log = []

    class Lib:
        def rt_frame_signal(self, dev, ptr, nbytes, word, value, stream):
            assert 0 <= word < abi.RT_FRAME_SIGNAL_WORDS
            words[word] = value
            log.append(("signal", dev, word, value))
            return 0

        def rt_frame_wait(self, dev, ptr, nbytes, first, n, value, timeout_ms, stream):
            assert first >= 0 and n >= 1 and first + n <= abi.RT_FRAME_SIGNAL_WORDS
            pending = [first + i for i in range(n) if words[first + i] < value]
            log.append(("wait", dev, first, n, value, tuple(pending)))
            return 0

        def rt_frame_release(self, *a):
            return 0

    monkeypatch.setattr(D, "lib", lambda: Lib())
    monkeypatch.setattr(D.PeerFrame, "_stream", lambda self: None)
    h, w, world = 16, 8, 2
    r0 = D.PeerFrame(0x1000, h, w, 0, world, 0, imported=False)
    r1 = D.PeerFrame(0x1000, h, w, 1, world, 1, imported=True)

    def pending_of_last_wait():
        return [e for e in log if e[0] == "wait"][-1][5]

    for frame_no in (1, 2, 3):
        r0.begin()                      # rank 0 never waits here
        r1.begin()                      # frame 1: nothing to wait for; later: rank 0 released frame_no - 1 below
        if frame_no > 1:
            assert pending_of_last_wait() == ()
        # start line: both publish "ready", then both wait -- a real run may order the two signals either way
        for f in (r0, r1):
            D.check(D.lib().rt_frame_signal(f.local_device, None, f.nbytes, D._WORD_READY0 + f.rank, f.epoch, None))
        r0.align()
        assert pending_of_last_wait() == ()
        r1.align()
        assert pending_of_last_wait() == ()
        assert r1.collect() is None     # a peer signals "done" and does not wait
        monkeypatch.setattr(D.torch, "as_tensor", lambda *a, **k: "frame")
        assert r0.collect() == "frame"  # rank 0 signals, then waits for both "done" words
        assert pending_of_last_wait() == ()
        r0.release()
        assert words[abi.RT_FRAME_WORD_RELEASED] == frame_no
        assert words[0] == words[1] == frame_no and words[D._WORD_READY0] == words[D._WORD_READY0 + 1] == frame_no
    # the three word ranges do not overlap for any supported world size
    assert D._WORD_READY0 + D._WORD_READY_MAX_RANKS <= abi.RT_FRAME_WORD_RELEASED


def test_peer_frame_protocol_does_not_deadlock_under_concurrency(monkeypatch):
    """The same protocol with REAL blocking: four "ranks" run bench.py's timed-step sequence
    (begin, align, render, collect, release) concurrently in threads against one shared word array whose
    waits block until satisfied (or fail after 5 s).  Random delays shuffle the interleavings; every
    rank must get through every frame, and rank 0 must never collect a frame a peer has not signalled."""
    import random
    import threading
    import time
    from rust_tracer_b200 import abi, distributed as D

    world, frames = 4, 25
    words = [0] * abi.RT_FRAME_SIGNAL_WORDS
    cond = threading.Condition()
    rendered = [0] * world          # frames each rank has "rendered"
    errors = []

    class Lib:
        def rt_frame_signal(self, dev, ptr, nbytes, word, value, stream):
            with cond:
                words[word] = value
                cond.notify_all()
            return 0

        def rt_frame_wait(self, dev, ptr, nbytes, first, n, value, timeout_ms, stream):
            deadline = time.monotonic() + 5.0
            with cond:
                while any(words[first + i] < value for i in range(n)):
                    left = deadline - time.monotonic()
                    if left <= 0:
                        errors.append("timeout waiting for words [%d, %d) >= %d: %s" % (first, first + n, value, words[first:first + n]))
                        return abi.RT_ERR_TIMEOUT
                    cond.wait(left)
            return 0

        def rt_last_error(self):
            return b"timeout"

    the_lib = Lib()
    monkeypatch.setattr(D, "lib", lambda: the_lib)
    monkeypatch.setattr(D, "check", lambda rc: (_ for _ in ()).throw(RuntimeError("rc %d" % rc)) if rc < 0 else rc)
    monkeypatch.setattr(D.PeerFrame, "_stream", lambda self: None)
    monkeypatch.setattr(D.torch, "as_tensor", lambda *a, **k: "frame")

    def rank_main(rank):
        rng = random.Random(rank)
        f = D.PeerFrame(0x1000, 8, 8, rank, world, rank, imported=rank != 0)
        try:
            for i in range(1, frames + 1):
                time.sleep(rng.random() * 0.002)        # L2 flush, host work between steps
                f.begin()
                if rank != 0 and i > 1:                 # a peer may not touch the pixels before rank 0 has consumed frame i - 1
                    assert words[abi.RT_FRAME_WORD_RELEASED] >= i - 1
                f.align()
                time.sleep(rng.random() * 0.003)        # the render call
                rendered[rank] = i
                full = f.collect()
                if rank == 0:
                    assert full == "frame"
                    assert all(r >= i for r in rendered), (i, list(rendered))  # every rank's rows of frame i are in
                f.release()
        except Exception as e:  # noqa: BLE001
            errors.append("rank %d: %r" % (rank, e))

    threads = [threading.Thread(target=rank_main, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(30)
    assert not errors, errors
    assert rendered == [frames] * world and words[abi.RT_FRAME_WORD_RELEASED] == frames
