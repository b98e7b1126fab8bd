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
