"""torchrun --nproc-per-node N tools/check_multi_gpu.py : the row-tiled frame of N GPUs -- collected
through the peer-mapped frame (PeerFrame) and through the NCCL gather -- must equal rank 0's own
single-GPU frame bit for bit (pixels are independent; collecting the rows is the only exchange,
rust_tracer_b200/distributed.py)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import rust_tracer_b200 as rt
from rust_tracer_b200.distributed import PeerFrame, gather_rows

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
ok = True
for name, w, h, depth, spp in (("default", 1920, 1080, 3, 1), ("m10k", 960, 540, 5, 1), ("s1k", 640, 360, 8, 4)):
    sc = rt.Scene(rt.SceneDesc.builtin(name), device=local)
    n_rows = len(rt.tile_rows(h, rank, world, 8))
    tile = torch.empty((n_rows, w, 3), dtype=torch.uint8, device=dev)
    sc.render_device(w, h, depth, d_rgb_u8=tile.data_ptr(), spp=spp, seed=3, rank=rank, world=world, rows_per_tile=8)
    full = gather_rows(tile, h, w, rank, world, 8)
    # the same through the peer-mapped frame: every rank stores its rows into rank 0's buffer
    frame = PeerFrame.create(h, w, rank, world, local) if world > 1 else None
    same_peer = None
    # two frames in a row (different depths): the second may only overwrite the first after rank 0's release()
    for d in ((1, depth) if frame is not None else ()):
        frame.begin()
        sc.render_device(w, h, d, d_rgb_u8=frame.ptr, spp=spp, seed=3, rank=rank, world=world, rows_per_tile=8,
                         flags=frame.flags)
        peer = frame.collect()
        if rank == 0:
            ref = torch.empty((h, w, 3), dtype=torch.uint8, device=dev)
            sc.render_device(w, h, d, d_rgb_u8=ref.data_ptr(), spp=spp, seed=3)
            same_peer = (same_peer is not False) and bool((peer == ref).all())
        frame.release()
    if rank == 0:
        ref = torch.empty((h, w, 3), dtype=torch.uint8, device=dev)
        sc.render_device(w, h, depth, d_rgb_u8=ref.data_ptr(), spp=spp, seed=3)
        same = bool((full == ref).all())
        ok &= same and same_peer is not False
        print("%s %dx%d d%d spp%d world %d: gathered == single-GPU frame: %s; peer-mapped frame: %s" % (
            name, w, h, depth, spp, world, same, same_peer), flush=True)
    if frame is not None:
        if world > 1:
            dist.barrier()
        frame.close()
    sc.close()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
sys.exit(0 if ok else 1)
