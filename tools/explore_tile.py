"""Developer tool: what every rank's share of a c4 frame costs on ONE GPU, for an N-way split, over tile
heights and a few sweep knobs (stand-in for an N-GPU run: the frame time of the N-GPU job is the slowest
rank's).   python tools/explore_tile.py [world]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rust_tracer_b200 as rt  # noqa: E402

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
sc = rt.Scene(rt.SceneDesc.builtin("m10k"))
W, H, D = 3840, 2160, 5


def tile_ms(rank, rows_per_tile, reps=2):
    best = None
    for rep in range(reps + 1):
        st = sc.render_device(W, H, D, rank=rank, world=world, rows_per_tile=rows_per_tile)
        if rep and (best is None or st["total_ms"] < best["total_ms"]):
            best = st
    return best


def all_ranks(rows_per_tile, label):
    tot, sw, rays = [], [], []
    for r in range(world):
        st = tile_ms(r, rows_per_tile)
        tot.append(st["total_ms"]); sw.append(st["sweep_ms"]); rays.append(st["rays"])
    print("%-28s tile %3d: max %.2f mean %.2f ms | sweeps mean %.2f | rays min %.2fM max %.2fM | per-rank %s" % (
        label, rows_per_tile, max(tot), sum(tot) / world, sum(sw) / world, min(rays) / 1e6, max(rays) / 1e6,
        " ".join("%.2f" % t for t in tot)), flush=True)


st = None
for rep in range(3):
    st = sc.render_device(W, H, D)
print("whole frame: %.2f ms (sweeps %.2f), /%d = %.2f ms" % (st["total_ms"], st["sweep_ms"], world, st["total_ms"] / world), flush=True)
for tr in (4, 8, 16, 32, 64, 135):
    all_ranks(tr, "defaults")
for knobs in ({"tail_pct": 0}, {"want_light": 2}, {"node_bits": 7}, {"tail_pct": 0, "want_light": 2, "node_bits": 7}):
    with rt.tuned(**knobs):
        for tr in (8, 32):
            all_ranks(tr, " ".join("%s=%s" % kv for kv in knobs.items()))
