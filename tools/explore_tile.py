"""Developer tool: what one rank's share of a c4 frame costs on ONE GPU, for every rank of an
N-way split, over tile heights and sweep knobs (stand-in for an N-GPU run: the frame time of the
N-GPU job is the slowest rank's).   python tools/explore_tile.py [world]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rust_tracer_b200 as rt

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
sc = rt.Scene(rt.SceneDesc.builtin("m10k"))
W, H, D = 3840, 2160, 5


def tile_ms(rank, rows_per_tile, reps=3):
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
    print("%-34s tile %3d: max %.2f mean %.2f ms | sweeps mean %.2f | rays min %.2fM max %.2fM | per-rank %s" % (
        label, rows_per_tile, max(tot), sum(tot) / world, sum(sw) / world, min(rays) / 1e6, max(rays) / 1e6,
        " ".join("%.2f" % t for t in tot)), flush=True)


def with_env(**kv):
    for k, v in kv.items():
        if v is None: os.environ.pop(k, None)
        else: os.environ[k] = str(v)


full = tile_ms(0, 8) if world == 1 else None
one = rt.Scene(rt.SceneDesc.builtin("m10k"))
st = None
for rep in range(3):
    st = one.render_device(W, H, D)
print("whole frame: %.2f ms (sweeps %.2f), /%d = %.2f ms" % (st["total_ms"], st["sweep_ms"], world, st["total_ms"] / world), flush=True)
one.close()

for rpt in (None, 2):
    with_env(RT_B200_RAYS_PER_THREAD=rpt)
    for tr in (8, 16, 32, 64):
        all_ranks(tr, "rays/thread %s" % (rpt or "auto"))
with_env(RT_B200_RAYS_PER_THREAD=None)
for want in (2, 8, 16):
    with_env(RT_B200_WANT=want)
    all_ranks(8, "WANT %d" % want)
with_env(RT_B200_WANT=None)
for im in (8, 24, 32):
    with_env(RT_B200_ITEM_MAX=im)
    all_ranks(8, "ITEM_MAX %d" % im)
with_env(RT_B200_ITEM_MAX=None)
