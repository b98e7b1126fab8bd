"""Developer tool: cost of CONTIGUOUS row bands of a c4 frame on one GPU (ms and ms per million rays, band by
band), against the interleaved 8-row tiles of the same share.   python tools/explore_bands.py [bands]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rust_tracer_b200 as rt  # noqa: E402

bands = int(sys.argv[1]) if len(sys.argv) > 1 else 8
sc = rt.Scene(rt.SceneDesc.builtin("m10k"))
W, H, D = 3840, 2160, 5


def best_of(rank, world, rows_per_tile, reps=2):
    best = None
    for rep in range(reps + 1):
        st = sc.render_device(W, H, D, rank=rank, world=world, rows_per_tile=rows_per_tile)
        if rep and (best is None or st["total_ms"] < best["total_ms"]):
            best = st
    return best


def show(label, world, rows_per_tile):
    ms, rays = [], []
    for r in range(world):
        st = best_of(r, world, rows_per_tile)
        ms.append(st["total_ms"]); rays.append(st["rays"] / 1e6)
    print("%-26s sum %.2f ms  max %.2f  | ms: %s | Mrays: %s | ms/Mray: %s" % (
        label, sum(ms), max(ms), " ".join("%.2f" % t for t in ms), " ".join("%.1f" % t for t in rays),
        " ".join("%.3f" % (t / max(r, 1e-9)) for t, r in zip(ms, rays))), flush=True)


st = best_of(0, 1, 8)
print("whole frame %.2f ms, %.1f Mrays, %.3f ms/Mray" % (st["total_ms"], st["rays"] / 1e6, st["total_ms"] / (st["rays"] / 1e6)), flush=True)
show("interleaved 8-row tiles", bands, 8)
show("contiguous bands", bands, (H + bands - 1) // bands)
for nb in (6, 7):
    with rt.tuned(node_bits=nb):
        show("contiguous, node_bits=%d" % nb, bands, (H + bands - 1) // bands)
show("contiguous, 2x bands", 2 * bands, (H + 2 * bands - 1) // (2 * bands))
