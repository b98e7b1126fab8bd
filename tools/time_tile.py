"""Device time of one rank's share of a c4 frame, untraced (developer tool): python tools/time_tile.py [world]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rust_tracer_b200 as rt
sc = rt.Scene(rt.SceneDesc.builtin("m10k"))
for world in [int(a) for a in sys.argv[1:]] or [1, 2, 4, 8]:
    best = None
    for rep in range(4):
        st = sc.render_device(3840, 2160, 5, rank=0, world=world)
        if rep and (best is None or st["total_ms"] < best["total_ms"]):
            best = st
    print("world %d: total %.2f ms sweeps %.2f ms launches %d  (x%d = %.1f ms)" % (
        world, best["total_ms"], best["sweep_ms"], best["kernel_launches"], world, world * best["total_ms"]))
