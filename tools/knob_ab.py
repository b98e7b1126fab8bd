"""Developer tool: device time of rank 0's share of a c4 frame for a list of env-knob settings, in
one process.   python tools/knob_ab.py WORLDS "K=V,K2=V2" "K=V" ...   (WORLDS e.g. 1,8; "" = defaults)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rust_tracer_b200 as rt
worlds = [int(x) for x in sys.argv[1].split(",")]
sc = rt.Scene(rt.SceneDesc.builtin("m10k"))
for cfg in sys.argv[2:] or [""]:
    kv = dict(x.split("=") for x in cfg.split(",") if x)
    for k, v in kv.items(): os.environ[k] = v
    out = []
    for world in worlds:
        best = None
        for rep in range(4):
            st = sc.render_device(3840, 2160, 5, rank=0, world=world)
            if rep and (best is None or st["total_ms"] < best["total_ms"]): best = st
        out.append("w%d %.2f (sweeps %.2f)" % (world, best["total_ms"], best["sweep_ms"]))
    print("%-44s %s" % (cfg or "defaults", " | ".join(out)), flush=True)
    for k in kv: os.environ.pop(k, None)
