"""Per-sweep-launch timings of one rank's share of a c4 frame (developer tool)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["RT_B200_TRACE"] = "1"
import rust_tracer_b200 as rt
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
sc = rt.Scene(rt.SceneDesc.builtin("m10k"))
for rep in range(2):
    print("--- rep", rep, file=sys.stderr)
    st = sc.render_device(3840, 2160, 5, rank=0, world=world)
print("world", world, "total %.2f sweep %.2f" % (st["total_ms"], st["sweep_ms"]))
