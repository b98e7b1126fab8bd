"""Short m10k render used as the ncu target for the sweep kernel (tools/gpu_round.sh)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rust_tracer_b200 as rt

w, h, depth = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (1920, 1080, 2)))
name = sys.argv[4] if len(sys.argv) > 4 else "m10k"
sc = rt.Scene(rt.SceneDesc.builtin(name))
st = sc.render_device(w, h, depth)
print(name, w, h, depth, "total %.2f ms sweep %.2f ms rays %d exact/ray %.2f" % (
    st["total_ms"], st["sweep_ms"], st["rays"], st["exact_tests"] / max(st["rays"], 1)))
sc.close()
