"""Device timings of the big configs (developer tool)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rust_tracer_b200 as rt
names = sys.argv[1:] or ["s1k", "m10k"]
for name in names:
    sc = rt.Scene(rt.SceneDesc.builtin(name))
    w, h, d = (1920, 1080, 3) if name == "default" else (3840, 2160, 5)
    n = {"default": 6, "s1k": 1001, "m10k": 10001}[name]
    for rep in range(3):
        st = sc.render_device(w, h, d)
    print(name, "total %.1f ms sweep %.1f ms exact/ray %.1f pairs/s %.3fT Mrays/s %.1f" % (
        st["total_ms"], st["sweep_ms"], st["exact_tests"] / st["rays"], st["rays"] * n / st["sweep_ms"] / 1e9,
        st["rays"] / st["total_ms"] / 1e3), flush=True)
    sc.close()
