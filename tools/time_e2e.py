"""Wall-clock breakdown of the host-buffer call sequence of bench.py's e2e leg (developer tool)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import rust_tracer_b200 as rt
desc = rt.SceneDesc.builtin("m10k")
buf = torch.empty((2160, 3840, 3), dtype=torch.uint8).pin_memory()
for rep in range(4):
    t0 = time.perf_counter(); sc = rt.Scene(desc); t1 = time.perf_counter()
    st = sc.render_device(3840, 2160, 5); t2 = time.perf_counter()
    sc.close(); t3 = time.perf_counter()
    print("create %.2f ms  render_device %.2f ms (device %.2f)  destroy %.2f ms" % ((t1-t0)*1e3, (t2-t1)*1e3, st["total_ms"], (t3-t2)*1e3))
for rep in range(4):
    t0 = time.perf_counter(); sc = rt.Scene(desc); t1 = time.perf_counter()
    st = sc.render(3840, 2160, 5, want_f32=False, out_u8=buf.numpy()); t2 = time.perf_counter()
    sc.close(); t3 = time.perf_counter()
    print("create %.2f ms  render (host RGB8) %.2f ms  destroy %.2f ms  total %.2f ms" % ((t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, (t3-t0)*1e3))
