"""Phase times of rt_scene_create / rt_render / rt_scene_destroy on the c4 scene (developer tool)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import rust_tracer_b200 as rt
desc = rt.SceneDesc.builtin("m10k")
buf = torch.empty((2160, 3840, 3), dtype=torch.uint8).pin_memory()
for rep in range(3):
    if rep == 2:
        os.environ["RT_B200_TRACE"] = "1"
    t0 = time.perf_counter(); sc = rt.Scene(desc); t1 = time.perf_counter()
    st = sc.render(3840, 2160, 5, want_f32=False, out_u8=buf.numpy())[2]; t2 = time.perf_counter()
    os.environ.pop("RT_B200_TRACE", None)
    sc.close(); t3 = time.perf_counter()
    print("create %.2f ms  render (host RGB8) %.2f ms (device %.2f)  destroy %.2f ms" % ((t1-t0)*1e3, (t2-t1)*1e3, st["total_ms"], (t3-t2)*1e3))
