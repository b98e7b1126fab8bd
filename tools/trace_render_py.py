"""Where the wall-clock of Scene.render goes on the Python side (developer tool)."""
import os, sys, time, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import rust_tracer_b200 as rt
from rust_tracer_b200 import abi
from rust_tracer_b200.api import camera, _params, tile_rows
from rust_tracer_b200._lib import lib, check
desc = rt.SceneDesc.builtin("m10k")
buf = torch.empty((2160, 3840, 3), dtype=torch.uint8).pin_memory()
sc = rt.Scene(desc)
L = lib()
for rep in range(4):
    t0 = time.perf_counter()
    cam = camera(3840, 2160); p = _params(5); n_rows = len(tile_rows(2160, 0, 1, 8))
    t1 = time.perf_counter()
    u = np.zeros((n_rows, 3840, 3), dtype=np.uint8)
    t2 = time.perf_counter()
    u = buf.numpy()
    t3 = time.perf_counter()
    st = abi.rt_stats()
    check(L.rt_render(sc._h, C.byref(cam), C.byref(p), None, u.ctypes.data, C.byref(st)))
    t4 = time.perf_counter()
    d = st.as_dict()
    t5 = time.perf_counter()
    print("params %.3f  zeros %.3f  drop+numpy() %.3f  rt_render %.3f (device %.3f)  as_dict %.3f ms" % (
        (t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, (t4-t3)*1e3, d["total_ms"], (t5-t4)*1e3))
for rep in range(3):
    t0 = time.perf_counter(); r = sc.render(3840, 2160, 5, want_f32=False, out_u8=buf.numpy()); t1 = time.perf_counter()
    print("Scene.render %.3f ms (device %.3f)" % ((t1-t0)*1e3, r[2]["total_ms"]))
