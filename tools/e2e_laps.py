"""Developer tool: wall-clock laps of the one-shot multi-device call (rt_multi_create / rt_multi_render into a pinned
host frame / rt_multi_destroy) on the first N devices, c4.   python tools/e2e_laps.py [N ...]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import rust_tracer_b200 as rt  # noqa: E402

W, H, D = 3840, 2160, 5
desc = rt.SceneDesc.builtin("m10k")
frame = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory().numpy()
for n in [int(a) for a in sys.argv[1:]] or [1]:
    if n > rt.device_count():
        break
    laps = []
    for rep in range(6):
        t0 = time.perf_counter()
        ms = rt.MultiScene(desc, list(range(n)))
        t1 = time.perf_counter()
        _, _, st = ms.render(W, H, D, want_f32=False, out_u8=frame)
        t2 = time.perf_counter()
        ms.close()
        t3 = time.perf_counter()
        laps.append((t1 - t0, t2 - t1, t3 - t2, st["total_ms"]))
    best = min(laps[2:], key=lambda x: x[0] + x[1] + x[2])
    print("%d devices: create %.2f ms, render %.2f ms (device %.2f ms), destroy %.2f ms, sum %.2f ms" % (
        n, best[0] * 1e3, best[1] * 1e3, best[3], best[2] * 1e3, sum(best[:3]) * 1e3), flush=True)
    for knobs in ({"trace": 1},):
        with rt.tuned(**knobs):
            ms = rt.MultiScene(desc, list(range(n)))
            ms.render(W, H, D, want_f32=False, out_u8=frame)
            ms.close()
