"""Developer check run on the GPU box: parity numbers vs the oracle + first timings."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import rust_tracer_b200 as rt
from rust_tracer_b200 import abi
import orc


def compare(name, desc, w, h, depth, flags=0, spp=1, seed=0, mode=0, rows=None):
    sc = rt.Scene(desc)
    t = time.time()
    f, u, st = sc.render(w, h, depth, spp=spp, seed=seed, flags=flags | (abi.RT_FLAG_FOREST_SHADING if mode else 0))
    dt = time.time() - t
    cam = abi.default_camera(w, h)
    of, ou, oc = orc.render(desc.desc, cam, depth, rows=rows, mode=mode, spp=spp, seed=seed)
    if rows is not None:
        f, u = f[rows], u[rows]
    d8 = np.abs(u.astype(int) - ou.astype(int))
    bad = (d8.max(-1) > 1).mean()
    nan_mismatch = (np.isnan(f) != np.isnan(of)).sum()
    fd = np.nanmax(np.abs(f - of)) if f.size else 0
    exact = (u == ou).all(-1).mean()
    print(f"[{name}] {w}x{h} d{depth} flags={flags}: u8 max|d|={d8.max()} frac>1={bad:.2e} exact_px={exact:.6f} "
          f"f32 max|d|={fd:.3e} nan_mismatch={nan_mismatch} | rays gpu={st['rays']} orc={oc.rays if rows is None else 'n/a'} "
          f"nodes gpu={st['nodes']} orc={oc.nodes if rows is None else 'n/a'} | gpu {st['total_ms']:.2f} ms (sweep {st['sweep_ms']:.2f}) "
          f"exact_tests={st['exact_tests']} wall={dt:.2f}s", flush=True)
    sc.close()
    return bad


if __name__ == "__main__":
    print("devices", rt.lib().rt_device_count())
    print("fp32 peak", rt.measure_fp32_peak(0))
    d = rt.SceneDesc.builtin("default")
    compare("default", d, 256, 256, 8)
    compare("default-audit", d, 256, 256, 8, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS)
    compare("default-forest", d, 256, 256, 8, mode=1)
    compare("default-c1", d, 512, 512, 8)
    compare("default-c2", d, 1920, 1080, 3)
    compare("default-spp4", d, 160, 120, 5, spp=4, seed=7)
    s = rt.SceneDesc.synthetic(300, 0, 1000, 0.05, 0.2)
    compare("s300", s, 256, 144, 5)
    compare("s300-audit", s, 256, 144, 5, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS)
    m = rt.SceneDesc.synthetic(400, 400, 10000, 0.03, 0.1)
    compare("m800", m, 256, 144, 5)
    compare("m800-audit", m, 256, 144, 5, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS)
    # timings on the big configs (GPU only)
    for name, w, h, depth in (("s1k", 3840, 2160, 5), ("m10k", 3840, 2160, 5)):
        sc = rt.Scene(rt.SceneDesc.builtin(name))
        for rep in range(2):
            st = sc.render_device(w, h, depth)
            print(f"[{name}] {w}x{h} d{depth}: {st['total_ms']:.1f} ms sweep {st['sweep_ms']:.1f} rays {st['rays']} "
                  f"-> {st['rays']/st['total_ms']/1e3:.1f} Mrays/s, alg {st['algorithmic_flops']/st['total_ms']/1e9:.2f} TFLOP/s "
                  f"exact/ray {st['exact_tests']/st['rays']:.1f}", flush=True)
        sc.close()
