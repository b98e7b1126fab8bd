"""Device time of one workload's frame under scheduling knobs (rt_debug_set), for A/B runs on the GPU box.

    python tools/frame_time.py c4 [lib=build/variants/X/librt_b200.so] [world=8 rank=0] [knob=value ...] [-- knob=value ...]
lib= loads another build of the library (tools/build_variant.sh) for the whole run.  Each `--` starts another variant; prints ms per frame (total, sweeps) and the sweeps' counters for every variant.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import rust_tracer_b200 as rt  # noqa: E402


def main():
    wl = bench.WORKLOADS[sys.argv[1]]
    variants, cur = [], {}
    for a in sys.argv[2:]:
        if a.startswith("lib="):
            from rust_tracer_b200 import _lib
            _lib.LIB_PATH = os.path.join(ROOT, a[4:])
            print("library:", a[4:], flush=True)
        elif a == "--":
            variants.append(cur)
            cur = {}
        else:
            k, v = a.split("=")
            cur[k] = int(v)
    variants.append(cur)
    import torch
    desc = rt.SceneDesc.builtin(wl["scene"])
    out = torch.empty((wl["height"], wl["width"], 3), dtype=torch.uint8, device="cuda")
    for knobs in variants:
        knobs = dict(knobs)
        world, rank, reps = knobs.pop("world", 1), knobs.pop("rank", 0), knobs.pop("reps", 5)
        with rt.tuned(**knobs):
            sc = rt.Scene(desc)
            best = None
            for _ in range(reps):
                st = sc.render_device(wl["width"], wl["height"], wl["depth"], d_rgb_u8=out.data_ptr(), spp=wl["spp"], seed=1,
                                      rank=rank, world=world, rows_per_tile=8)
                if best is None or st["total_ms"] < best["total_ms"]:
                    best = st
            sc.close()
        print("%-44s frame %.3f ms  sweeps %.3f ms  exact/ray %.2f  slot evals/ray %.0f  hier evals/ray %.0f  launches %d" % (
            " ".join("%s=%s" % kv for kv in knobs.items()) + (" world=%d" % world if world > 1 else "") or "(defaults)",
            best["total_ms"], best["sweep_ms"], (best["exact_tests"] - best["plane_tests"]) / best["rays"],
            (best["bound_evals_sphere"] + best["bound_evals_triangle"] + best["bound_evals_cube"]) / best["rays"],
            best["hier_evals"] / best["rays"], best["kernel_launches"]), flush=True)


if __name__ == "__main__":
    main()
