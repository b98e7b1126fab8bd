#!/usr/bin/env python
"""bench.py -- the headline measurement of the render path (driver contract).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c3|c2|c5|c5s|c2f] [--impl ours|reference]

A "step" is one full frame of the workload: primary rays -> nearest-hit sweeps -> shading ->
RGB8, rendered by the CUDA path of librt_b200.so.  With N > 1 (one process per GPU, torchrun)
the frame is split into interleaved row tiles and the finished tiles are gathered to rank 0
over NCCL inside the step (strong scaling: the frame is fixed).

Prints ONE JSON line (rank 0).  `value` = Mrays/s with the scene resident in HBM, timed with
CUDA events on the launching stream; `e2e` = the same metric through the host-buffer call of the
C ABI (scene upload + render + RGB8 read-back inside the timed region; at N > 1 that is ONE
rt_render_multi call from rank 0 driving all N GPUs).  `frame_crc32` is the CRC of the RGB8 frame
rank 0 collected in the last timed step: the same at every N.  The oracle is executed only by the
cpu_baseline leg and by --impl reference (it is never on the product path), and --impl reference
never loads the product library.  --workload c2f times the RayForest calls instead (build, eval,
filtered eval: the reference's `bench --filter`, src/main.rs:153-216).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# BASELINE.json configs.  c4 is the north_star target ("4K, 10k-primitive, depth-5 ... on 1
# GPU, >= 7x at 8 GPUs") and fits one GPU, so it is the default at every N.
WORKLOADS = {
    "c2": dict(scene="default", width=1920, height=1080, depth=3, spp=1, desc="my_scene.rs 1920x1080 depth 3"),
    "c3": dict(scene="s1k", width=3840, height=2160, depth=5, spp=1, desc="1,000 spheres + plane, 4 lights, 3840x2160 depth 5"),
    "c4": dict(scene="m10k", width=3840, height=2160, depth=5, spp=1, desc="10k mixed spheres/triangles, 4 lights, 3840x2160 depth 5"),
    "c5": dict(scene="m10k", width=7680, height=4320, depth=8, spp=4, desc="8K, 10k prims, 4 spp jittered, depth 8 (whole-box config)"),
    "c5s": dict(scene="m10k", width=3840, height=2160, depth=8, spp=4, desc="config 5 at 4K instead of 8K: 10k prims, 4 spp, depth 8"),
    "c2f": dict(scene="default", width=1920, height=1080, depth=8, spp=1, desc="RayForest: my_scene.rs 1920x1080 depth 8, build / eval / eval filtered on the blue sphere"),
}
ROWS_PER_TILE = 8
L2_NOTE = "256 MiB buffer written between timed steps; ray pools (>1 GB) exceed L2 anyway"


def base_config(args, wl):
    """The same keys and values in both arms (--impl ours / reference)."""
    return {"workload": args.workload + ": " + wl["desc"], "scene": wl["scene"], "width": wl["width"], "height": wl["height"],
            "depth": wl["depth"], "spp": wl["spp"], "rows_per_tile": ROWS_PER_TILE, "l2": L2_NOTE}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU every 100 ms while the timed region runs."""

    REASONS = {  # nvmlClocksEventReason* bits
        0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
        0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
        0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
    }

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        while not self.stop_flag and self.nv is not None:
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                bits = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for b, name in self.REASONS.items():
                    if bits & b and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.1)

    def result(self):
        self.stop_flag = True
        if self.is_alive():
            self.join(timeout=1.0)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def oracle_rate(desc, wl, seconds, threads):
    """Times the CPU oracle (the port of the reference's own algorithm) on a bounded row sample
    of the workload; returns (Mrays/s, description, rays, seconds).  desc: an rt_scene_desc struct."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc

    cam = orc.abi.default_camera(wl["width"], wl["height"])
    h = wl["height"]
    probe = np.array([h // 3, (2 * h) // 3], dtype=np.int32)
    t0 = time.perf_counter()
    _, _, c = orc.render(desc, cam, wl["depth"], rows=probe, spp=wl["spp"], seed=1, threads=threads, want_f32=False)
    dt = max(time.perf_counter() - t0, 1e-6)
    n_rows = int(max(threads, min(h // 2, round(len(probe) * seconds / dt))))
    rows = np.unique(np.linspace(7, h - 1, n_rows).astype(np.int32))
    t0 = time.perf_counter()
    _, _, c = orc.render(desc, cam, wl["depth"], rows=rows, spp=wl["spp"], seed=1, threads=threads, want_f32=False)
    dt = time.perf_counter() - t0
    sample = "%d of %d rows of the %dx%d frame (%d rays, %.1f s)" % (len(rows), h, wl["width"], h, c.rays, dt)
    return c.rays / dt / 1e6, sample, int(c.rays), dt


def oracle_one_thread(desc, wl):
    """The oracle on one thread over the middle row of the frame: (Mrays/s, description)."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc

    cam = orc.abi.default_camera(wl["width"], wl["height"])
    rows = np.array([wl["height"] // 2], dtype=np.int32)
    t0 = time.perf_counter()
    _, _, c = orc.render(desc, cam, wl["depth"], rows=rows, spp=wl["spp"], seed=1, threads=1, want_f32=False)
    dt = max(time.perf_counter() - t0, 1e-6)
    return c.rays / dt / 1e6, "row %d of the %dx%d frame (%d rays, %.1f s)" % (rows[0], wl["width"], wl["height"], c.rays, dt)


_REAL_STDOUT = None


def quiet_stdout():
    """The contract is ONE JSON line on stdout; libraries (NCCL prints its version banner on
    fd 1) must not add to it.  Everything written to fd 1 from here on goes to stderr; emit()
    writes to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = json.dumps(obj) + "\n"
    if _REAL_STDOUT is None:
        sys.stdout.write(line)
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, line.encode())


def run_reference(args, wl):
    """--impl reference: the reference's own CPU algorithm (the oracle port -- the Rust crate
    cannot be built here) on all host threads; each step renders a bounded row
    sample of the same workload.  Rank 0 only.  The product package is never imported here: the
    scene comes from the committed scene description files under tests/golden/ (or, for the
    reference's own scene, from the oracle's restatement of src/my_scene.rs)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    os.environ["RT_ORC_STANDALONE"] = "1"
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc
    assert "rust_tracer_b200" not in sys.modules
    if wl["scene"] == "default":
        arrays = orc.default_scene()
    else:
        arrays = orc.load_rtscene(os.path.join(ROOT, "tests", "golden", wl["scene"] + ".rtscene.xz"))
    threads = os.cpu_count() or 1
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    rates, sample = [], ""
    for i in range(args.warmup + args.steps):
        _, sample, rays, dt = oracle_rate(arrays.desc, wl, per_step, threads)
        if i >= args.warmup:
            rates.append((rays, dt))
    total_rays = sum(r for r, _ in rates)
    total_s = sum(t for _, t in rates)
    value = total_rays / total_s / 1e6
    with open("/proc/self/maps") as fh:
        assert "librt_b200" not in fh.read() and "rust_tracer_b200" not in sys.modules, "the reference arm must not load the product"
    emit({
        "impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_s / max(1, args.steps),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": base_config(args, wl),
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    })


def scene_bytes(desc):
    """Bytes of the flat scene description rt_scene_create copies to the device."""
    import ctypes as C
    from rust_tracer_b200 import abi
    d = desc.desc
    return (d.n_shapes * C.sizeof(abi.rt_shape) + d.n_materials * C.sizeof(abi.rt_material)
            + d.n_lights * C.sizeof(abi.rt_light) + 12)


def run_ours(args, wl):
    import numpy as np
    import torch
    import torch.distributed as dist

    import rust_tracer_b200 as rt
    from rust_tracer_b200.distributed import PeerFrame, gather_rows

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the render path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    w, h, depth, spp = wl["width"], wl["height"], wl["depth"], wl["spp"]
    desc = rt.SceneDesc.builtin(wl["scene"])
    scene = rt.Scene(desc, device=local)
    stream = torch.cuda.Stream(device=dev)
    scene.set_stream(stream.cuda_stream)
    n_rows = len(rt.tile_rows(h, rank, world, ROWS_PER_TILE))
    tile = torch.empty((n_rows, w, 3), dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > the 126 MB L2
    pinned = torch.empty((h, w, 3), dtype=torch.uint8).pin_memory() if rank == 0 else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # N > 1: the frame lives on rank 0's GPU, mapped into every rank; the tiles are written straight into
    # it over NVLink (RT_FLAG_FRAME_ROWS); completion travels through signal words in the same memory (PeerFrame).  NCCL send/recv where mapping fails.
    frame = PeerFrame.create(h, w, rank, world, local) if world > 1 and not os.environ.get("RT_B200_NCCL_GATHER") else None

    frame_kind = "peer" if frame is not None else "nccl"

    def step(begun=False):
        """One frame with the scene resident: render this rank's row tiles, collect the frame on rank 0.
        begun: the caller has already called frame.begin() for this frame."""
        with torch.cuda.stream(stream):
            if frame is not None:
                if not begun:
                    frame.begin()
                st = scene.render_device(w, h, depth, d_rgb_u8=frame.ptr, spp=spp, seed=1, rank=rank, world=world,
                                         rows_per_tile=ROWS_PER_TILE, flags=frame.flags)
                full = frame.collect()  # rank 0: returns once every rank's rows are in its frame
                frame.release()
            else:
                st = scene.render_device(w, h, depth, d_rgb_u8=tile.data_ptr(), spp=spp, seed=1, rank=rank, world=world,
                                         rows_per_tile=ROWS_PER_TILE)
                full = gather_rows(tile, h, w, rank, world, ROWS_PER_TILE)
        return st, full

    for _ in range(args.warmup):
        flush.zero_()
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    spans, stats_list = [], []
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if frame is not None:
            # N > 1: a device-side start line.  Without it a peer's span would begin before it may touch the
            # frame (it first waits for rank 0 to release the previous one) and so include rank 0's L2 flush
            # and host work BETWEEN steps, which the N = 1 span above excludes too.  With it every rank starts
            # frame i together, and the longest span over the ranks is the time of the frame.
            with torch.cuda.stream(stream):
                frame.begin()
                frame.align()
        e0.record(stream)
        st, full = step(begun=frame is not None)
        e1.record(stream)
        spans.append((e0, e1))
        stats_list.append(st)
    barrier()
    clocks = sampler.result()
    # CRC of the RGB8 frame rank 0 holds after the last timed step: the same number at every N
    frame_crc32 = None
    if rank == 0:
        import zlib
        last = full if world > 1 else tile
        frame_crc32 = zlib.crc32(last.cpu().numpy().tobytes()) & 0xFFFFFFFF
    barrier()
    dev_ms = sum(a.elapsed_time(b) for a, b in spans)
    if os.environ.get("RT_B200_BENCH_DEBUG"):  # per-rank split: the library's own device time vs the whole step
        sys.stderr.write("[bench] rank %d: render %.3f ms/step (sweeps %.3f), step incl. gather %.3f ms\n" % (
            rank, sum(s["total_ms"] for s in stats_list) / max(1, args.steps),
            sum(s["sweep_ms"] for s in stats_list) / max(1, args.steps), dev_ms / max(1, args.steps)))

    # whole-job numbers: rays of all ranks / max-over-ranks device time
    rays_step = float(sum(s["rays"] for s in stats_list)) / max(1, args.steps)
    flops_step = float(sum(s["algorithmic_flops"] for s in stats_list)) / max(1, args.steps)
    agg = torch.tensor([rays_step, flops_step, float(stats_list[-1]["kernel_launches"])], dtype=torch.float64, device=dev)
    tmax = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(agg, op=dist.ReduceOp.SUM)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    rays_step, flops_step, launches = (float(x) for x in agg.tolist())
    dev_ms = float(tmax.item())
    ms_per_step = dev_ms / max(1, args.steps)
    value = rays_step / (ms_per_step * 1e-3) / 1e6

    # ---- e2e: the host-buffer call a user of the reference would make -----------------------
    # every step: scene description in (H2D), RGB8 frame out on the host (D2H), through the C ABI.
    #   N = 1: rt_scene_create + rt_render + rt_scene_destroy
    #   N > 1: ONE rt_render_multi call from rank 0 -- the library builds the tables once, drives all N
    #          devices from its own worker threads and every device copies its rows home; the other
    #          ranks of the torchrun job have released their GPUs' work and wait on the TCP store
    #          (a host-side wait: an NCCL barrier would park a spinning kernel on every GPU).
    scene.close()
    del tile
    if frame is not None:
        barrier()
        frame.close()
        frame = None
    torch.cuda.empty_cache()
    h2d = scene_bytes(desc) * world
    d2h = w * h * 3
    e2e_steps = max(1, min(args.steps, 3))
    e2e_crc32 = None
    e2e_s = 0.0
    store = dist.distributed_c10d._get_default_store() if world > 1 else None
    if rank == 0:
        import zlib
        devices = list(range(world))

        def e2e_step():
            if world == 1:
                sc = rt.Scene(desc, device=local)
                sc.render(w, h, depth, spp=spp, seed=1, want_f32=False, out_u8=pinned.numpy())
                sc.close()
            else:
                rt.render_multi(desc, devices, w, h, depth, spp=spp, seed=1, rows_per_tile=ROWS_PER_TILE, out_u8=pinned.numpy())

        e2e_step()  # warm-up (contexts, worker threads, pool caches on every device)
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        e2e_s = time.perf_counter() - t0
        e2e_crc32 = zlib.crc32(pinned.numpy().tobytes()) & 0xFFFFFFFF
        if store is not None:
            store.set("rt_b200_e2e_done", "1")
    elif store is not None:
        store.wait(["rt_b200_e2e_done"])
    e2e_value = rays_step * e2e_steps / e2e_s / 1e6 if e2e_s > 0 else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (sweep_kernel): FP32, not HBM and not tensor --------
    # frac = what the FP32 pipe was really asked to do / peak (culling is subtracted -- a pair
    # rejected by a 4- or 7-FMA bounding-sphere test earns those FMAs, not an intersection test):
    #   executed = filter flops (2 x 4|7 FMA per sphere evaluated: bound slots, group and quad spheres)
    #            + exact tests x their 8d price (54 sphere, 30 triangle, 393 cube, 47|14 plane)
    n_lights = desc.desc.n_lights
    sweep_ms = sum(s["sweep_ms"] for s in stats_list) / max(1, args.steps)          # rank 0, CUDA events on its stream
    total_ms = sum(s["total_ms"] for s in stats_list) / max(1, args.steps)
    r0 = stats_list[-1]
    all_pairs_flops = r0["algorithmic_flops"] - r0["nodes"] * (60.0 + 70.0 * n_lights)   # every pair, rank 0's share
    pairs = r0["sphere_tests"] + r0["triangle_tests"] + r0["cube_tests"] + r0["plane_tests"]   # rays x shapes
    tested_share = r0["tested_pairs"] / pairs if pairs else 0.0
    measured_peak, sms, khz = rt.measure_fp32_peak(local)
    nominal_peak = sms * 128 * 2 * khz * 1e3 / 1e12
    secs = sweep_ms * 1e-3
    executed = r0["executed_flops"] / secs / 1e12 if secs > 0 else 0.0
    credit = all_pairs_flops * tested_share / secs / 1e12 if secs > 0 else 0.0
    all_pairs = all_pairs_flops / secs / 1e12 if secs > 0 else 0.0
    traffic, ncu = None, None  # from the committed ncu --set full capture of the longest sweep launch
    try:
        with open(os.path.join(ROOT, "profiles", "sweep_ncu_summary.json")) as fh:
            prof = json.load(fh)
        if prof.get("workload") == args.workload:
            traffic = prof.get("dram_bytes_per_launch")
            ncu = {k: prof.get(k) for k in ("fma_pipe_pct", "issue_active_pct", "warps_active_pct", "kernel_ms_avg",
                                            "launches_per_step", "duration", "kernel", "code_state")}
    except (OSError, ValueError):
        pass
    roofline = {
        "bound": "fp32", "kernel": "sweep_kernel", "achieved": executed, "peak": measured_peak, "unit": "TFLOP/s",
        "frac": executed / measured_peak if measured_peak else None, "traffic": traffic,
        "peak_source": "of measured: rt_measure_fp32_peak FFMA micro-kernel on this GPU (MEASURED_PEAKS.json has no FP32 "
                       "figure); nominal %d SMs x 128 x 2 x %.0f MHz = %.1f TFLOP/s" % (sms, khz / 1e3, nominal_peak),
        "achieved_is": "flop the sweeps executed / sweep time from CUDA events: 2 x (4 or 7 FMA) per bounding sphere the filters "
                       "evaluated (bound slots, group and quad spheres; rt_stats.filter_flops) + exact tests x their SURVEY 8(d) "
                       "price (rt_stats.exact_flops); pairs culled wholesale earn nothing, pairs rejected by their own bound earn "
                       "the 4|7 FMA that rejected them",
        "filter_tflops": r0["filter_flops"] / secs / 1e12 if secs > 0 else None,
        "exact_tflops": r0["exact_flops"] / secs / 1e12 if secs > 0 else None,
        "counts_per_step": {k: r0[k] for k in ("rays", "bound_evals_sphere", "bound_evals_triangle", "bound_evals_cube", "hier_evals",
                                               "exact_sphere_tests", "exact_triangle_tests", "exact_cube_tests", "plane_tests")},
        "ncu": ncu,
        "algorithmic_credit_frac": credit / measured_peak if measured_peak else None,
        "algorithmic_credit_is": "NOT the roofline fraction: SURVEY 8(d) price of a full intersection test (54 / 30 / 393) for every "
                                 "pair whose own bound was evaluated or that was tested exactly (rt_stats.tested_pairs), / peak",
        "tested_share_of_pairs": tested_share,
        "all_pairs_equivalent_tflops": all_pairs,
        "all_pairs_equivalent_is": "the 8(d) price for EVERY ray x shape pair / sweep time, i.e. the rate a brute-force scan would need",
        "pairs_per_s": pairs / secs if secs > 0 else None,
        "sweep_ms_per_step": sweep_ms, "sweep_share_of_step": sweep_ms / total_ms if total_ms else None,
    }

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, sample, _, _ = oracle_rate(desc.desc, wl, args.cpu_seconds, threads)
        cpu = {"value": v, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample}
        # the reference itself is single-threaded (Rc<RefCell<..>> everywhere, src/render.rs:31-38): the same
        # port on ONE thread, on one row of the frame (SURVEY 8d asks for both figures)
        v1, sample1 = oracle_one_thread(desc.desc, wl)
        cpu["one_thread"] = {"value": v1, "unit": "Mrays/s", "cores": 1, "sample": sample1}

    partition = ("single GPU" if world == 1 else "interleaved row tiles, " + (
        "written straight into rank 0's frame over NVLink (peer-mapped, RT_FLAG_FRAME_ROWS); each rank signals a word in that frame, rank 0 waits for all on the device"
        if frame_kind == "peer" else "NCCL send/recv gather to rank 0") + ", inside the step"
        + ("; every timed step starts at a device-side start line over all ranks (PeerFrame.align), outside the span like the L2 flush" if frame_kind == "peer" else ""))
    emit({
        "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": base_config(args, wl),
        "rays_per_frame": rays_step, "partition": partition, "frame_crc32": frame_crc32,
        "roofline": roofline, "cpu_baseline": cpu,
        "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "frame_crc32": e2e_crc32,
                "what": ("rt_scene_create + rt_render into a pinned host RGB8 buffer + rt_scene_destroy" if world == 1 else
                         "one rt_render_multi call on rank 0 (tables built once, uploaded to %d devices, rendered by the library's "
                         "worker threads, every device copies its own rows into the pinned host RGB8 frame)" % world)},
        "gpu_launches": int(launches) * args.steps, "clocks": clocks,
        "frame_ms": ms_per_step,
    })
    if world > 1:
        dist.destroy_process_group()


def run_forest(args, wl):
    """--workload c2f: the RayForest calls the reference's own harness times (src/main.rs:153-216:
    generate_ray_forest, render_forest, render_forest_filter on the blue sphere), one GPU."""
    import numpy as np
    import torch

    import rust_tracer_b200 as rt

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the render path has no CPU fallback")
    if int(os.environ.get("RANK", "0")) != 0:
        return
    w, h, depth = wl["width"], wl["height"], wl["depth"]
    desc = rt.SceneDesc.builtin(wl["scene"])
    scene = rt.Scene(desc, device=0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
    blue = 1  # the shape the reference's `bench --filter` mutates (src/main.rs:176-187: "blue")
    build_ms, eval_ms, filt_ms, rays, nodes, touched = [], [], [], 0, 0, 0
    sampler = None
    for i in range(args.warmup + args.steps):
        if i == args.warmup:
            sampler = ClockSampler(0)
            sampler.start()
        flush.zero_()
        torch.cuda.synchronize()
        forest = scene.build_forest(w, h, depth)
        bs = forest.build_stats
        flush.zero_()
        torch.cuda.synchronize()
        f, u, es = forest.eval()
        flush.zero_()
        torch.cuda.synchronize()
        n, fs = forest.eval_filtered([blue], f, u)
        if i >= args.warmup:
            build_ms.append(bs["total_ms"]); eval_ms.append(es["total_ms"]); filt_ms.append(fs["total_ms"])
            rays, nodes, touched = bs["rays"], bs["nodes"], n
        forest.close()
    clocks = sampler.result() if sampler else None
    scene.close()
    mean = lambda v: float(sum(v) / max(1, len(v)))
    n_point = sum(1 for i in range(desc.desc.n_lights) if desc.desc.lights[i].kind == 0)
    tree_rays = rays - bs["shadow_rays"]
    # algorithmic bytes of one shading pass (shade_kernel): per tree ray its ray, key and parent link
    # (32 + 8 + 4 B); per node its shadow base, two child colours, one shadow key per point light and
    # the colour it hands up (4 + 32 + 8 L + 16 B); the scene tables stay in L2
    eval_bytes = tree_rays * 44 + nodes * (52 + 8 * n_point)
    hbm = None
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            hbm = json.load(fh).get("hbm_gbs")
    except (OSError, ValueError):
        pass
    hbm_src = "MEASURED_PEAKS.json hbm_gbs" if hbm else "fallback 7700 GB/s nominal (B200_PROFILING.md)"
    hbm = hbm or 7700.0
    gbs = eval_bytes / (mean(eval_ms) * 1e-3) / 1e9
    emit({
        "metric": "Mrays/s", "value": rays / (mean(build_ms) * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": mean(build_ms), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": base_config(args, wl),
        "what": "value = rays cast by rt_forest_build / its device time (CUDA events inside the library)",
        "forest": {
            "build_ms": mean(build_ms), "eval_ms": mean(eval_ms), "eval_filtered_ms": mean(filt_ms),
            "trees": w * h, "trees_touched": int(touched), "trees_touched_pct": 100.0 * touched / (w * h),
            "rays": int(rays), "tree_rays": int(tree_rays), "nodes": int(nodes),
            "eval_Mnodes_per_s": nodes / (mean(eval_ms) * 1e-3) / 1e6,
            "eval_filtered_Mtrees_per_s": touched / (mean(filt_ms) * 1e-3) / 1e6 if touched else None,
            "mutated_shape": blue,
        },
        "roofline": {"bound": "hbm", "kernel": "shade_kernel (rt_forest_eval)", "achieved": gbs, "peak": hbm, "unit": "GB/s",
                     "frac": gbs / hbm, "traffic": None, "peak_source": hbm_src,
                     "achieved_is": "algorithmic bytes of one shading pass over the kept trees (44 B per tree ray + (52 + 8 L) B per node) / rt_forest_eval device time"},
        "cpu_baseline": None, "e2e": None, "gpu_launches": None, "clocks": clocks,
    })


def main():
    args = parse_args()
    quiet_stdout()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    elif args.workload == "c2f":
        run_forest(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
