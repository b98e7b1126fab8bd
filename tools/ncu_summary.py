"""Summarise an `ncu --set full` capture of sweep_kernel into profiles/ (run here, no GPU needed).

    python tools/ncu_summary.py gpurun_out/sweep_full.ncu-rep r02 c4 [gpurun_out/launches.csv [frames]] [code-state]
writes profiles/<tag>_sweep_kernel_ncu_full.csv (selected raw metrics, one column per captured
launch), profiles/<tag>_sweep_kernel_hot_loop.txt (SASS of the filter loop with per-instruction
execution counts and stall samples) and profiles/sweep_ncu_summary.json (read by bench.py for
roofline.traffic).
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, tag, workload = sys.argv[1], sys.argv[2], sys.argv[3]
launch_csv = sys.argv[4] if len(sys.argv) > 4 else None  # ncu --metrics gpu__time_duration.sum list of the same command
frames = int(sys.argv[5]) if len(sys.argv) > 5 else 1     # frames that command rendered
code_state = sys.argv[6] if len(sys.argv) > 6 else None

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed.avg.per_cycle_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
]

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, launches = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
out = os.path.join(ROOT, "profiles", "%s_sweep_kernel_ncu_full.csv" % tag)
with open(out, "w", newline="") as fh:
    wr = csv.writer(fh)
    wr.writerow(["metric", "unit"] + ["launch%d" % i for i in range(len(launches))])
    wr.writerow(["Kernel Name", ""] + [l[ix["Kernel Name"]] for l in launches])
    for k in KEYS + sorted(h for h in hdr if "issue_stalled" in h and h.endswith("per_issue_active.ratio")):
        if k in ix:
            wr.writerow([k, units[ix[k]]] + [l[ix[k]] for l in launches])
print("wrote", out)


def num(l, k):
    try:
        return float(l[ix[k]].replace(",", ""))
    except (KeyError, ValueError):
        return None


def to_bytes(l, k):
    v = num(l, k)
    if v is None:
        return None
    u = units[ix[k]].lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)


def launch_list_stats(path):
    """(sweep launches per frame, their mean duration in ms, sweep share of all kernel time) from a launch list."""
    rows = [r for r in csv.reader(open(path)) if r]
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, body = rows[start], rows[start + 1:]
    k, v, u = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    scale = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}
    per = [(r[k], float(r[v].replace(",", "")) * scale.get(r[u], 1e-6)) for r in body if len(r) > v and "fp32_peak" not in r[k]]
    sweeps = [t for name, t in per if "sweep_kernel" in name]
    total = sum(t for _, t in per)
    return len(sweeps) / max(frames, 1), sum(sweeps) / max(len(sweeps), 1), sum(sweeps) / total if total else None


big = max(launches, key=lambda l: num(l, "gpu__time_duration.sum") or 0)
extra = {}
if launch_csv:
    n, mean_ms, share = launch_list_stats(launch_csv)
    extra = {"launches_per_step": n, "kernel_ms_avg": mean_ms, "sweep_share_of_kernel_time": share,
             "launch_list": os.path.basename(launch_csv)}
summary = {
    "workload": workload, "kernel": big[ix["Kernel Name"]], "source": os.path.basename(rep),
    "duration": "%s %s" % (big[ix["gpu__time_duration.sum"]], units[ix["gpu__time_duration.sum"]]),
    "dram_bytes_per_launch": (to_bytes(big, "dram__bytes_read.sum") or 0) + (to_bytes(big, "dram__bytes_write.sum") or 0),
    "fma_pipe_pct": num(big, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
    "issue_active_pct": num(big, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "warps_active_pct": num(big, "sm__warps_active.avg.pct_of_peak_sustained_active"),
    "registers_per_thread": num(big, "launch__registers_per_thread"),
    "code_state": code_state,
    "note": "per-launch figures of the longest captured sweep_kernel launch (ncu --set full, cold cache, serialised)",
}
summary.update(extra)
with open(os.path.join(ROOT, "profiles", "sweep_ncu_summary.json"), "w") as fh:
    json.dump(summary, fh, indent=1)
print(json.dumps(summary))

# ---- hot loop with per-instruction counts
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
blocks = src.split('"Kernel Name"')
rows = list(csv.reader(('"Kernel Name"' + blocks[-1]).splitlines()))
h2 = rows[1]
jx = {h: i for i, h in enumerate(h2)}
data = rows[2:]
cnt = [float(r[jx["Instructions Executed"]] or 0) for r in data]
mx = max(cnt)
hot = [i for i, c in enumerate(cnt) if c >= 0.4 * mx]
lo, hi = min(hot), max(hot)
tot_inst, tot_smp = sum(cnt), sum(float(r[jx["# Samples"]] or 0) for r in data)
with open(os.path.join(ROOT, "profiles", "%s_sweep_kernel_hot_loop.txt" % tag), "w") as fh:
    fh.write("# filter loop of %s (instructions executed >= 40%% of the hottest), from %s\n" % (rows[0][1], os.path.basename(rep)))
    fh.write("# %d of %d SASS instructions; they account for %.1f%% of instructions executed and %.1f%% of stall samples\n" % (
        len(hot), len(data), 100 * sum(cnt[i] for i in hot) / tot_inst,
        100 * sum(float(data[i][jx["# Samples"]] or 0) for i in hot) / tot_smp))
    fh.write("# executed  samples  instruction\n")
    for i in range(lo, hi + 1):
        if cnt[i] >= 0.4 * mx:
            fh.write("%10.3e %7s  %s\n" % (cnt[i], data[i][jx["# Samples"]], data[i][jx["Source"]].strip()))
print("hot loop:", len(hot), "instructions")
