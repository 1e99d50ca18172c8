"""Turns the ncu artefacts of one gpurun call (gpurun_out/) into the tracked summaries under profiles/.

  python tools/summarize_profile.py <tag> [--launches gpurun_out/launches_horse.csv]
                                          [--rep gpurun_out/nn_filter_full.ncu-rep --kernel nn_filter]

Writes  profiles/<tag>_launches.csv       the raw per-launch list (gpu__time_duration.sum)
        profiles/<tag>_launches.md        per-kernel count / mean / share of the step
        profiles/<tag>_<kernel>_metrics.csv  the metrics of the --set full capture the design cites
        profiles/<kernel>_traffic.json    dram bytes per launch (read by bench.py -> roofline.traffic)
"""
import argparse
import collections
import csv
import json
import os
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROF = os.path.join(ROOT, "profiles")

KEEP = ("gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit", "launch__grid_size",
        "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg",
        "sm__pipe_fma_cycles_active.avg", "sm__pipe_fmaheavy_cycles_active.avg", "sm__inst_executed_pipe_alu.avg",
        "sm__pipe_alu_cycles_active.avg", "sm__inst_executed_pipe_lsu.avg", "sm__inst_executed_pipe_fp64.avg",
        "sm__pipe_fp64_cycles_active.avg", "sm__pipe_tensor", "sm__inst_issued.avg", "sm__issue_active.avg",
        "smsp__issue_active.avg", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg",
        "gpu__dram_throughput.avg", "lts__t_bytes.sum", "lts__t_sector_hit_rate", "l1tex__t_sector_hit_rate", "l1tex__throughput.avg", "lts__throughput.avg", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.avg", "smsp__cycles_active.avg", "sm__throughput.avg",
        "smsp__average_warps_issue_stalled", "smsp__thread_inst_executed_per_inst_executed",
        "sm__sass_thread_inst_executed_op_ffma_pred_on.sum", "sm__sass_thread_inst_executed_op_fadd_pred_on.sum",
        "sm__sass_thread_inst_executed_op_fmul_pred_on.sum", "sm__sass_thread_inst_executed_op_dfma_pred_on.sum",
        "derived__sm__sass_thread_inst_executed_op_ffma_pred_on_x2")


def launches(tag, path):
    lines = open(path).read().splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
    rows = list(csv.DictReader(lines[start:]))
    shutil.copyfile(path, os.path.join(PROF, f"{tag}_launches.csv"))
    agg = collections.OrderedDict()
    for r in rows:
        k = r["Kernel Name"].split("(")[0].replace("icpb::<unnamed>::", "")
        agg.setdefault(k, []).append(float(r["Metric Value"].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    out = [f"# {tag}: ncu launch list ({len(rows)} launches; gpu__time_duration.sum, --clock-control none)", "",
           "Per-launch times are cold-cache and serialised: compare SHARES, not absolutes.", "",
           "| kernel | launches | mean us | min us | max us | share of listed time |", "|---|---:|---:|---:|---:|---:|"]
    for k, v in agg.items():
        out.append(f"| `{k[:70]}` | {len(v)} | {sum(v) / len(v) / 1e3:.2f} | {min(v) / 1e3:.2f} | {max(v) / 1e3:.2f} | "
                   f"{sum(v) / tot * 100:.1f} % |")
    open(os.path.join(PROF, f"{tag}_launches.md"), "w").write("\n".join(out) + "\n")
    print("\n".join(out))


def full(tag, rep, kernel, workload=None):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    picked = []
    for li, vals in enumerate(rows[2:]):
        d = dict(zip(hdr, vals))
        for h, u in zip(hdr, units):
            if any(h.startswith(k) for k in KEEP):
                picked.append((li, d.get("Kernel Name", "").split("(")[0], h, u, d[h]))
    with open(os.path.join(PROF, f"{tag}_{kernel}_metrics.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["launch", "kernel", "metric", "unit", "value"])
        w.writerows(picked)
    d0 = dict(zip(hdr, rows[2]))
    u0 = dict(zip(hdr, units))

    def to_bytes(name):
        v = float(d0[name].replace(",", ""))
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u0[name]]
    traffic = {"kernel": kernel, "source": f"profiles/{tag}_{kernel}_metrics.csv (ncu --set full, one launch)",
               "dram_bytes_read": to_bytes("dram__bytes_read.sum"), "dram_bytes_write": to_bytes("dram__bytes_write.sum")}
    traffic["dram_bytes_per_launch"] = traffic["dram_bytes_read"] + traffic["dram_bytes_write"]
    path = os.path.join(PROF, f"{kernel}_traffic.json")
    if workload:        # one capture per workload: bench.py prints roofline.traffic only for the workloads listed here
        old = json.load(open(path)) if os.path.exists(path) else {}
        wl = dict(old.get("workloads", {}))
        wl[workload] = traffic["dram_bytes_per_launch"]
        src = dict(old.get("sources", {}))
        src[workload] = traffic["source"]
        traffic = dict(old, workloads=wl, sources=src) if old else dict(traffic, workloads=wl, sources=src)
    json.dump(traffic, open(os.path.join(PROF, f"{kernel}_traffic.json"), "w"), indent=1)
    print(json.dumps(traffic))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("tag")
    ap.add_argument("--launches")
    ap.add_argument("--rep")
    ap.add_argument("--kernel", default="nn_filter")
    ap.add_argument("--workload", default=None, help="bench workload the capture was taken on (key of <kernel>_traffic.json's `workloads`)")
    a = ap.parse_args()
    os.makedirs(PROF, exist_ok=True)
    if a.launches:
        launches(a.tag, a.launches)
    if a.rep:
        full(a.tag, a.rep, a.kernel, a.workload)
