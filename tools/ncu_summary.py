#!/usr/bin/env python
"""Summarise Nsight Compute reports of bench.py for profiles/.

    python tools/ncu_summary.py --tag r01_final --workload inst60000 gpurun_out/x.ncu-rep [...]

writes profiles/<tag>_ncu_<workload>.md (one row per captured launch group: duration, occupancy,
unit utilisation, DRAM/L2 bytes, top stall reasons) and merges the measured DRAM traffic per launch
(dram__bytes_read.sum + dram__bytes_write.sum) of every loop kernel into profiles/ncu_traffic.json,
which bench.py reports as roofline.traffic.  Needs `ncu` on PATH (reads the report, no GPU).
"""
import argparse
import csv
import io
import json
import os
import re
import subprocess
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

METRICS = [
    "gpu__time_duration.sum",
    "launch__grid_size",
    "launch__block_size",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "lts__t_sectors_srcunit_tex_op_read.sum",
    "lts__t_sectors_srcunit_tex_op_write.sum",
    "lts__t_sector_hit_rate.pct",
]

STALLS = "smsp__average_warps_issue_stalled_{}_per_issue_active.ratio"
STALL_NAMES = ["long_scoreboard", "short_scoreboard", "wait", "barrier", "math_pipe_throttle", "mio_throttle",
               "lg_throttle", "branch_resolving", "no_instruction", "not_selected", "dispatch_stall", "membar",
               "sleeping", "drain", "imc_miss", "tex_throttle"]


def label(kernel):
    """bench.py's name of the launch (recommender_system_b200/api.py KERNELS)."""
    if "k_update_L" in kernel or "k_dot_units" in kernel:      # slot 0: the by-user pass (split mode: its dots)
        return "k_update_L"
    if "k_acc" in kernel:                                        # split mode: every accumulation that is not a deep chain
        return "k_chain_bulk"
    if "k_dot" in kernel:
        return "k_dot"
    if "k_chain_ws" in kernel:
        return "k_chain_deep"
    m = re.search(r"k_chain<\s*\d+\s*,\s*(\d+)\s*>", kernel)
    if m:
        return "k_chain_deep" if int(m.group(1)) in (1, 3) else "k_chain_bulk"
    if "k_apply_R" in kernel:
        return "exchange_apply"
    return None


def to_float(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def read_report(path):
    names = METRICS + [STALLS.format(n) for n in STALL_NAMES]
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv", "--print-units", "base", "--metrics",
                          ",".join(names)], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}
    launches = []
    for r in rows[2:]:
        if len(r) != len(hdr):
            continue
        d = {"kernel": r[col["Kernel Name"]]}
        for n in names:
            if n in col:
                d[n] = to_float(r[col[n]])
        launches.append(d)
    return launches


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tag", required=True)
    ap.add_argument("--workload", required=True)
    ap.add_argument("reports", nargs="+")
    a = ap.parse_args()
    groups = defaultdict(list)
    for rep in a.reports:
        for l in read_report(rep):
            groups[re.sub(r"\(.*", "", l["kernel"]).strip()].append(l)

    def avg(ls, k):
        v = [l[k] for l in ls if l.get(k) is not None]
        return sum(v) / len(v) if v else float("nan")

    lines = [f"# ncu --set full, workload {a.workload} ({a.tag})", "",
             "Captured with `--clock-control none`; durations are cold-cache, serialised launches (the bench",
             "numbers come from CUDA events outside the profiler).  Averages over the captured launches.", "",
             "| kernel | launches | us | grid x block | regs | dyn smem | warps active % | issue active % | FP64 pipe % "
             "| L1TEX % | smem wavefronts % | L2 % | DRAM % | DRAM rd+wr MB | L2<-SM rd MB | L2 hit % |",
             "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
    traffic = {}
    stall_lines = []
    for k, ls in sorted(groups.items()):
        dram = avg(ls, "dram__bytes_read.sum") + avg(ls, "dram__bytes_write.sum")
        l2rd = avg(ls, "lts__t_sectors_srcunit_tex_op_read.sum") * 32
        lines.append("| `%s` | %d | %.1f | %d x %d | %d | %d | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f |" % (
            k, len(ls), avg(ls, "gpu__time_duration.sum") / 1e3, avg(ls, "launch__grid_size"),
            avg(ls, "launch__block_size"), avg(ls, "launch__registers_per_thread"),
            avg(ls, "launch__shared_mem_per_block_dynamic"),
            avg(ls, "sm__warps_active.avg.pct_of_peak_sustained_active"),
            avg(ls, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
            avg(ls, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
            avg(ls, "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
            avg(ls, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"),
            avg(ls, "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
            avg(ls, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
            dram / 1e6, l2rd / 1e6, avg(ls, "lts__t_sector_hit_rate.pct")))
        st = sorted(((avg(ls, STALLS.format(n)), n) for n in STALL_NAMES if avg(ls, STALLS.format(n)) == avg(ls, STALLS.format(n))),
                    reverse=True)[:4]
        stall_lines.append("- `%s`: " % k + ", ".join("%s %.2f" % (n, v) for v, n in st))
        lab = label(k)
        if lab and dram == dram:
            traffic[lab] = traffic.get(lab, 0) + int(dram)
    lines += ["", "Warp stall reasons (warps stalled per issue-active cycle, top 4):", ""] + stall_lines + [""]
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", f"{a.tag}_ncu_{a.workload}.md"), "w") as fh:
        fh.write("\n".join(lines))
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    allt = {}
    if os.path.exists(tp):
        with open(tp) as fh:
            allt = json.load(fh)
    allt[a.workload] = traffic
    allt.setdefault("_note", "dram__bytes_read.sum + dram__bytes_write.sum per launch, from ncu --set full "
                             "(tools/ncu_summary.py); cold-cache launches")
    with open(tp, "w") as fh:
        json.dump(allt, fh, indent=1, sort_keys=True)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
