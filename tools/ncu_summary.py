"""Summarise ncu captures into profiles/ (run here, no GPU needed).

    python tools/ncu_summary.py --rep gpurun_out/prof.ncu-rep --launches gpurun_out/launches.csv --name r01_humanoid_131072
"""
import argparse
import csv
import json
import os
import subprocess
import sys
from collections import Counter, defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RAW_KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__block_size", "launch__grid_size", "launch__waves_per_multiprocessor",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_blocks",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sector_hit_rate.pct",
]


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(out.splitlines()))


def raw_summary(rep):
    rows = ncu_csv(rep, "raw")
    hdr, units = rows[0], rows[1]
    kernels = []
    for r in rows[2:]:
        d = {"kernel": r[hdr.index("Kernel Name")][:120]}
        for k in RAW_KEYS:
            if k in hdr:
                i = hdr.index(k)
                d[k] = f"{r[i]} {units[i]}".strip()
        kernels.append(d)
    return kernels


def sass_summary(rep):
    rows = ncu_csv(rep, "source", ("--print-source", "sass"))
    hdr = rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    data = []
    for r in rows[2:]:
        if r and r[0] == "Kernel Name":
            break
        if len(r) >= len(hdr):
            data.append(r)
    f = lambda r, k: float(r[idx[k]] or 0)
    tot_i = sum(f(r, "Instructions Executed") for r in data)
    tot_s = sum(f(r, "# Samples") for r in data)
    reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    stalls = {h: sum(f(r, h) for r in data) for h in reasons}
    ops = Counter()
    for r in data:
        s = r[idx["Source"]].split()
        o = s[1] if s and s[0].startswith("@") else (s[0] if s else "")
        ops[o.split(".")[0]] += int(f(r, "Instructions Executed"))
    top = sorted(data, key=lambda r: -f(r, "# Samples"))[:12]
    return {
        "kernel": rows[0][1][:120], "sass_instructions": len(data), "warp_instructions_executed": tot_i,
        "samples": tot_s,
        "stall_pct": {k: round(100 * v / max(tot_s, 1), 1) for k, v in sorted(stalls.items(), key=lambda x: -x[1])[:8]},
        "opcode_pct": {k: round(100 * v / max(tot_i, 1), 1) for k, v in ops.most_common(14)},
        "bulk_copy_sass": sorted({k for k in ops if k.startswith(("UBLKCP", "UTMA", "SYNCS", "LDGSTS"))}),
        "hottest": [{"samples": int(f(r, "# Samples")), "executed": int(f(r, "Instructions Executed")),
                     "sass": r[idx["Source"]][:80]} for r in top],
    }


def launches_summary(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, mi, vi, ii = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
    gi = hdr.index("Grid Size") if "Grid Size" in hdr else None
    per = defaultdict(dict)
    for r in rows[1:]:
        try:
            # one row per (kernel, grid): a bench run launches the same kernel at several sizes (headline, extras, e2e chunks)
            name = r[ki][:100] + (" grid=" + r[gi].replace(" ", "") if gi is not None else "")
            per[(r[ii], name)][r[mi]] = float(r[vi].replace(",", ""))
        except ValueError:
            pass
    agg = defaultdict(lambda: defaultdict(list))
    for (_, k), m in per.items():
        for name, v in m.items():
            agg[k][name].append(v)
    out = {}
    for k, m in agg.items():
        out[k] = {name: {"launches": len(v), "mean": sum(v) / len(v), "min": min(v), "max": max(v)} for name, v in m.items()}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rep")
    ap.add_argument("--launches")
    ap.add_argument("--name", required=True)
    ap.add_argument("--note", default="")
    a = ap.parse_args()
    doc = {"name": a.name, "note": a.note}
    if a.launches:
        doc["launch_list"] = launches_summary(a.launches)
    if a.rep:
        doc["raw"] = raw_summary(a.rep)
        doc["sass"] = sass_summary(a.rep)
    path = os.path.join(ROOT, "profiles", a.name + ".json")
    with open(path, "w") as fh:
        json.dump(doc, fh, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
