#!/usr/bin/env python
"""tools/ncu_facts.py REPORT.ncu-rep CONFIG [SOURCE-LABEL]  ->  merges into profiles/ncu_facts.json

Per-kernel facts that bench.py cannot measure itself (it may not run under a profiler): DRAM bytes per launch
(dram__bytes_read.sum + dram__bytes_write.sum), FP64-pipe activity, issue-slot utilisation, registers -- read from one
`ncu --set full --clock-control none` capture of `python bench.py --config CONFIG --steps 1 --warmup 1 --no-e2e --no-cpu`.
Launches of one kernel are averaged (fit_kernel runs twice per step, the polyfit kernels once per round); the kernel
name is the function name without template arguments, as bench.py's stage table names them."""
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = {
    "time_ns": "gpu__time_duration.sum",
    "dram_rd": "dram__bytes_read.sum",
    "dram_wr": "dram__bytes_write.sum",
    "fp64": "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "issue": "sm__inst_issued.avg.pct_of_peak_sustained_active",
    "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "regs": "launch__registers_per_thread",
    "warps": "sm__warps_active.avg.pct_of_peak_sustained_active",
}
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
              "ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9, "usecond": 1e3, "msecond": 1e6, "nsecond": 1.0, "second": 1e9}


def num(x):
    try:
        return float(x.replace(",", ""))
    except Exception:
        return None


def main():
    rep, config = sys.argv[1], sys.argv[2]
    label = sys.argv[3] if len(sys.argv) > 3 else os.path.basename(rep)
    if rep.endswith(".csv"):      # already exported on the GPU box (`ncu -i X.ncu-rep --page raw --csv`): the report
        txt = open(rep).read()    # itself is too large to travel back
    else:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    head, units, body = rows[0], rows[1], rows[2:]
    col = {k: (head.index(v) if v in head else None) for k, v in COLS.items()}
    name_i = head.index("Kernel Name")
    per = {}
    for r in body:
        if len(r) <= name_i:
            continue
        name = re.sub(r"^(void\s+)?(shopt::)?", "", r[name_i])
        name = re.split(r"[<(]", name)[0]
        row = {}
        for k, i in col.items():
            if i is None:
                continue
            val = num(r[i])
            if val is not None:
                row[k] = val * UNIT_SCALE.get(units[i], 1.0)
        per.setdefault(name, []).append(row)
    facts = {}
    for name, launches in per.items():
        # the polyfit kernels run once per round on shrinking (or, for the downdate, much smaller) lists: the facts are
        # those of the LONGEST launch in the capture (round 1), which is the launch bench.py's stage line describes
        e = max(launches, key=lambda row: row.get("time_ns", 0.0))
        rd, wr = e.get("dram_rd"), e.get("dram_wr")
        facts[name] = {
            "launches_in_capture": len(launches),
            "ms_under_ncu": e.get("time_ns", 0.0) / 1e6,
            "dram_bytes_per_launch": (rd + wr) if rd is not None and wr is not None else None,
            "dram_read_bytes_per_launch": rd, "dram_write_bytes_per_launch": wr,
            "fp64_pipe_pct": e.get("fp64"), "issue_slot_pct": e.get("issue"), "dram_throughput_pct": e.get("dram_pct"),
            "registers": e.get("regs"), "warps_active_pct": e.get("warps"),
            "source": "ncu --set full --clock-control none, %s (longest launch of the kernel in the capture)" % label,
        }
    path = os.environ.get("NCU_FACTS_OUT") or os.path.join(ROOT, "profiles", "ncu_facts.json")   # the tests write elsewhere
    try:
        with open(path) as f:
            allf = json.load(f)
    except Exception:
        allf = {}
    allf[config] = facts
    with open(path, "w") as f:
        json.dump(allf, f, indent=1, sort_keys=True)
    for k, v in sorted(facts.items(), key=lambda kv: -kv[1]["ms_under_ncu"] * kv[1]["launches_in_capture"]):
        print("%-28s x%-3d %9.3f ms  dram %s  fp64 %s  issue %s  regs %s" % (
            k, v["launches_in_capture"], v["ms_under_ncu"],
            "%.3e" % v["dram_bytes_per_launch"] if v["dram_bytes_per_launch"] is not None else "-",
            "%.1f" % v["fp64_pipe_pct"] if v["fp64_pipe_pct"] is not None else "-",
            "%.1f" % v["issue_slot_pct"] if v["issue_slot_pct"] is not None else "-", v["registers"]))


if __name__ == "__main__":
    main()
