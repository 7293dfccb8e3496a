#!/usr/bin/env python
"""profiles/r2_roofline.md: one row per kernel from `ncu --set full` reports — duration, the three candidate bounds ncu measures
(DRAM throughput, issue slots, L1 wavefront pipe, all in % of peak) and which is the largest, plus DRAM bytes and lanes per
instruction.  Several reports may name the same kernel (different scenes / a mid-render launch): every (report, kernel) pair
gets a row, averaged over that kernel's launches in the report.

    python tools/roofline_table.py profiles/r2_roofline.md gpurun_out/r2_all_cornell-box.ncu-rep[:label] ...
"""
import collections
import csv
import subprocess
import sys

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.abspath(__file__)))
from ncu_summary import short  # noqa: E402

M = {"dur": "gpu__time_duration.sum", "dram": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
     "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1": "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
     "lts": "lts__throughput.avg.pct_of_peak_sustained_elapsed", "rd": "dram__bytes_read.sum", "wr": "dram__bytes_write.sum",
     "lanes": "smsp__thread_inst_executed_per_inst_executed.ratio", "warps": "sm__warps_active.avg.pct_of_peak_sustained_active",
     "inst": "smsp__inst_executed.sum", "regs": "launch__registers_per_thread", "grid": "launch__grid_size", "block": "launch__block_size"}
UNIT = {"nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6,
        "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def rows_of(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        out = {"kernel": short(d["Kernel Name"])}
        for k, name in M.items():
            try:
                v = float(d[name].replace(",", ""))
            except (KeyError, ValueError):
                v = float("nan")
            if k in ("dur", "rd", "wr"):
                v *= UNIT.get(u.get(name, ""), 1.0)
            out[k] = v
        yield out


def main():
    dst, specs = sys.argv[1], sys.argv[2:]
    lines = []
    for spec in specs:
        path, _, label = spec.partition(":")
        label = label or path.split("/")[-1].replace(".ncu-rep", "")
        per = collections.OrderedDict()
        for r in rows_of(path):
            per.setdefault(r["kernel"], []).append(r)
        for k, rs in per.items():
            avg = lambda f: sum(r[f] for r in rs) / len(rs)
            cand = {"hbm": avg("dram"), "issue": avg("issue"), "l1 wavefronts": avg("l1"), "l2": avg("lts")}
            bound = max(cand, key=lambda b: cand[b] if cand[b] == cand[b] else -1)
            lines.append("| %s | `%s` | %d | %.1f | %s | %.1f %% | %.1f | %.1f | %.1f | %.1f | %.3g | %.1f | %.0f | %d |" % (
                label, k, len(rs), avg("dur"), bound, cand[bound], avg("dram"), avg("issue"), avg("l1"), avg("lts"), avg("rd") + avg("wr"),
                avg("lanes"), avg("warps"), int(avg("regs"))))
    with open(dst, "w") as f:
        f.write("# Roofline position of every kernel (`ncu --set full --clock-control none`, one B200)\n\n"
                "bound = the largest of: DRAM throughput (`gpu__dram_throughput`), issue slots (`smsp__issue_active`), the L1 wavefront pipe "
                "(`l1tex__data_pipe_lsu_wavefronts`), L2 throughput (`lts__throughput`) — each in % of its peak, which is the kernel's roofline "
                "fraction on that axis.  Durations are ncu's (cold caches, serialised); bytes are per launch.\n\n"
                "| capture | kernel | launches | avg us | bound | frac of that peak | DRAM % | issue % | L1 wavefront % | L2 % | DRAM bytes / launch | lanes / inst | warps active % | regs |\n"
                "|---|---|---:|---:|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|\n")
        f.write("\n".join(lines) + "\n")
    print("wrote", dst, len(lines), "rows")


if __name__ == "__main__":
    main()
