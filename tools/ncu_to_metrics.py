#!/usr/bin/env python
"""Turns the ncu CSV logs of tools/profile_render.py runs into profiles/kernel_metrics.json (what bench.py's roofline reads)
and a per-kernel table (markdown) for profiles/.

    python tools/ncu_to_metrics.py --tag "r2 <date> <commit>" gpurun_out/prof_<workload>.csv:gpurun_out/prof_<workload>.json ...

For every kernel name: launches, summed duration, summed warp / thread instructions, summed DRAM bytes.  Per workload it
stores, for wf_trace / pt_megakernel: warp instructions per RAY (sum over all its launches / device-counted rays), active
lanes per instruction, DRAM bytes per launch; for wf_shade: warp instructions per SAMPLE.
"""
import argparse
import csv
import hashlib
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import KERNEL_SOURCES, kernel_source_hash  # noqa: E402


def short(name):
    m = re.match(r"(?:void )?(?:mcpt::)?([A-Za-z_0-9:]+)", name)
    n = m.group(1) if m else name
    return n.split("::")[-1] if "cub" not in name else "cub::" + n.split("::")[-1]


def read_csv(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.DictReader(lines)
    for r in rd:
        rows.append(r)
    return rows


def aggregate(rows):
    """-> {kernel: {"launches", metric sums...}}; the CSV has one row per (launch id, metric)."""
    per = {}
    seen = {}
    for r in rows:
        k = short(r["Kernel Name"])
        d = per.setdefault(k, {"launches": 0})
        key = (r["ID"], k)
        if key not in seen:
            seen[key] = 1
            d["launches"] += 1
        try:
            v = float(r["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        unit = r.get("Metric Unit", "")
        name = r["Metric Name"]
        if name == "gpu__time_duration.sum":
            v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}.get(unit, 1e-6)
        if name.startswith("dram__bytes"):
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "B": 1, "KB": 1e3, "MB": 1e6, "GB": 1e9}.get(unit, 1)
        d[name] = d.get(name, 0.0) + v
    return per


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("pairs", nargs="+", help="ncu.csv:run.json")
    ap.add_argument("--tag", default="")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "kernel_metrics.json"))
    ap.add_argument("--table", default=None)
    a = ap.parse_args()
    out = {"source_hash": kernel_source_hash(), "sources": KERNEL_SOURCES, "captured": a.tag, "workloads": {}}
    table = []
    for pair in a.pairs:
        c, j = pair.split(":")
        run = json.loads(open(j).read().strip().splitlines()[-1])
        per = aggregate(read_csv(c))
        w = out["workloads"].setdefault(run["workload"], {"spp": run["spp"], "rays": run["rays"], "samples": run["samples"]})
        for k, d in sorted(per.items(), key=lambda kv: -kv[1].get("gpu__time_duration.sum", 0)):
            wi, ti = d.get("smsp__inst_executed.sum", 0.0), d.get("smsp__thread_inst_executed.sum", 0.0)
            dram = d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
            rec = {"launches": d["launches"], "ms_total": d.get("gpu__time_duration.sum", 0.0), "warp_inst": wi, "thread_inst": ti,
                   "lanes_per_inst": ti / wi if wi else None, "dram_bytes_per_launch": dram / d["launches"]}
            if k in ("wf_trace", "pt_megakernel", "trace_kernel"):
                rec["warp_inst_per_ray"] = wi / run["rays"]
            if k == "wf_shade":
                rec["warp_inst_per_sample"] = wi / run["samples"]
            w[k] = rec
            table.append((run["workload"], k, rec))
    json.dump(out, open(a.out, "w"), indent=1)
    if a.table:
        with open(a.table, "w") as f:
            f.write("| workload | kernel | launches | total ms (ncu, serialised) | warp inst | lanes / inst | DRAM bytes / launch |\n|---|---|---:|---:|---:|---:|---:|\n")
            for wl, k, r in table:
                f.write("| %s | %s | %d | %.3f | %.4g | %s | %.4g |\n" % (wl, k, r["launches"], r["ms_total"], r["warp_inst"],
                                                                          "%.1f" % r["lanes_per_inst"] if r["lanes_per_inst"] else "-", r["dram_bytes_per_launch"]))
    print("wrote", a.out)


if __name__ == "__main__":
    main()
