"""Turns ncu artefacts brought back in gpurun_out/ into the small tracked summaries under profiles/.

  python tools/ncu_summary.py launches <launches.csv> <out.md>     per-kernel totals of a gpu__time_duration launch list
  python tools/ncu_summary.py metrics <report.ncu-rep> <out.md>    key counters of every profiled launch in a --set full report
"""
import collections
import csv
import re
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__warps_eligible.avg.per_cycle_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
        "l1tex__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores"]


def short(name):
    name = re.sub(r"\(.*", "", name)
    name = re.sub(r"^void ", "", name)
    return re.sub(r"<.*", "", name).replace("mcpt::", "")


def launches(src, dst):
    rows = [r for r in csv.reader(open(src, errors="replace")) if len(r) > 14 and r[0].isdigit()]
    tot = collections.OrderedDict()
    for r in rows:
        k = short(r[4])
        n, t = tot.get(k, (0, 0.0))
        tot[k] = (n + 1, t + float(r[-1]))
    total = sum(t for _, t in tot.values())
    with open(dst, "w") as f:
        f.write("# ncu launch list summary (`--metrics gpu__time_duration.sum --clock-control none`)\n\n")
        f.write("Source: `%s` (%d launches, %.3f ms of kernel time; per-launch times are cold-cache and serialised — "
                "compare SHARES).\n\n| kernel | launches | total ms | share | avg us |\n|---|---:|---:|---:|---:|\n" %
                (src, len(rows), total * 1e-6))
        for k, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
            f.write("| `%s` | %d | %.3f | %.2f %% | %.1f |\n" % (k, n, t * 1e-6, 100 * t / total, t / n * 1e-3))


def metrics(src, dst):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    stall = [h for h in hdr if "smsp__average_warps_issue_stalled" in h and h.endswith("per_issue_active.ratio")]
    with open(dst, "w") as f:
        f.write("# ncu `--set full --clock-control none` key counters\n\nSource: `%s`\n" % src)
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            u = dict(zip(hdr, units))
            f.write("\n## launch %s: `%s`\n\n| metric | value | unit |\n|---|---:|---|\n" % (d.get("ID", "?"), short(d["Kernel Name"])))
            for k in KEYS:
                if k in d:
                    f.write("| %s | %s | %s |\n" % (k, d[k], u[k]))
            top = sorted(((float(d[h] or 0), h) for h in stall), reverse=True)[:6]
            f.write("\nTop warp stall reasons (warps per issue-active cycle): " + ", ".join(
                "%s %.2f" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v)
                for v, h in top) + "\n")


if __name__ == "__main__":
    {"launches": launches, "metrics": metrics}[sys.argv[1]](sys.argv[2], sys.argv[3])
