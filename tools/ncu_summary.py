"""Turns ncu artefacts brought back in gpurun_out/ into the small tracked summaries under profiles/.

  python tools/ncu_summary.py launches <launches.csv> <out.md>     per-kernel totals of a gpu__time_duration launch list
  python tools/ncu_summary.py metrics <report.ncu-rep> <out.md>    key counters of every profiled launch in a --set full report
  python tools/ncu_summary.py source <report.ncu-rep> <out.md>     per-instruction execution counts of the first launch (source
                                                                   view): share and active lanes per 20-instruction segment, opcode mix
  python tools/ncu_summary.py lines <report.ncu-rep> <out.md>      the same per CUDA source line (top lines by instructions + stall samples)
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


def source(src, dst, seg=20):
    raw = subprocess.run(["ncu", "-i", src, "--page", "source", "--csv", "--print-source", "sass"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    name = rows[0][1] if rows and len(rows[0]) > 1 else "?"
    hdr = rows[1]
    isrc, ie, it, iss = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    data = []
    for r in rows[2:]:
        if len(r) < 10 or r[0] == "Kernel Name":
            break
        if r[0] != "Address":
            data.append((r[isrc].strip(), int(r[ie]), int(r[it]), int(r[iss])))
    tot = sum(d[1] for d in data)
    ops = collections.Counter()
    for s_, e, _, _ in data:
        tok = s_.split()
        op = tok[1] if tok and tok[0].startswith("@") and len(tok) > 1 else (tok[0] if tok else "?")
        ops[op.split(".")[0]] += e
    with open(dst, "w") as f:
        f.write("# ncu source view: executed warp instructions per SASS segment\n\nSource: `%s`, first launch of `%s`: %d SASS instructions, "
                "%d warp instructions executed, %.1f active lanes per instruction.\n\n" %
                (src, short(name), len(data), tot, sum(d[2] for d in data) / max(tot, 1)))
        f.write("Opcode mix (share of executed warp instructions): " + ", ".join("%s %.1f %%" % (k, 100 * v / tot) for k, v in ops.most_common(14)) + "\n\n")
        f.write("| SASS instructions | share | active lanes | stall samples | first instruction |\n|---|---:|---:|---:|---|\n")
        for k in range(0, len(data), seg):
            g = data[k:k + seg]
            e = sum(d[1] for d in g)
            if e * 1000 < tot:
                continue
            f.write("| %d-%d | %.1f %% | %.1f | %d | `%s` |\n" % (k, k + len(g) - 1, 100 * e / tot, sum(d[2] for d in g) / max(e, 1), sum(d[3] for d in g), g[0][0][:48]))


def lines(src, dst, top=45):
    """Per CUDA source line (needs -lineinfo and --import-source on): executed warp instructions, active lanes, stall samples."""
    raw = subprocess.run(["ncu", "-i", src, "--page", "source", "--csv", "--print-source", "cuda,sass"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    per, fname, hdr, kernel = [], "?", None, "?"
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
        elif r[0] == "Function Name":
            kernel = r[1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and r[0].isdigit() and len(r) > 8 and r[2] == "-":
            d = dict(zip(hdr[4:], r[4:]))
            try:
                per.append((fname, int(r[0]), r[1].strip(), int(d["Instructions Executed"] or 0), int(d["Thread Instructions Executed"] or 0), int(d["# Samples"] or 0)))
            except (KeyError, ValueError):
                pass
    tot = sum(p[3] for p in per) or 1
    smp = sum(p[5] for p in per) or 1
    byfile = collections.Counter()
    for p in per:
        byfile[p[0]] += p[3]
    with open(dst, "w") as f:
        f.write("# ncu source view by CUDA line\n\nSource: `%s`, `%s`: %d warp instructions executed, %.1f active lanes per instruction, %d stall samples.\n\n" %
                (src, short(kernel), tot, sum(p[4] for p in per) / tot, smp))
        f.write("Share of executed warp instructions by file: " + ", ".join("%s %.1f %%" % (k, 100 * v / tot) for k, v in byfile.most_common()) + "\n\n")
        f.write("| file:line | inst share | lanes | sample share | source |\n|---|---:|---:|---:|---|\n")
        for p in sorted(per, key=lambda p: -(p[3] / tot + p[5] / smp))[:top]:
            f.write("| %s:%d | %.1f %% | %.1f | %.1f %% | `%s` |\n" % (p[0], p[1], 100 * p[3] / tot, p[4] / max(p[3], 1), 100 * p[5] / smp, p[2][:90].replace("|", "¦")))


if __name__ == "__main__":
    {"launches": launches, "metrics": metrics, "source": source, "lines": lines}[sys.argv[1]](sys.argv[2], sys.argv[3])
