#!/usr/bin/env python
"""Turn gpurun_out/launches.csv (ncu --metrics gpu__time_duration.sum) and an ncu --set full report into the small
text summaries committed under profiles/.  Usage: summarize_ncu.py <tag> [launches.csv] [report.ncu-rep]"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]


def launches(path, out):
    txt = open(path).read()
    r = csv.DictReader(io.StringIO(txt[txt.index('"ID"'):]))
    agg = collections.OrderedDict()
    for row in r:
        if row["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v = {"ns": v / 1e3, "us": v, "ms": v * 1e3}.get(row["Metric Unit"], v)
        agg.setdefault(row["Kernel Name"].split("(")[0][-80:], []).append(v)
    tot = sum(sum(v) for v in agg.values())
    out.write("kernel | launches | mean us | total ms | share\n---|---|---|---|---\n")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        out.write(f"{k} | {len(v)} | {sum(v) / len(v):.1f} | {sum(v) / 1e3:.2f} | {100 * sum(v) / tot:.1f}%\n")


def full(path, out):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    import json
    for r in rows[2:]:
        if "search_p1_kernel" in r[hdr.index("Kernel Name")]:
            def gb(k):
                v, u = float(r[hdr.index(k)].replace(",", "")), units[hdr.index(k)]
                return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            with open("profiles/search_traffic.json", "w") as tf:
                json.dump({"kernel": "search_p1_kernel", "source": path,
                           "dram_bytes_per_launch": gb("dram__bytes_read.sum") + gb("dram__bytes_write.sum")}, tf)
    for r in rows[2:]:
        out.write(f"\n### {r[hdr.index('Kernel Name')][:120]}\n")
        for k in KEYS:
            if k in hdr:
                out.write(f"- {k}: {r[hdr.index(k)]} {units[hdr.index(k)]}\n")


if __name__ == "__main__":
    tag = sys.argv[1]
    lp = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/launches.csv"
    fp = sys.argv[3] if len(sys.argv) > 3 else "gpurun_out/prof_search.ncu-rep"
    with open(f"profiles/{tag}_launches.md", "w") as f:
        f.write(f"# {tag}: ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`) of "
                "`python bench.py --steps 2 --warmup 3 --no-cpu-baseline`\n\nCold-cache, serialised per-launch times: "
                "compare shares, not absolutes.\n\n")
        launches(lp, f)
    with open(f"profiles/{tag}_search_full.md", "w") as f:
        f.write(f"# {tag}: `ncu --set full --clock-control none` of the nearest-code search kernels\n")
        full(fp, f)
    print("wrote profiles/%s_*.md" % tag)
