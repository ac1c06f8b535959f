#!/usr/bin/env python3
"""Turn ncu outputs under gpurun_out/ into the small text summaries committed under profiles/.
usage: summarize.py launches <launches.csv> <proves in the run>   |   summarize.py raw <report.ncu-rep | raw-page.csv>"""
import collections, csv, re, subprocess, sys

def launches(path, proves):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        k = re.sub(r"\(.*", "", row["Kernel Name"])
        k = re.sub(r"g16::|void |Curve<|Field<|FqParams>>?|cub::|radix::policy_hub<.*|scan::policy_hub<.*", "", k)[:60]
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1e3 if row["Metric Unit"] == "ns" else v * 1e3 if row["Metric Unit"] == "ms" else v
        a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
    tot = sum(a[1] for a in agg.values())
    print("kernel | launches | total us | us per prove | share")
    print("---|---|---|---|---")
    for k, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%s | %d | %.0f | %.0f | %.1f%%" % (k, c, v, v / proves, 100 * v / tot))

def raw(path):
    if path.endswith(".csv"):   # already exported on the GPU box: ncu -i x.ncu-rep --page raw --csv
        out = open(path).read()
    else:
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    want = ["Kernel Name", "launch__grid_size", "launch__registers_per_thread", "gpu__time_duration.sum",
            "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__pcsamp_warps_issue_stalled_math_pipe_throttle", "smsp__pcsamp_warps_issue_stalled_wait",
            "smsp__pcsamp_warps_issue_stalled_no_instructions", "smsp__pcsamp_warps_issue_stalled_long_scoreboard",
            "smsp__pcsamp_warps_issue_stalled_barrier", "smsp__pcsamp_warps_issue_stalled_short_scoreboard",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
            "smsp__pcsamp_warps_issue_stalled_not_selected", "smsp__pcsamp_warps_issue_stalled_selected"]
    idx = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        print("---")
        for w in want:
            if w in idx:
                print("%s = %s %s" % (w, r[idx[w]][:100], units[idx[w]]))

if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], int(sys.argv[3]))
    else:
        raw(sys.argv[2])
