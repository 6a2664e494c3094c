#!/usr/bin/env python
"""Turn ncu outputs into the small text summaries committed under profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches.csv   > profiles/rNN_launches.txt
    python tools/ncu_summary.py full gpurun_out/prof.ncu-rep       > profiles/rNN_top_kernels.txt
"""
import collections
import csv
import io
import subprocess
import sys


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1000 if u == "ns" else v * 1000 if u == "ms" else v
        k = row["Kernel Name"]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    print("# ncu --metrics gpu__time_duration.sum --clock-control none : per-kernel totals (cold-cache, serialised)")
    print("# total %.1f us over %d launches" % (tot, sum(v[0] for v in agg.values())))
    print("%10s %6s %6s %10s  kernel" % ("us", "share", "n", "avg_us"))
    for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%10.1f %5.1f%% %6d %10.2f  %s" % (t, 100 * t / tot, c, t / c, k[:110]))


WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__grid_size", "launch__block_size", "launch__cluster_dim_x", "launch__registers_per_thread",
        "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_bytes.sum"]


def full(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, units = rows[0], rows[1]
    idx = [(w, h.index(w)) for w in WANT if w in h]
    kn = h.index("Kernel Name")
    print("# ncu --set full --clock-control none : one line per captured launch")
    for r in rows[2:]:
        print(r[kn][:120])
        for w, i in idx:
            print("    %-62s %s %s" % (w, r[i], units[i]))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
