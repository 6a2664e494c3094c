#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv` per kernel: SASS lines with stall samples.

    ncu -i prof.ncu-rep --page source --csv > src.csv ; python tools/ncu_src.py src.csv [section] [min_samples]
"""
import csv
import sys


def sections(path):
    rows = list(csv.reader(open(path)))
    secs, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "rows": []}
            secs.append(cur)
        elif cur is not None and r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and r:
            cur["rows"].append(r)
    return secs


def main():
    secs = sections(sys.argv[1])
    if len(sys.argv) < 3:
        for i, s in enumerate(secs):
            print(i, s["name"][:100], len(s["rows"]))
        return
    s = secs[int(sys.argv[2])]
    mins = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    h = s["hdr"]
    ci = {n: i for i, n in enumerate(h)}
    stall = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
    tot = sum(int(r[ci["# Samples"]] or 0) for r in s["rows"])
    print(s["name"], "total samples", tot)
    for r in s["rows"]:
        n = int(r[ci["# Samples"]] or 0)
        if n < mins:
            continue
        st = sorted(((int(r[ci[k]] or 0), k[6:]) for k in stall), reverse=True)[:3]
        st = " ".join("%s=%d" % (k, v) for v, k in st if v)
        print("%6d %5.1f%% ex=%-8s %-70s %s" % (n, 100.0 * n / max(tot, 1), r[ci["Instructions Executed"]], r[ci["Source"]][:70], st))


if __name__ == "__main__":
    main()
