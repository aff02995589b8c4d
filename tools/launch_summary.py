#!/usr/bin/env python
"""Sum an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel: launches, total us, share.  The last `--steps`
worth of launches can be selected with --from-launch (the list of a bench_train run holds warm-up + timed steps).
    python tools/launch_summary.py gpurun_out/train_launches.csv [--from-launch N] [--grid]"""
import csv, sys, collections, re

def main():
    path = sys.argv[1]
    start = int(sys.argv[sys.argv.index("--from-launch") + 1]) if "--from-launch" in sys.argv else 0
    by_grid = "--grid" in sys.argv
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]; col = {h: i for i, h in enumerate(hdr)}
    agg = collections.OrderedDict()
    total = 0.0
    n = 0
    for r in rows[1:]:
        if r[col["Metric Name"]] != "gpu__time_duration.sum" or int(r[col["ID"]]) < start:
            continue
        name = re.sub(r"\(.*", "", r[col["Kernel Name"]])
        if by_grid:
            name += " grid=" + r[col["Grid Size"]]
        us = float(r[col["Metric Value"]].replace(",", "")) / 1e3
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1; a[1] += us; total += us; n += 1
    print(f"{n} launches, {total:.1f} us")
    for k, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{us:12.1f} us  {100 * us / total:5.1f}%  x{c:<5d} {k}")

main()
