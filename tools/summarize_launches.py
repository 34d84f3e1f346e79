"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel: launches, total, mean, share.
usage: python tools/summarize_launches.py gpurun_out/launches.csv > profiles/<round>_launches_summary.csv"""
import csv
import re
import sys
from collections import OrderedDict


def main(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    iname, ival, iunit = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = OrderedDict()
    for r in rows[1:]:
        if len(r) <= ival:
            continue
        name = re.sub(r"\(.*", "", r[iname]).replace("<unnamed>::", "").replace("mpasb200::", "")
        v = float(r[ival].replace(",", ""))
        scale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3}.get(r[iunit], 1.0)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v * scale
    tot = sum(a[1] for a in agg.values())
    w = csv.writer(sys.stdout)
    w.writerow(["kernel", "launches", "total_us", "mean_us", "share_pct"])
    for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        w.writerow([name, n, "%.1f" % t, "%.1f" % (t / n), "%.2f" % (100 * t / tot)])
    w.writerow(["TOTAL", sum(a[0] for a in agg.values()), "%.1f" % tot, "", "100.00"])


main(sys.argv[1])
