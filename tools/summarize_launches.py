"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel: launches, total, mean, share.
usage: python tools/summarize_launches.py gpurun_out/launches.csv > profiles/<round>_launches_summary.csv
A capture with several metrics per launch (the dram-traffic pass of tools/dram_traffic.py) is fine: only the
gpu__time_duration.sum rows are used.  A per-kernel table written by tools/dram_traffic.py (kernel, launches, ..., mean_us)
is accepted as well and gives the same summary."""
import csv
import re
import sys
from collections import OrderedDict


def main(path):
    first = open(path).readline()
    if first.startswith("kernel,launches"):  # per-kernel table of tools/dram_traffic.py
        agg = OrderedDict()
        for r in csv.DictReader(open(path)):
            agg[r["kernel"]] = [int(r["launches"]), int(r["launches"]) * float(r["mean_us"])]
        return emit(agg)
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    imetric = hdr.index("Metric Name")
    iname, ival, iunit = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = OrderedDict()
    for r in rows[1:]:
        if len(r) <= ival or r[imetric] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[iname]).replace("<unnamed>::", "").replace("mpasb200::", "")
        v = float(r[ival].replace(",", ""))
        scale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3}.get(r[iunit], 1.0)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v * scale
    emit(agg)


def emit(agg):
    tot = sum(a[1] for a in agg.values())
    w = csv.writer(sys.stdout)
    w.writerow(["kernel", "launches", "total_us", "mean_us", "share_pct"])
    for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        w.writerow([name, n, "%.1f" % t, "%.1f" % (t / n), "%.2f" % (100 * t / tot)])
    w.writerow(["TOTAL", sum(a[0] for a in agg.values()), "%.1f" % tot, "", "100.00"])


main(sys.argv[1])
