"""Hot SASS instructions from `ncu -i X.ncu-rep --page source --csv --kernel-name regex:K` output (first kernel in the file).
usage: python tools/ncu_sass_hot.py src.csv [top_n] [--all]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 30
hdr = rows[1]
i_s, i_i, i_src = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
stalls = [(hdr.index(n), n.replace("stall_", "")) for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
data = []
for n, r in enumerate(rows[2:]):
    if len(r) <= i_s or r[0] == "Address" or r[0] == "Kernel Name":
        if r and r[0] == "Kernel Name" and data:
            break
        continue
    st = sorted(((int(r[i]), s) for i, s in stalls if r[i].isdigit()), reverse=True)[:2]
    data.append((n, int(r[i_s]), int(r[i_i]), r[i_src].strip(), st))
tot, toti = sum(d[1] for d in data), sum(d[2] for d in data)
print("instructions:", len(data), "samples:", tot, "warp-instructions executed:", toti)
if "--all" in sys.argv:
    for d in data:
        print("%4d %5.2f%% %5.2f%%  %-60s %s" % (d[0], 100 * d[1] / tot, 100 * d[2] / toti, d[3][:60], d[4][0] if d[1] else ""))
else:
    for d in sorted(data, key=lambda d: -d[1])[:topn]:
        print("%4d %5.2f%% samp %5.2f%% inst  %-70s %s" % (d[0], 100 * d[1] / tot, 100 * d[2] / toti, d[3][:70], d[4]))
