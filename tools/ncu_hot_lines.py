#!/usr/bin/env python
"""CUDA source lines of one profiled kernel ranked by warp instructions executed (needs -lineinfo and --import-source on).

  python tools/ncu_hot_lines.py report.ncu-rep KERNEL_REGEX [LAUNCH_INDEX] [--per N] [--top K]

Runs `ncu -i report --page source --csv --print-source cuda,sass --kernel-id ::regex:KERNEL_REGEX:LAUNCH_INDEX` and sums the
per-line rows.  --per N divides the counts by N (the number of mesh elements = warps of the launch), which gives warp
instructions per element and line -- the figure the instruction-bound kernels are tuned against."""
import csv
import subprocess
import sys


def main():
    a = sys.argv[1:]
    per, top = 1.0, 40
    if "--per" in a:
        i = a.index("--per"); per = float(a[i + 1]); del a[i:i + 2]
    if "--top" in a:
        i = a.index("--top"); top = int(a[i + 1]); del a[i:i + 2]
    rep, rx = a[0], a[1]
    idx = a[2] if len(a) > 2 else "1"
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-id", "::regex:%s:%s" % (rx, idx)],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    fname, hdr, lines, seen = None, None, [], set()
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            fname = r[1].split("/")[-1]
        elif r and r[0] == "Function Name":
            print(r[1])
        elif r and r[0] == "Line No":
            hdr = r
        elif hdr and r and r[0].isdigit() and len(r) > 10:
            iex, ismp, ithr = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
            try:
                key = (fname, int(r[0]))
                if key in seen:
                    continue
                seen.add(key)
                lines.append((int(r[iex]), int(r[ismp]), int(r[ithr]), fname, int(r[0]), r[1].strip()))
            except ValueError:
                pass
    tot = sum(l[0] for l in lines) or 1
    smp = sum(l[1] for l in lines) or 1
    thr = sum(l[2] for l in lines)
    print("warp instructions %.0f (%.1f per element), active threads per instruction %.1f, samples %d" % (tot, tot / per, thr / tot, smp))
    print("%9s %6s %6s  %s" % ("instr/el", "inst%", "smpl%", "line"))
    for ex, sm, th, f, ln, src in sorted(lines, reverse=True)[:top]:
        print("%9.1f %6.1f %6.1f  %s:%d  %s" % (ex / per, 100.0 * ex / tot, 100.0 * sm / smp, f, ln, src[:110]))


main()
