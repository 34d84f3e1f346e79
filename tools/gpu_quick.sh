#!/bin/bash
# On the GPU box: parity tests, per-kernel times of one bench run and executed-instruction counts of the kernels matching $1.
#   gpurun -- 'tools/gpu_quick.sh "dt_edge_kernel|dt_cell_kernel" [notest]'
PAT="$1"
mkdir -p gpurun_out
if [ "$2" != "notest" ]; then timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3; fi
timeout 300 python bench.py --steps 5 --warmup 3 --ktime --no-cpu-baseline --no-e2e > gpurun_out/quick.log 2>&1; echo "bench rc=$?"
grep -E "KTIME" gpurun_out/quick.log | grep -E "$PAT" | cut -c1-150
tail -1 gpurun_out/quick.log | cut -c1-110
timeout 300 ncu --metrics sm__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__throughput.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active \
  --clock-control none --kernel-name regex:"$PAT" --launch-skip 30 --launch-count ${3:-6} --csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph 2>/dev/null \
  | grep -E "^\"[0-9]" | python -c "
import csv,sys,collections
d=collections.OrderedDict()
for r in csv.reader(sys.stdin):
    d.setdefault((r[0],r[4].replace('<unnamed>::','').split('(')[0]),{})[r[12]]=r[14]
for (i,k),m in d.items():
    print('%-40s time=%8.1fus inst=%6.1fM issue=%5.1f%% l1=%5.1f%% occ=%5.1f%%'%(k[:40],float(m['gpu__time_duration.sum'])/1e3,float(m['sm__inst_executed.sum'])/1e6,float(m['smsp__issue_active.avg.pct_of_peak_sustained_active']),float(m['l1tex__throughput.avg.pct_of_peak_sustained_active']),float(m['sm__warps_active.avg.pct_of_peak_sustained_active'])))
"
