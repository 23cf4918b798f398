#!/bin/bash
# per-kernel instruction counts / durations of one P frame (run under gpurun): 1 stream, frame 3
TAG=${1:-x}
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -s 30 -c 12 --csv --log-file gpurun_out/${TAG}_insts.csv python bench.py --steps 1 --warmup 1 --streams 1 --frames-per-step 4 --no-cpu-baseline > gpurun_out/${TAG}_insts.log 2>&1
python - <<PY
import csv,collections
rows=list(csv.reader(l for l in open('gpurun_out/${TAG}_insts.csv') if l.startswith('"')))
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ii=h.index("ID")
d=collections.OrderedDict()
for r in rows[1:]:
    d.setdefault((r[ii],r[ki].split('(')[0]),{})[r[mi]]=r[vi]
for (i,k),m in d.items():
    print(f"{i:>4} {k:14s} dur_us={float(m['gpu__time_duration.sum'])/1000:8.1f} inst={float(m['smsp__inst_executed.sum'])/1e6:7.2f}M sm%={m['sm__throughput.avg.pct_of_peak_sustained_elapsed'][:5]} issue%={m['smsp__issue_active.avg.pct_of_peak_sustained_active'][:5]}")
PY
