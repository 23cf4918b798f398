#!/bin/bash
# Round-2 profile recipe (run under gpurun): plain run first, then the ncu launch list, then one --set full capture per hot kernel,
# then the per-launch instruction counts of one stream's P frame.  Short runs on a 24-frame store (--pool: ncu runs only).
set -x
TAG=${1:-r02}
B="python bench.py --steps 1 --warmup 1 --streams 2 --frames-per-step 6 --no-cpu-baseline --pool 24"
$B > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || exit 1
CC_DEBUG_CYCLES=1 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_cycles.json 2> gpurun_out/${TAG}_cycles.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/${TAG}_ncu1.log 2>&1
for kn in k_mb_encode k_deblock k_p_code k_me_l1 k_me_l0 k_cavlc_size k_cavlc_write k_hpel; do
  ncu --set full --clock-control none --import-source on -k regex:$kn -s 8 -c 2 -f -o gpurun_out/${TAG}_$kn $B > gpurun_out/${TAG}_ncu_$kn.log 2>&1
done
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -s 45 -c 15 --csv --log-file gpurun_out/${TAG}_insts.csv python bench.py --steps 1 --warmup 1 --streams 1 --frames-per-step 4 --no-cpu-baseline --pool 24 > gpurun_out/${TAG}_insts.log 2>&1
ls -la gpurun_out | tail -30
