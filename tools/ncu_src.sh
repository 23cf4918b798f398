#!/bin/bash
# source-level captures of the given kernels (run under gpurun): tools/ncu_src.sh TAG k_me_l0 k_p_code ...
TAG=$1; shift
B="python bench.py --steps 1 --warmup 1 --streams 2 --frames-per-step 6 --no-cpu-baseline --pool 24"
for kn in "$@"; do
  ncu --set full --clock-control none --import-source on -k regex:$kn -s 8 -c 2 -f -o gpurun_out/${TAG}_$kn $B > gpurun_out/${TAG}_ncu_$kn.log 2>&1
done
ls -la gpurun_out/*.ncu-rep
