#!/bin/bash
# bench lines per CC_OVERLAP setting (bit 0: filter next to k_mb_encode, bit 1: k_mb_encode next to k_p_code), release point of the
# filter (CC_TRAIL_PCT) and library variant (run under gpurun; the e2e leg is the one to read: the device-resident leg profiles its
# stages, which turns the overlaps off):
#   VARIANTS="base" OVERLAPS="0 1" PCTS="55" STREAMS_LIST="1 8 16" tools/overlap_sweep.sh
for v in ${VARIANTS:-base}; do for o in ${OVERLAPS:-0 1}; do for p in $([ $o = 0 ] && echo 0 || echo ${PCTS:-55}); do for s in ${STREAMS_LIST:-8 16}; do
  if [ $v = base ]; then L=ceracoder_b200/libceracuda.so; else L=build/var_$v/libceracuda.so; fi
  CC_OVERLAP=$o CC_TRAIL_PCT=$p CERACUDA_LIB=$L timeout 300 python bench.py --no-cpu-baseline --streams $s --steps ${STEPS:-6} > gpurun_out/o_${v}_${o}_${p}_$s.json 2>/dev/null
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/o_${v}_${o}_${p}_$s.json').read().strip().splitlines()[-1])
    print('%-6s overlap %d at %3d%%  %2d streams  fps %5.0f  e2e %5.0f'%('$v',$o,$p,$s,d['value'],d['e2e']['value']))
except Exception as ex: print('$v $o $p $s failed', ex)
PY
done; done; done; done
