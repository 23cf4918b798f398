#!/bin/bash
# bench lines of library variants built by tools/build_variant.sh (run under gpurun): VARIANTS="base a b" STREAMS_LIST="8 16" tools/variant_bench.sh
for v in ${VARIANTS:-base}; do for s in ${STREAMS_LIST:-8 16}; do
  if [ $v = base ]; then L=ceracoder_b200/libceracuda.so; else L=build/var_$v/libceracuda.so; fi
  CERACUDA_LIB=$L python bench.py --no-cpu-baseline --streams $s --steps ${STEPS:-6} > gpurun_out/v_${v}_$s.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open('gpurun_out/v_${v}_$s.json').read().strip().splitlines()[-1]);st=d['roofline']['stage_ms_per_frame']
print('%-8s %2d  fps %5.0f e2e %5.0f  '%('$v',$s,d['value'],d['e2e']['value'])+' '.join('%s %.3f'%(n[:6],st[n]) for n in ('me_coarse','me_fine','p_code','mb_encode','pack','deblock','hpel')))
PY
done; done
