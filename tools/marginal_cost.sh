#!/bin/bash
# Marginal cost of each stage with the GPU shared by S streams (run under gpurun): the stage's kernel is launched twice
# (CC_DBG_DUP, idempotent kernels; same bytes) or left out (CC_DBG_SKIP: the in-loop filter, whose output then differs).
# Δ(1/fps) against the plain run = what the stage costs per frame where it matters: in the mix.
# CC_OVERLAP=0: the launch order with the filter after the macroblock kernel (next to it, a second k_mb_encode cannot be launched
# between the two).
for S in ${STREAMS_LIST:-8 16}; do
for v in base dup:2 dup:3 dup:4 dup:5 dup:9 skip:7; do
  k=${v%%:*}; b=${v##*:}
  D=0; K=0
  [ $k = dup ] && D=$((1<<b)); [ $k = skip ] && K=$((1<<b))
  CC_OVERLAP=0 CC_DBG_DUP=$D CC_DBG_SKIP=$K python bench.py --no-cpu-baseline --streams $S --steps 6 > gpurun_out/mc_${S}_$k$b.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open('gpurun_out/mc_${S}_$k$b.json').read().strip().splitlines()[-1]);st=d['roofline']['stage_ms_per_frame']
print('$S $v', 'fps %.0f us/frame %.1f'%(d['value'],1e6/d['value']), ' '.join('%s %.3f'%(n[:6],st[n]) for n in ('me_coarse','me_fine','p_code','mb_encode','pack','deblock','hpel')))
PY
done; done
