for kb in ${KBS:-0 64 120 180}; do for s in ${STREAMS_LIST:-8 16}; do
CC_WAVEFRONT_SMEM_KB=$kb python bench.py --no-cpu-baseline --streams $s --steps 6 > gpurun_out/sm_${kb}_$s.json 2>/dev/null
python - <<PY
import json
d=json.loads(open('gpurun_out/sm_${kb}_$s.json').read().strip().splitlines()[-1]);st=d['roofline']['stage_ms_per_frame']
print('smem %3d KB %2d  fps %5.0f e2e %5.0f  '%($kb,$s,d['value'],d['e2e']['value'])+' '.join('%s %.3f'%(n[:6],st[n]) for n in ('me_coarse','me_fine','p_code','mb_encode','pack','deblock','hpel')))
PY
done; done
