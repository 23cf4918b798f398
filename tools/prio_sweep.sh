for pr in ${PRS:-0 2 1 3}; do for s in ${STREAMS_LIST:-8 16}; do
CC_PRIO=$pr python bench.py --no-cpu-baseline --streams $s --steps 6 > gpurun_out/pr_${pr}_$s.json 2>/dev/null
python - <<PY
import json
d=json.loads(open('gpurun_out/pr_${pr}_$s.json').read().strip().splitlines()[-1]);st=d['roofline']['stage_ms_per_frame']
print('prio %d %2d  fps %5.0f e2e %5.0f  '%($pr,$s,d['value'],d['e2e']['value'])+' '.join('%s %.3f'%(n[:6],st[n]) for n in ('me_coarse','wait_ref','me_fine','p_code','mb_encode','pack','deblock','hpel')))
PY
done; done
