for v in base db2 mb128 both mb160; do for s in 8 16; do
  if [ $v = base ]; then L=ceracoder_b200/libceracuda.so; else L=build/var_$v/libceracuda.so; fi
  CERACUDA_LIB=$L python bench.py --no-cpu-baseline --streams $s --steps 6 > gpurun_out/v_${v}_$s.json 2>/dev/null
  python -c "
import json,sys;d=json.loads(open('gpurun_out/v_${v}_$s.json').read().strip().splitlines()[-1]);st=d['roofline']['stage_ms_per_frame'];print('$v',$s, round(d['value']),round(d['e2e']['value']), 'mb %.3f db %.3f fine %.3f pcode %.3f'%(st['mb_encode'],st['deblock'],st['me_fine'],st['p_code']))"
done; done
