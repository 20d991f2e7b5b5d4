timeout 600 python -m pytest tests/test_gpu_zz_precision.py tests/test_gpu_kernels.py -m gpu -x -q -k "precision or fused_update" 2>&1 | tail -3
CTN_PRECISION=fp16 python bench.py --steps 5 --warmup 3 --no-cpu --no-train 2>gpurun_out/x11.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('fp16', round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'], d['dtype'])
for k,v in d['kernel_breakdown'].items(): print('   ',k,v['launches'],round(v['ms']*1e3/v['launches'],1),'us')"
python bench.py --steps 5 --warmup 3 --no-cpu --no-train 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bf16x3', round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'])"
