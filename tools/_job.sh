timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py -m gpu -x -q -k "decoder or encoder or fixture or baseline" 2>&1 | tail -2
timeout 200 python tools/bench_decoder.py 2>&1 | tail -4
python bench.py --steps 5 --warmup 3 --no-cpu --no-train 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'])
for k,v in d['kernel_breakdown'].items(): print('   ',k,v['launches'],round(v['ms']*1e3/v['launches'],1),'us')"
