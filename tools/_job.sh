timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_gpu_properties.py tests/test_gpu_train.py tests/test_gpu_bn.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 8 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'], d['train_step']['ms_per_step'])
for k,v in d['kernel_breakdown'].items(): print('   ',k,v['launches'],round(v['ms']*1e3/v['launches'],1),'us')"
