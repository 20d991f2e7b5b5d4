python -m pytest tests/test_gpu_backward_kernels.py tests/test_gpu_train.py tests/test_gpu_kernels.py -q -m gpu -x 2>&1 | tail -3
for i in 1 2; do python tools/profile_step_phases.py 3 | egrep "total|fwd|bwd"; done
python bench.py --steps 4 --warmup 3 2>&1 | tail -1 > gpurun_out/x22.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/x22.json')); t=d['train_step']
print('fwd', round(d['value']), 'train ms', round(t['ms_per_step'],3), 'serial', round(t['serial_kernel_ms'],3), 'launches', t['gpu_launches_per_step'])
for k,v in sorted(t['kernel_breakdown'].items(), key=lambda x:-x[1]['ms']): print('  %-28s %4d  %.3f ms  %.1f us/launch' % (k, v['launches'], v['ms'], 1e3*v['ms']/v['launches']))
PY
