python -m pytest tests/test_gpu_kernels.py tests/test_gpu_train.py tests/test_gpu_optim.py -q -m gpu -x 2>&1 | tail -3
python tools/profile_step_phases.py 3
python bench.py --steps 4 --warmup 3 2>&1 | tail -1 > gpurun_out/x17.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/x17.json')); t=d['train_step']
print('fwd', round(d['value']), 'train ms', round(t['ms_per_step'],3), 'serial', round(t['serial_kernel_ms'],3))
PY
