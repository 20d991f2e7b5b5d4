python -m pytest tests/test_gpu_train.py tests/test_gpu_bn.py tests/test_gpu_model.py -q -m gpu -x 2>&1 | tail -4
for r in 1 2; do
python bench.py --steps 4 --warmup 3 2>&1 | tail -1 > gpurun_out/x16.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/x16.json')); t=d['train_step']
print('fwd', round(d['value']), 'train ms', round(t['ms_per_step'],3), 'serial', round(t['serial_kernel_ms'],3))
PY
done
