O=gpurun_out
python -m pytest tests/test_gpu_train.py tests/test_gpu_kernels.py -q -m gpu -x 2>&1 | tail -4
export CTN_ALLOW_EXPERIMENT_BUILD=1
for r in 1 2; do
for l in libctn_base.so libctn_b200.so; do
echo "== $l run $r"; CTN_LIB_FILE=$l python bench.py --steps 4 --warmup 3 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); t=d['train_step']
print('fwd', round(d['value']), 'train ms', round(t['ms_per_step'],3), 'serial', round(t['serial_kernel_ms'],3))
for k in t.get('all_kernels', t.get('kernels', []))[:12]: print('   ', k)
"
done; done
