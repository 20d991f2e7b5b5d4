timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "fused_depthwise" 2>&1 | tail -2
for i in 1 2; do
CTN_DILS=1,8,128 timeout 200 python tools/bench_fused.py 64 2>&1 | grep dwpw
CTN_ALLOW_EXPERIMENT_BUILD=1 CTN_LIB_FILE=libctn_exp.so CTN_DILS=1,8,128 timeout 200 python tools/bench_fused.py 64 2>&1 | grep dwpw
done
