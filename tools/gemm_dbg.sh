timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_train.py -q -x -k "gemm or wgrad or gradients" -p no:cacheprovider > gpurun_out/gemm_tests.log 2>&1
for d in 0; do echo "== DBG $d"; CTN_GEMM_DBG=$d timeout 200 python tools/bench_gemm.py 16 1 2; done > gpurun_out/bench_gemm.log 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench.log 2>&1
