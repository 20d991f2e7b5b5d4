timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_kernels.py -q -x -p no:cacheprovider -s > gpurun_out/tests.log 2>&1
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-train > gpurun_out/bench_a.log 2>&1
