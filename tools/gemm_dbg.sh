timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py -q -x -p no:cacheprovider > gpurun_out/tests.log 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-train > gpurun_out/bench.log 2>&1
