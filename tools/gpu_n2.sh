mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -x -p no:cacheprovider > gpurun_out/multi.log 2>&1
tail -n 5 gpurun_out/multi.log
