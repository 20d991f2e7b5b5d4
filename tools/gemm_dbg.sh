timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/tests.log 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench.log 2>&1
