timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/tests.log 2>&1
