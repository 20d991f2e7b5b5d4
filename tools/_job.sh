timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_optim.py tests/test_gpu_bn.py -m gpu -x -q 2>&1 | tail -12
timeout 200 python tools/exp_split_train.py 3 2>&1 | head -1
