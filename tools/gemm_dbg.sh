timeout 900 python -m pytest tests/test_gpu_backward_kernels.py tests/test_gpu_train.py -q -x -p no:cacheprovider > gpurun_out/tests.log 2>&1
python tools/profile_train.py 3 > gpurun_out/ptrain.log 2>&1;  python tools/profile_train.py 16 >> gpurun_out/ptrain.log 2>&1
