timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "fused" 2>&1 | tail -8
