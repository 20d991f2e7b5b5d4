timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "fused_update" 2>&1 | tail -5
CTN_DBGS=0,1,2,3,8,11 timeout 120 python tools/bench_updc1.py 64
