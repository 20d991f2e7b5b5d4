set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 300 python tests/gpu_diag_gemm.py > gpurun_out/diag.log 2>&1; echo "diag exit $?" >> gpurun_out/diag.log
timeout 900 python -m pytest tests/test_gpu_kernels.py -q --maxfail=40 --tb=short -p no:cacheprovider > gpurun_out/kernels.log 2>&1; echo "exit $?" >> gpurun_out/kernels.log
timeout 900 python -m pytest tests/test_gpu_model.py -q --maxfail=40 --tb=short -p no:cacheprovider -s > gpurun_out/model.log 2>&1; echo "exit $?" >> gpurun_out/model.log
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.log 2>&1; echo "exit $?" >> gpurun_out/bench.log
tail -5 gpurun_out/diag.log gpurun_out/kernels.log gpurun_out/model.log gpurun_out/bench.log
