set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -q --maxfail=30 --tb=short -p no:cacheprovider > gpurun_out/kernels.log 2>&1; echo "exit $?" >> gpurun_out/kernels.log
timeout 600 python -m pytest tests/test_gpu_model.py -q --maxfail=30 --tb=short -p no:cacheprovider -s > gpurun_out/model.log 2>&1; echo "exit $?" >> gpurun_out/model.log
timeout 900 python -m pytest tests/test_gpu_train.py -q --maxfail=30 --tb=short -p no:cacheprovider -s > gpurun_out/train.log 2>&1; echo "exit $?" >> gpurun_out/train.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>&1; echo "exit $?" >> gpurun_out/bench.log
for mb in 4 8 32; do
  timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --micro-batch $mb > gpurun_out/bench_mb$mb.log 2>&1
done
tail -4 gpurun_out/kernels.log gpurun_out/model.log gpurun_out/train.log
