set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_model.py -q --tb=short -p no:cacheprovider > gpurun_out/model.log 2>&1; echo "exit $?" >> gpurun_out/model.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>&1; echo "exit $?" >> gpurun_out/bench.log
for mb in 2 4 8 32; do
  timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --micro-batch $mb > gpurun_out/bench_mb$mb.log 2>&1
done
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --micro-batch 16 --graph 0 > gpurun_out/bench_nograph.log 2>&1
timeout 300 python tools/profile_forward.py 16 3 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 110 -c 200 --csv --log-file gpurun_out/launches.csv python tools/profile_forward.py 16 3 > gpurun_out/ncu1.log 2>&1
timeout 300 python tools/profile_forward.py 16 2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 70 -c 3 -o gpurun_out/prof_gemm python tools/profile_forward.py 16 2 > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dwconv -s 35 -c 2 -o gpurun_out/prof_dw python tools/profile_forward.py 16 2 > gpurun_out/ncu3.log 2>&1
ls -la gpurun_out
