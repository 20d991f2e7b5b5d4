# First GPU call of round 2 (one box): confirm the round-1 state, then the numbers round 2 starts from.
#   gpurun --timeout 900 -- 'bash tools/gpu_round2_start.sh'
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/tests_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tests_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/bench_n1_r2_start.json 2> gpurun_out/bench_n1_r2_start.err
CTN_PRECISION=fp16 timeout 300 python bench.py --no-cpu > gpurun_out/bench_n1_fp16_r2_start.json 2> gpurun_out/bench_n1_fp16_r2_start.err
timeout 600 python tools/bench_configs.py > gpurun_out/bench_configs_r2_start.jsonl 2> gpurun_out/bench_configs_r2_start.err
tail -3 gpurun_out/tests_gpu.log; tail -1 gpurun_out/smoke.log
cut -c1-260 gpurun_out/bench_n1_r2_start.json; cut -c1-260 gpurun_out/bench_n1_fp16_r2_start.json; cut -c1-170 gpurun_out/bench_configs_r2_start.jsonl
