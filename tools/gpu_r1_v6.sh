# Round-1 re-entry validation of HEAD (one box): tests, smoke, bench (both arms), other BASELINE configs
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/tests_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tests_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/bench_n1_v6.json 2> gpurun_out/bench_n1_v6.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_v6.json 2> gpurun_out/bench_ref_v6.err
timeout 600 python tools/bench_configs.py > gpurun_out/bench_configs_v6.jsonl 2> gpurun_out/bench_configs_v6.err
tail -3 gpurun_out/tests_gpu.log; tail -2 gpurun_out/smoke.log; cut -c1-400 gpurun_out/bench_n1_v6.json; cat gpurun_out/bench_configs_v6.jsonl; tail -5 gpurun_out/bench_configs_v6.err
