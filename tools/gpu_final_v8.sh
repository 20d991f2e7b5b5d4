# Round-1 final validation of the committed state (one box): tests, smoke, both bench arms, the other
# BASELINE configurations, launch list of a training step
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/tests_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tests_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/bench_n1_v8.json 2> gpurun_out/bench_n1_v8.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_v8.json 2> gpurun_out/bench_ref_v8.err
timeout 600 python tools/bench_configs.py > gpurun_out/bench_configs_v8.jsonl 2> gpurun_out/bench_configs_v8.err
timeout 300 python tools/profile_train_once.py 3 2 > gpurun_out/train_once.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_train_v8.csv python tools/profile_train_once.py 3 2 > gpurun_out/ncu_train.log 2>&1
tail -3 gpurun_out/tests_gpu.log; tail -1 gpurun_out/smoke.log; cut -c1-300 gpurun_out/bench_n1_v8.json; cut -c1-200 gpurun_out/bench_ref_v8.json; cut -c1-160 gpurun_out/bench_configs_v8.jsonl
