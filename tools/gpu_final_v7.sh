# Round-1 final validation + profiles of the committed kernels (one box)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/tests_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tests_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/bench_n1_v7.json 2> gpurun_out/bench_n1_v7.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_v7.json 2> gpurun_out/bench_ref_v7.err
timeout 600 python tools/bench_configs.py > gpurun_out/bench_configs_v7.jsonl 2> gpurun_out/bench_configs_v7.err
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu --no-train > gpurun_out/bench_short.json 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2400 -c 820 --csv --log-file gpurun_out/launches_bench_v7.csv python bench.py --steps 1 --warmup 3 --no-cpu --no-train > gpurun_out/ncu_list.log 2>&1
timeout 300 python tools/profile_forward.py 64 2 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|dwpw|resid_update|encoder|decoder" -s 100 -c 8 -o gpurun_out/prof_fwd_v7 python tools/profile_forward.py 64 2 > gpurun_out/ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"decoder" -s 1 -c 1 -o gpurun_out/prof_dec_v7 python tools/profile_forward.py 64 2 > gpurun_out/ncu_full2.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_train_v7.csv python tools/profile_train_once.py 3 2 > gpurun_out/ncu_train.log 2>&1
tail -3 gpurun_out/tests_gpu.log; tail -1 gpurun_out/smoke.log; ls -la gpurun_out | tail -8
