# Round-1 final validation + profiles (one box): tests, smoke, bench (both arms), ncu launch list and full-set captures
set -x
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/tests_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tests_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_final.json 2> gpurun_out/bench_ref_final.err
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu --no-train > gpurun_out/bench_short.json 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2400 -c 820 --csv --log-file gpurun_out/launches_bench_v5.csv python bench.py --steps 1 --warmup 3 --no-cpu --no-train > gpurun_out/ncu_list.log 2>&1
timeout 300 python tools/profile_forward.py 64 2 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|dwpw|dwconv_march|resid_update" -s 99 -c 24 -o gpurun_out/prof_fwd_v5 python tools/profile_forward.py 64 2 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/tests_gpu.log; cat gpurun_out/smoke.log | tail -2; cut -c1-300 gpurun_out/bench_final.json; ls -la gpurun_out | tail -12
