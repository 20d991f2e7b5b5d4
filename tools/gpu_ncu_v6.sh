# v6 profiles: ncu launch list of the bench command, full-set capture of the forward kernels (micro-batch 64)
set -x
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu --no-train > gpurun_out/bench_short.json 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2400 -c 820 --csv --log-file gpurun_out/launches_bench_v6.csv python bench.py --steps 1 --warmup 3 --no-cpu --no-train > gpurun_out/ncu_list.log 2>&1
timeout 300 python tools/profile_forward.py 64 2 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|dwpw|resid_update|encoder|decoder" -s 100 -c 14 -o gpurun_out/prof_fwd_v6 python tools/profile_forward.py 64 2 > gpurun_out/ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"decoder" -s 1 -c 1 -o gpurun_out/prof_dec_v6 python tools/profile_forward.py 64 2 > gpurun_out/ncu_full2.log 2>&1
ls -la gpurun_out | tail -8
