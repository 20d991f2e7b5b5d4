set -x
mkdir -p gpurun_out
timeout 300 python tools/profile_forward.py 16 2 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 100 --csv --log-file gpurun_out/launches_fwd_v3.csv python tools/profile_forward.py 16 2 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 68 -c 2 -o gpurun_out/prof_gemm_v3 python tools/profile_forward.py 16 2 > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dwconv -s 34 -c 1 -o gpurun_out/prof_dw_v3 python tools/profile_forward.py 16 2 > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:decoder -s 1 -c 1 -o gpurun_out/prof_dec_v3 python tools/profile_forward.py 16 2 > gpurun_out/ncu4.log 2>&1
timeout 300 python tools/profile_train_once.py 3 3 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bwd_dw_march -s 40 -c 1 -o gpurun_out/prof_bwd_dw_v3 python tools/profile_train_once.py 16 2 > gpurun_out/ncu5.log 2>&1
ls -la gpurun_out
