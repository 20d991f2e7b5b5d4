# v8: full-set ncu capture of the training step's kernels at batch 3 (one launch of each family, mid-network)
set -x
mkdir -p gpurun_out
timeout 300 python tools/profile_train_once.py 3 2 > gpurun_out/train_once.log 2>&1 &&
timeout 700 ncu --set full --clock-control none --import-source on -k regex:"bwd_dw_march|wgrad_tc|bwd_prelu_norm|norm_split|reduce_partials|dwconv_march" -s 60 -c 8 -o gpurun_out/prof_train_v8 python tools/profile_train_once.py 3 2 > gpurun_out/ncu_train_full.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc" -s 120 -c 5 -o gpurun_out/prof_train_gemm_v8 python tools/profile_train_once.py 3 2 > gpurun_out/ncu_train_full2.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/ncu_train_full.log
