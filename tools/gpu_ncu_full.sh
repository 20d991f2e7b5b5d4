set -x
mkdir -p gpurun_out
# forward launches per pass (mb16, paper): enc, bott, 32 x (conv1x1, dw, pointwise), mask, dec = 100; pass 1 = warm-up
timeout 300 python tools/profile_forward.py 16 2 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 70 -c 2 -o gpurun_out/prof_gemm python tools/profile_forward.py 16 2 > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dwconv -s 40 -c 1 -o gpurun_out/prof_dw python tools/profile_forward.py 16 2 > gpurun_out/ncu3.log 2>&1
ls -la gpurun_out
