export CTN_KINDS=pointwise
for d in 0 4 6 36 38; do echo "== DBG $d"; CTN_GEMM_DBG=$d timeout 200 python tools/bench_gemm.py 16 2; done > gpurun_out/bench_gemm.log 2>&1
