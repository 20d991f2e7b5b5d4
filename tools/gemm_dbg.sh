timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-train --micro-batch 64 > gpurun_out/bench_a.log 2>&1
