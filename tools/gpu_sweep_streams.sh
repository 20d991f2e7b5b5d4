# forward throughput with 1/2/3 micro-batch chains in flight, micro-batch 32/64, and the folded residual update
mkdir -p gpurun_out
for cfg in "--streams 1" "--streams 2" "--streams 3" "--streams 2 --micro-batch 32" "--streams 4 --micro-batch 32" "--streams 1 --fuse-update 1" "--streams 2 --fuse-update 1"; do
  echo "== $cfg"
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-train $cfg 2> gpurun_out/sweep.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']), d['clocks'])"
done
