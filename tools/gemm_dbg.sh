timeout 600 python -m pytest tests/test_gpu_kernels.py -q -x -p no:cacheprovider -k "fused_update" 2>&1 | tail -3
for pf in 0 1 0 1; do timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-train --fuse-update $pf > gpurun_out/bench_fu$pf.log 2>&1; python - <<PY
import json
d=json.loads(open("gpurun_out/bench_fu$pf.log").read().strip().splitlines()[-1])
print("fuse-update $pf", round(d["value"]), d["clocks"]["sm_mhz"], {k:round(v["ms"],3) for k,v in d["kernel_breakdown"].items()})
PY
done
