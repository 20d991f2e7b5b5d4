timeout 300 python -m pytest tests/test_gpu_kernels.py -q -x -p no:cacheprovider -k "fused_depthwise" 2>&1 | grep -v "^$" | tail -3
timeout 300 python -m pytest tests/test_gpu_model.py -q -x -p no:cacheprovider 2>&1 | tail -2
CTN_DBGS=0,1,5 CTN_DILS=1,128 timeout 300 python tools/bench_fused.py 64 2>&1 | grep "dwpw\|unfused"
for f in 1073741824; do timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-train --fuse $f > gpurun_out/bench_fuse$f.log 2>&1; python - <<PY
import json
d=json.loads(open("gpurun_out/bench_fuse$f.log").read().strip().splitlines()[-1])
print("fuse $f", round(d["value"]), d["clocks"]["sm_mhz"], {k:round(v["ms"],3) for k,v in d["kernel_breakdown"].items()})
PY
done
