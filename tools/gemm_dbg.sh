timeout 300 python -m pytest tests/test_gpu_kernels.py -q -x -p no:cacheprovider -k "encoder" 2>&1 | tail -3
timeout 300 python -m pytest tests/test_gpu_model.py tests/test_gpu_train.py -q -x -p no:cacheprovider 2>&1 | tail -2
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-train > gpurun_out/bench_x.log 2>&1; python - <<PY
import json
d=json.loads(open("gpurun_out/bench_x.log").read().strip().splitlines()[-1])
print(round(d["value"]), d["clocks"]["sm_mhz"], {k:round(v["ms"],3) for k,v in d["kernel_breakdown"].items()})
PY
