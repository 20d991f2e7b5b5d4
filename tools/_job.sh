O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_gpu_properties.py tests/test_gpu_bn.py tests/test_gpu_zz_precision.py -m gpu -x -q 2>&1 | tail -5
python bench.py --steps 5 --warmup 3 --no-cpu --no-train > $O/x5_bench.json 2> $O/x5_bench.err; echo rc=$?
python - <<'PY'
import json
d=json.loads(open("gpurun_out/x5_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["clocks"])
for k,v in d["kernel_breakdown"].items(): print("   ",k,v["launches"],round(v["ms"]*1e3/v["launches"],1),"us",round(v["share"],3))
PY
