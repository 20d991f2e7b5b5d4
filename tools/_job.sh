O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_properties.py tests/test_gpu_bn.py tests/test_gpu_stream.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 8 --warmup 3 --no-cpu --no-train > $O/x10_bench.json 2> $O/x10_bench.err; echo rc=$?
python - <<'PY'
import json
d=json.loads(open("gpurun_out/x10_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["clocks"], d["gpu_launches"])
for k,v in d["kernel_breakdown"].items(): print("   ",k,v["launches"],round(v["ms"]*1e3/v["launches"],1),"us",round(v["share"],3))
r=d["roofline"]; print({k:r[k] for k in r if k not in ("all_kernels","how")})
PY
timeout 300 python tools/bench_configs.py 2>&1 | tail -8
