mkdir -p gpurun_out
for n in 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n${n}_v6.json 2> gpurun_out/bench_n${n}_v6.err
done
for n in 8; do python - <<PY
import json
l=[x for x in open("gpurun_out/bench_n${n}_v6.json") if x.startswith("{")][-1]
d=json.loads(l); print(d["n_gpus"], round(d["value"]), round(d["e2e"]["value"]), d["clocks"], d["train_step"]["ms_per_step"], d["train_step"]["global_batch"])
PY
done
