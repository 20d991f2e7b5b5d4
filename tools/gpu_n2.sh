mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/smi.txt
timeout 600 python bench.py --gpus 1 --steps 3 --warmup 3 > gpurun_out/bench_n1.log 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.log 2>&1
tail -n 2 gpurun_out/bench_n1.log gpurun_out/bench_n2.log | cut -c1-400
