# weak-scaled training step at N = $2 (default 1), 2*$2, .. $1 GPUs of the box
set -x
N=${1:-2}
mkdir -p gpurun_out
n=${2:-1}
while [ $n -le $N ]; do
  if [ $n -eq 1 ]; then
    timeout 200 python tools/bench_train_dp.py cfg2 >> gpurun_out/train_dp.jsonl 2> gpurun_out/train_dp_$n.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 tools/bench_train_dp.py cfg2 >> gpurun_out/train_dp.jsonl 2> gpurun_out/train_dp_$n.err
  fi
  echo "n=$n rc=$?"; tail -3 gpurun_out/train_dp_$n.err
  n=$((n * 2))
done
cat gpurun_out/train_dp.jsonl
