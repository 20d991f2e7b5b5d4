#!/bin/bash
# The GPU-side jobs of a round, one entry point:  gpurun --timeout S -- 'bash tools/gpu_job.sh <job> [tag]'
#   validate   pytest -m gpu, smoke, both bench arms (no profiler)
#   ncu        validate-lite (plain runs) && launch lists + --set full captures of the hot kernels (forward mb128, train batch 3)
#   sanitize   compute-sanitizer memcheck over the small-fixture GPU tests (racecheck: job "racecheck")
#   multi      the 2-GPU NCCL tests + 2-rank bench (needs gpurun --gpus 2)
# Everything is written under gpurun_out/<tag>_*; summaries are made HERE afterwards (tools/ncu_summary.py ...).
job=${1:-validate}; tag=${2:-r2}
O=gpurun_out; mkdir -p $O
export PYTHONUNBUFFERED=1
case $job in
validate)
  timeout 1500 python -m pytest tests -m gpu -x -q > $O/${tag}_tests_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${tag}_tests_gpu.log
  tail -5 $O/${tag}_tests_gpu.log
  timeout 300 python __graft_entry__.py --smoke > $O/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/${tag}_smoke.log
  timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > $O/${tag}_bench_ref.json 2> $O/${tag}_bench_ref.err; echo "ref rc=$?"
  timeout 900 python bench.py --steps 10 --warmup 3 > $O/${tag}_bench_n1.json 2> $O/${tag}_bench_n1.err; echo "bench rc=$?"
  tail -c 600 $O/${tag}_bench_n1.err; cut -c1-400 $O/${tag}_bench_n1.json
  ;;
ncu)
  # forward, one 128-utterance micro-batch, 2 passes of 68 launches (enc, bott, conv1x1, 31 x (dwpw, update+conv1x1), dwpw,
  # update+mask, dec)
  timeout 300 python tools/profile_forward.py 128 2 > $O/${tag}_plain_fwd.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -s 68 -c 68 --csv --log-file $O/${tag}_launches_fwd.csv \
      python tools/profile_forward.py 128 2 > $O/${tag}_ncu_fwd_list.log 2>&1
  for k in dwpw_kernel:33:2 updc1_kernel:33:2 gemm_tc_kernel:2:2 encoder_fwd:1:1 decoder_fwd:1:1; do
    IFS=: read name skip cnt <<< "$k"
    ncu --set full --clock-control none --import-source on -k regex:$name -s $skip -c $cnt -o $O/${tag}_prof_fwd_$name -f \
        python tools/profile_forward.py 128 2 > $O/${tag}_ncu_fwd_$name.log 2>&1
  done
  # training, batch 3 (BASELINE configs[1]), 2 steps; the second step's launches
  timeout 300 python tools/profile_train_once.py 3 2 > $O/${tag}_plain_train.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/${tag}_launches_train.csv \
      python tools/profile_train_once.py 3 2 > $O/${tag}_ncu_train_list.log 2>&1
  for k in wgrad_tc_kernel:70:2 gemm_tc_kernel:80:4 bwd_dw_march:34:1 dwpw_kernel:34:1 updc1_kernel:34:1; do
    IFS=: read name skip cnt <<< "$k"
    ncu --set full --clock-control none --import-source on -k regex:$name -s $skip -c $cnt -o $O/${tag}_prof_train_$name -f \
        python tools/profile_train_once.py 3 2 > $O/${tag}_ncu_train_$name.log 2>&1
  done
  ls -la $O | tail -30
  ;;
sanitize|racecheck)
  tool=memcheck; [ $job = racecheck ] && tool=racecheck
  # the small-fixture tests: every kernel family at sizes the sanitizer finishes in minutes
  timeout 2400 compute-sanitizer --tool $tool --error-exitcode 99 --log-file $O/${tag}_sanitizer_$tool.log \
      python -m pytest -x -q -m gpu tests/test_gpu_kernels.py tests/test_gpu_backward_kernels.py \
      "tests/test_gpu_model.py::test_forward_matches_reference_fixture" "tests/test_gpu_model.py::test_loss_matches_reference_fixture" \
      "tests/test_gpu_train.py::test_gradients_match_reference_fixture" tests/test_gpu_optim.py::test_matches_torch_clip_and_adam \
      > $O/${tag}_sanitizer_${tool}_pytest.log 2>&1
  echo "sanitizer($tool) rc=$?"; tail -3 $O/${tag}_sanitizer_${tool}_pytest.log; tail -5 $O/${tag}_sanitizer_$tool.log
  ;;
multi)
  nvidia-smi -L
  timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -rA > $O/${tag}_tests_multi.log 2>&1; echo "multi rc=$?"
  tail -15 $O/${tag}_tests_multi.log
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
      bench.py --gpus 2 --steps 10 --warmup 3 > $O/${tag}_bench_n2.json 2> $O/${tag}_bench_n2.err; echo "bench n2 rc=$?"
  cut -c1-300 $O/${tag}_bench_n2.json
  ;;
*) echo "unknown job $job"; exit 2;;
esac
