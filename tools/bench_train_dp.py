"""Data-parallel training step, weak scaling: every rank trains on its own 3 utterances of 4 s
(BASELINE configs[1], run.sh's batch size per GPU), one NCCL all-reduce of the flat gradient
buffer per step.  Launch with torchrun (one process per GPU) or plainly for one GPU:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tools/bench_train_dp.py [cfg2|cfg5] [--steps K] [--warmup W]

Timed region per step: host inputs -> device, forward, PIT SI-SNR loss, backward (with the all-reduce
inside), clip(5) + Adam, loss -> host.  Barrier + device events, max over ranks; rank 0 prints one JSON line."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "conv-tasnet_b200"), os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

import bench_configs as bc  # noqa: E402


def main():
    args = sys.argv[1:]
    name = "cfg5" if "cfg5" in args else "cfg2"
    steps = int(args[args.index("--steps") + 1]) if "--steps" in args else 30
    warm = int(args[args.index("--warmup") + 1]) if "--warmup" in args else 8
    cfg, M, T = (bc.BEST, 8, 32000) if name == "cfg5" else (bc.PAPER, 3, 32000)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import pit_criterion
    from ctn_b200 import _dist
    from ctn_b200.optim import FusedAdam
    model = bc.build(cfg, train=True)
    if world > 1:
        _dist.attach(model)
    opt = FusedAdam(model.parameters(), lr=1e-3, max_norm=5.0)
    g = torch.Generator().manual_seed(77 + rank)
    h_src = (0.1 * torch.randn(M, cfg["C"], T, generator=g)).pin_memory()
    h_mix = h_src.sum(1).pin_memory()
    lengths = torch.full((M,), T, dtype=torch.int64, device="cuda")
    h_loss = torch.empty((), dtype=torch.float32).pin_memory()
    losses = []

    def step():
        mix = h_mix.cuda(non_blocking=True)
        src = h_src.cuda(non_blocking=True)
        est = model(mix)
        loss, _, _, _ = pit_criterion.cal_loss(src, est, lengths)
        opt.zero_grad()
        loss.backward()
        opt.step()
        h_loss.copy_(loss.detach(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        losses.append(float(h_loss))

    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda")
    if world > 1:
        dist.barrier()
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        # replicas must stay identical: same reduced gradient, same update
        w = model.state_dict()["separator.network.1.weight"].double().sum().reshape(1)
        lo, hi = w.clone(), w.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        in_sync = bool((lo == hi).item())
    else:
        in_sync = True
    if rank == 0:
        ms = float(ms)
        print(json.dumps({"config": "%s train step, weak scaling: %d x %d utterances of 4 s" % (name, world, M),
                          "n_gpus": world, "ms_per_step": ms, "global_batch": world * M,
                          "audio_s_per_s": world * M * T / bc.SR / (ms / 1e3),
                          "algorithmic_tflops": 3.0 * bc.fwd_flops(cfg, T) * M * world / (ms / 1e3) / 1e12,
                          "steps": steps, "warmup": warm, "replicas_in_sync": in_sync,
                          "loss_first_rank0": losses[0], "loss_last_rank0": losses[-1],
                          "collective": "one ncclAllReduce(sum) of the flat fp32 gradient per step" if world > 1 else None,
                          "timed": "host inputs -> device, fwd + PIT SI-SNR loss + bwd (+ all-reduce) + clip(5) + Adam, "
                                   "loss -> host; barrier + device events, max over ranks"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
