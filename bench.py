#!/usr/bin/env python
"""Benchmark of the Conv-TasNet separation path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload separate|train]

Default workload = BASELINE.json's headline metric, "separated audio-sec/s" on the paper
configuration (N256 L20 B256 H512 P3 X8 R4 C2 gLN, 4 s @ 8 kHz): bulk inference of 512
synthetic mixtures (BASELINE configs[3]), sharded by utterance over the N GPUs (strong
scaling, no data-path collective).  One "step" = one pass over the 512 mixtures.

  value   : audio-seconds separated per second, inputs already resident in HBM
  e2e     : the same through the public API ConvTasNet.forward with HOST buffers: pinned
            host -> device copy of the mixtures and device -> host copy of the estimates
            inside the timed region
  roofline: the dominant kernel (tcgen05 1x1-conv GEMM), algorithmic FLOPs / CUDA-event
            time per launch, against the measured bf16 peak in MEASURED_PEAKS.json
  cpu_baseline / --impl reference: the oracle port of the reference's CPU path
            (torch CPU ops, all host threads) on a bounded sample of the same workload

Rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "conv-tasnet_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

PAPER = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
SR = 8000
T_4S = 4 * SR
GLOBAL_BATCH = 512
METRIC = "separated audio-sec/s (N256 L20 B256 H512 P3 X8 R4 gLN, 4 s @ 8 kHz)"
UNIT = "audio-s/s"


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return dict(hbm=d.get("hbm_gbs", 6650.0), tf_burst=d.get("bf16_tflops", 1590.0),
                    tf_sustained=d.get("bf16_tflops_sustained", 1400.0), source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


def forward_flops(cfg, T):
    N, L, B, H, P, X, R, C = (cfg[k] for k in "NLBHPXRC")
    K = (T - L) // (L // 2) + 1
    return 2.0 * K * (N * L + N * B + R * X * (2 * B * H + H * P) + B * C * N + C * N * L)


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU while the timed region runs."""

    def __init__(self, index, period=0.1):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nvml = None

    def run(self):
        if self.nvml is None:
            return
        n = self.nvml
        names = {
            getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(n, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM))
                try:
                    mask = n.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(self.period)

    def finish(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def physical_gpu_index(local):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local])
        except Exception:
            return local
    return local


# ------------------------------------------------------------------ CPU arm
def oracle_cpu_throughput(cfg, T, n_utt, steps, warmup):
    """Times the oracle port of the reference forward on the host cores."""
    from oracle import convtasnet_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    params = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=1234)
    g = torch.Generator().manual_seed(2025)
    mix = 0.1 * torch.randn(n_utt, T, generator=g)
    with torch.no_grad():
        for _ in range(warmup):
            O.forward(mix, params, cfg)
        t0 = time.perf_counter()
        for _ in range(steps):
            O.forward(mix, params, cfg)
        dt = time.perf_counter() - t0
    return n_utt * steps * (T / SR) / dt, dt / steps, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_utt = 4
    steps = max(1, min(args.steps, 10))
    warm = max(1, min(args.warmup, 2))
    v, step_s, cores = oracle_cpu_throughput(PAPER, T_4S, n_utt, steps, warm)
    sample = "%d x 4 s mixtures per step (bounded sample of the 512-mixture workload), fp32, torch CPU ops" % n_utt
    out = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "bulk separation, paper config, 4 s @ 8 kHz mixtures", "sample_utterances": n_utt},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


# ------------------------------------------------------------------ GPU arm
def run_b200(args):
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import conv_tasnet
    from oracle import convtasnet_oracle as O

    cfg = PAPER
    T = T_4S
    n_local = GLOBAL_BATCH // world
    params = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=1234)  # same weights as the CPU arm
    model = conv_tasnet.ConvTasNet(**cfg)
    model.load_state_dict(params)
    model = model.to(dev).eval()
    eng = model.engine()
    eng.micro_batch = args.micro_batch
    eng.fuse_max_dilation = args.fuse
    eng.fuse_update = bool(args.fuse_update)
    eng.use_graph = bool(args.graph)
    eng.n_streams = args.streams

    g = torch.Generator().manual_seed(2025 + rank)
    host_mix = (0.1 * torch.randn(n_local, T, generator=g)).pin_memory()
    host_est = torch.empty(n_local, cfg["C"], T).pin_memory()
    mix = host_mix.to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        barrier()
        return float(ms.item()) / 1e3

    def step_resident():
        with torch.no_grad():
            model(mix)

    def step_e2e():
        # the public host-to-host call: pinned mixtures in, pinned estimates out, copies of neighbouring
        # micro-batches overlapped with the kernels (ConvTasNet.separate -> SeparationEngine.forward_host)
        model.separate(host_mix, out=host_est)

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(physical_gpu_index(local))
    sampler.start()
    l0 = eng.launches
    sec = timed(step_resident, args.steps)
    launches = eng.launches - l0
    clocks = sampler.finish()
    audio_s = GLOBAL_BATCH * (T / SR) * args.steps
    value = audio_s / sec

    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    sec_e2e = timed(step_e2e, args.steps)
    e2e = {"value": audio_s / sec_e2e, "unit": UNIT, "h2d_bytes_per_step": GLOBAL_BATCH * T * 4,
           "d2h_bytes_per_step": GLOBAL_BATCH * cfg["C"] * T * 4, "ms_per_step": sec_e2e / args.steps * 1e3}

    # per-kernel times of one micro-batch pass, CUDA events around every launch (rank 0)
    roof, breakdown = None, None
    if rank == 0:
        pk = peaks()
        mb = min(args.micro_batch, n_local)
        plan = eng._plan({k: p.detach() for k, p in model._params().items()}, mb, T, dev)
        plan.mix.copy_(mix[:mb])
        plan.launch()
        torch.cuda.synchronize(dev)
        acc = {}
        for _ in range(3):
            for kind, fl, by, ms in plan.launch_timed():
                a = acc.setdefault(kind, [0.0, 0.0, 0.0, 0])
                a[0] += fl
                a[1] += by
                a[2] += ms
                a[3] += 1
        tot_ms = sum(a[2] for a in acc.values())
        breakdown = {k: {"launches": a[3] // 3, "ms": a[2] / 3, "share": a[2] / tot_ms,
                         "tflops": a[0] / a[2] / 1e9, "gbs": a[1] / a[2] / 1e6} for k, a in acc.items()}
        # one roofline entry per kernel kind: tensor-bound kinds against the sustained bf16 peak
        # (algorithmic FLOPs = ONE product; the hi/lo split executes three MMAs per product, reported
        # as executed_*), HBM-bound kinds against the measured copy bandwidth.  traffic = DRAM bytes per
        # launch from the committed ncu --set full capture (valid at the captured micro-batch only).
        traffic = {}
        try:
            with open(os.path.join(ROOT, "profiles", "r1", "ncu_traffic_v7.json")) as f:
                tj = json.load(f)
            if tj.get("micro_batch") == mb:
                traffic = {k: v["traffic_bytes"] for k, v in tj["kernels"].items()}
        except Exception:
            traffic = {}
        names = {"gemm_conv1x1": "gemm_tc_kernel<256,pair,plain> (conv1x1 + PReLU statistics)",
                 "gemm_pointwise": "gemm_tc_kernel<256,pair,affine> (pointwise + folded gLN + residual)",
                 "gemm_dwpw_fused": "fb::dwpw_kernel (depthwise producers + pointwise GEMM)",
                 "gemm_update_conv1x1": "c1::updc1_kernel (residual-update producers + conv1x1 GEMM + PReLU statistics)",
                 "gemm_bottleneck": "gemm_tc_kernel (bottleneck)", "gemm_mask": "gemm_tc_kernel (mask conv)",
                 "dwconv": "dwconv_march_kernel", "resid_update": "fb::resid_update_kernel",
                 "encoder": "encoder_fwd_kernel", "decoder": "decoder_fwd_blocked_kernel"}
        roofs = {}
        for k, a in acc.items():
            n = a[3]
            ms = a[2] / n
            tensor = k.startswith("gemm")
            if tensor:
                ach = a[0] / a[2] / 1e9
                r = {"bound": "tensor", "achieved": ach, "peak": pk["tf_sustained"], "unit": "TFLOP/s",
                     "frac": ach / pk["tf_sustained"], "executed_tflops": 3.0 * ach,
                     "executed_frac": 3.0 * ach / pk["tf_sustained"],
                     "peak_source": pk["source"] + " (sustained bf16)", "algorithmic_flops_per_launch": a[0] / n,
                     "hbm_gbs": a[1] / a[2] / 1e6, "hbm_frac": a[1] / a[2] / 1e6 / pk["hbm"]}
            else:
                ach = a[1] / a[2] / 1e6
                r = {"bound": "hbm", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"],
                     "peak_source": pk["source"] + " (copy bandwidth)"}
            r.update({"kernel": names.get(k, k), "avg_launch_ms": ms, "algorithmic_bytes_per_launch": a[1] / n,
                      "share_of_step": a[2] / tot_ms, "traffic": traffic.get(k)})
            roofs[k] = r
        dom = max(roofs, key=lambda k: roofs[k]["share_of_step"])
        roof = dict(roofs[dom])
        roof["how"] = ("dominant kernel by share of the step; CUDA events around every launch of one %d-utterance "
                       "micro-batch pass, 3 repeats; traffic from profiles/r1/ncu_traffic_v7.json" % mb)
        roof["all_kernels"] = roofs

    train = None
    if not args.no_train and _precision() == "bf16x3":     # the fp16 inference precision has no training path
        train = train_step(args, world, rank, dev, timed)
        train["ms_per_step_torch_optimizer"] = train_step(args, world, rank, dev, timed, fused_optim=False)["ms_per_step"]

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, step_s, cores = oracle_cpu_throughput(cfg, T, 4, 3, 1)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "4 x 4 s mixtures per step, 3 steps after 1 warm-up, fp32 torch CPU ops (oracle port)"}

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None,
            "dtype": ("bf16 hi+lo split operands (3 MMAs), fp32 accumulate; fp32 elsewhere" if _precision() == "bf16x3"
                      else "fp16 operands (1 MMA), fp32 accumulate; fp32 elsewhere (CTN_PRECISION=fp16, opt-in)"),
            "data": "synthetic",
            "config": {"workload": "bulk separation of %d x 4 s @ 8 kHz mixtures, paper config "
                                   "(BASELINE configs[3]), sharded by utterance" % GLOBAL_BATCH,
                       "global_batch": GLOBAL_BATCH, "per_gpu": n_local, "micro_batch": args.micro_batch,
                       "cuda_graph": bool(args.graph), "streams": args.streams, "parallelism": "utterance-sharded x%d, no collective" % world,
                       "l2": "activation working set per step (%.1f GB) >> 126 MB L2" % (n_local * 33e6 / 1e9),
                       "x_realtime_per_gpu": value / world},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "algorithmic_tflops": forward_flops(cfg, T) * GLOBAL_BATCH * args.steps / sec / 1e12,
            "roofline": roof, "kernel_breakdown": breakdown,
        }
        if cpu is not None:
            out["cpu_baseline"] = cpu
        if train is not None:
            out["train_step"] = train
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _precision():
    from ctn_b200 import _native as nat
    return nat.PRECISION


def train_step(args, world, rank, dev, timed, fused_optim=True):
    """BASELINE configs[1]: paper config, batch 3 x 4 s per GPU (weak scaling), forward +
    cal_loss + backward (ONE NCCL all-reduce of the flat gradient when world > 1) + clip(5) +
    Adam(1e-3), as solver.py:178-186.  Host mixtures/sources are copied in every step and the
    loss is read back (solver.py:188 does loss.item())."""
    import conv_tasnet
    import pit_criterion
    from ctn_b200 import _dist
    from oracle import convtasnet_oracle as O
    cfg, T, M = PAPER, T_4S, 3
    params = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=1234)
    model = conv_tasnet.ConvTasNet(**cfg)
    model.load_state_dict(params)
    model = model.to(dev).train()
    if world > 1:
        _dist.attach(model)
    if fused_optim:   # clip(5) + Adam in two launches (ctn_clip_adam)
        from ctn_b200.optim import FusedAdam
        opt = FusedAdam(model.parameters(), lr=1e-3, max_norm=5.0)
    else:
        opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    g = torch.Generator().manual_seed(77 + rank)
    h_src = (0.1 * torch.randn(M, cfg["C"], T, generator=g)).pin_memory()
    h_mix = h_src.sum(1).pin_memory()
    lengths = torch.full((M,), T, dtype=torch.int64, device=dev)
    h_loss = torch.empty((), dtype=torch.float32).pin_memory()
    eng = model.train_engine()

    def step():
        mix = h_mix.to(dev, non_blocking=True)
        src = h_src.to(dev, non_blocking=True)
        est = model(mix)
        loss, _, _, _ = pit_criterion.cal_loss(src, est, lengths)
        opt.zero_grad()
        loss.backward()
        if not fused_optim:
            torch.nn.utils.clip_grad_norm_(model.parameters(), 5.0)
        opt.step()
        h_loss.copy_(loss.detach(), non_blocking=True)

    for _ in range(max(3, args.warmup)):
        step()
    l0 = eng.launches
    steps = max(args.steps, 5)
    sec = timed(step, steps)
    return {"ms_per_step": sec / steps * 1e3, "steps": steps, "utterances_per_gpu": M, "global_batch": M * world,
            "scaling": "weak", "audio_s_per_s": M * world * (T / SR) * steps / sec,
            "algorithmic_tflops": 3.0 * forward_flops(cfg, T) * M * world * steps / sec / 1e12,
            "gpu_launches_per_step": (eng.launches - l0) // steps, "loss": float(h_loss),
            "optimizer": "ctn_b200.optim.FusedAdam (clip + Adam, 2 launches)" if fused_optim
                         else "torch clip_grad_norm_ + torch.optim.Adam",
            "config": "BASELINE configs[1]: paper config, 3 x 4 s per GPU, fwd + PIT SI-SNR loss + bwd + "
                      "grad all-reduce + clip(5) + Adam, host inputs copied every step",
            "h2d_bytes_per_step": M * T * 4 * (1 + cfg["C"]), "d2h_bytes_per_step": 4}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--micro-batch", type=int, default=64)
    ap.add_argument("--fuse", type=int, default=1 << 30,
                    help="blocks with dilation <= this use the fused depthwise+pointwise kernel (0 = off)")
    ap.add_argument("--fuse-update", type=int, default=0,
                    help="fold the deferred residual update into the next conv1x1 (ctn_updc1_fwd)")
    ap.add_argument("--graph", type=int, default=1)
    ap.add_argument("--streams", type=int, default=1,
                    help="micro-batches in flight on separate streams (engine.n_streams)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-train", action="store_true", help="skip the train-step leg")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
