#!/usr/bin/env python
"""Benchmark of the Conv-TasNet separation path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Workload = BASELINE.json's headline metric, "separated audio-sec/s" on the paper configuration
(N256 L20 B256 H512 P3 X8 R4 C2 gLN, 4 s @ 8 kHz): bulk inference of 512 synthetic mixtures
(BASELINE configs[3]), sharded by utterance over the N GPUs (strong scaling, no data-path
collective).  One "step" = one pass over the 512 mixtures.  The metric's second half, "train
step ms" (BASELINE configs[1], 3 x 4 s per GPU, solver.py:178-186), is the `train_step` object.

  value        : audio-seconds separated per second, inputs already resident in HBM
  e2e          : the same through the public host-to-host call ConvTasNet.separate: pinned host ->
                 device copy of every step's mixtures and device -> host copy of its estimates inside
                 the timed region
  roofline     : the dominant kernel of the step: algorithmic FLOPs (bytes) per launch / CUDA-event
                 time per launch against MEASURED_PEAKS.json; train_step carries its own
  cpu_baseline : the reference's CPU path on the host cores (bounded sample), forward and train step
  --impl reference : the same, as the reference arm: the UNMODIFIED reference modules staged under
                 baseline/_ref/src by __graft_entry__.build() (kind "reference"); the oracle port
                 (kind "port") only where they are absent

Rank 0 prints ONE JSON line.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "conv-tasnet_b200")
REF_SRC = os.path.join(ROOT, "baseline", "_ref", "src")

import torch  # noqa: E402

PAPER = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
SR = 8000
T_4S = 4 * SR
GLOBAL_BATCH = 512
TRAIN_BATCH = 3
WEIGHT_SEED = 1234      # torch.manual_seed(WEIGHT_SEED); ConvTasNet(**PAPER): same weights from the reference
                        # class and from the drop-in (tests/test_host_cpu.py: same-seed state_dict bit-equal)
METRIC = "separated audio-sec/s (N256 L20 B256 H512 P3 X8 R4 gLN, 4 s @ 8 kHz); train step ms"
UNIT = "audio-s/s"
CONFIG = {"workload": "bulk separation of %d x 4 s @ 8 kHz mixtures, paper config (BASELINE configs[3]), "
                      "sharded by utterance; train_step = BASELINE configs[1], 3 x 4 s per GPU" % GLOBAL_BATCH,
          "global_batch": GLOBAL_BATCH, "seq_len": T_4S,
          "l2": "inputs larger than L2: the activation working set of a step (~33 MB per utterance, 17 GB per 512) "
                ">> 126 MB"}


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


# the files the forward / backward kernels of the timed path live in
HOT_SOURCES = ("backward.cu", "bwd_dw_march.cuh", "common.cuh", "decoder.cu", "dwconv.cu", "dwconv_march.cuh",
               "encoder.cu", "f32x2.cuh", "fused_block.cu", "fused_c1.cu", "gemm.cu", "ptx.cuh", "wgrad.cu")


def kernel_source_hash():
    """sha256 over the CUDA sources: ncu-derived numbers (roofline.traffic) are only valid for the kernels they
    were captured from, so the committed json carries this hash and a mismatch reports traffic as null."""
    h = hashlib.sha256()
    d = os.path.join(PKG, "csrc")
    for name in HOT_SOURCES:
        with open(os.path.join(d, name), "rb") as f:
            h.update(name.encode() + b"\0" + f.read())
    return h.hexdigest()[:16]


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


# ------------------------------------------------------------------ CPU arm (reference on the host cores)
class _CpuArm:
    """The reference's CPU implementation of the path.  kind "reference": the unmodified modules
    (conv_tasnet.py, pit_criterion.py, utils.py as staged under baseline/_ref/src) with the harness-side shim
    for the hard-coded `.cuda()` of utils.py:40; kind "port": the oracle's restatement."""

    def __init__(self, cfg):
        self.cfg = cfg
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        have = all(os.path.exists(os.path.join(REF_SRC, f)) for f in ("conv_tasnet.py", "pit_criterion.py", "utils.py"))
        if have:
            self.kind = "reference"
            sys.path.insert(0, REF_SRC)
            torch.Tensor.cuda = lambda self, *a, **k: self      # utils.py:40: keep the index table on the CPU
            import conv_tasnet as ref_ct
            import pit_criterion as ref_pit
            assert os.path.dirname(os.path.abspath(ref_ct.__file__)) == REF_SRC
            torch.manual_seed(WEIGHT_SEED)
            self.model = ref_ct.ConvTasNet(*(cfg[k] for k in "NLBHPXRC"), norm_type=cfg["norm_type"],
                                           causal=cfg["causal"], mask_nonlinear=cfg["mask_nonlinear"])
            self.cal_loss = ref_pit.cal_loss
        else:
            self.kind = "port"
            sys.path.insert(0, ROOT)
            from oracle import convtasnet_oracle as O
            self.O = O
            self.params = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=WEIGHT_SEED)

    def forward_throughput(self, n_utt, steps, warmup, T=T_4S):
        g = torch.Generator().manual_seed(2025)
        mix = 0.1 * torch.randn(n_utt, T, generator=g)
        if self.kind == "reference":
            self.model.eval()
            fn = lambda: self.model(mix)                       # noqa: E731
        else:
            fn = lambda: self.O.forward(mix, self.params, self.cfg)  # noqa: E731
        with torch.no_grad():
            for _ in range(warmup):
                fn()
            t0 = time.perf_counter()
            for _ in range(steps):
                fn()
            dt = time.perf_counter() - t0
        return n_utt * steps * (T / SR) / dt, dt / steps

    def train_step_seconds(self, steps, warmup, M=TRAIN_BATCH, T=T_4S):
        """solver.py:178-186: forward, cal_loss, zero_grad, backward, clip_grad_norm_(5), Adam(1e-3) step."""
        g = torch.Generator().manual_seed(77)
        src = 0.1 * torch.randn(M, self.cfg["C"], T, generator=g)
        mix = src.sum(1)
        lengths = torch.full((M,), T, dtype=torch.int64)
        if self.kind == "reference":
            model = self.model.train()
            ps = list(model.parameters())
            opt = torch.optim.Adam(ps, lr=1e-3)

            def step():
                est = model(mix)
                loss, *_ = self.cal_loss(src, est, lengths)
                opt.zero_grad()
                loss.backward()
                torch.nn.utils.clip_grad_norm_(ps, 5.0)
                opt.step()
        else:
            p = {k: v.clone().requires_grad_(v.is_floating_point()) for k, v in self.params.items()}
            ps = [v for v in p.values() if v.requires_grad]
            opt = torch.optim.Adam(ps, lr=1e-3)

            def step():
                est = self.O.forward(mix, p, self.cfg)
                loss = self.O.cal_loss(src, est, lengths)[0]
                opt.zero_grad()
                loss.backward()
                torch.nn.utils.clip_grad_norm_(ps, 5.0)
                opt.step()
        for _ in range(warmup):
            step()
        t0 = time.perf_counter()
        for _ in range(steps):
            step()
        return (time.perf_counter() - t0) / steps


def run_reference(args):
    """The reference arm: the reference's own CPU implementation of the path on the box's host cores, all of
    them, on the b200 arm's metric / unit / config; each step separates a bounded SAMPLE of the 512-mixture
    workload (--sample utterances) so that --steps K --warmup W ends within minutes.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    arm = _CpuArm(PAPER)
    n_utt = max(1, args.sample)
    steps, warm = max(1, args.steps), max(0, args.warmup)
    v, step_s = arm.forward_throughput(n_utt, steps, warm)
    what = ("unmodified reference modules (baseline/_ref/src)" if arm.kind == "reference" else "oracle port") + \
        ", torch CPU fp32, %d threads" % arm.cores
    sample = "%d x 4 s mixtures per step (bounded sample of the %d-mixture workload), %d steps after %d warm-up; %s" % (
        n_utt, GLOBAL_BATCH, steps, warm, what)
    out = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": dict(CONFIG),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if not args.no_train:
        t_steps, t_warm = max(1, min(steps, 20)), max(1, min(warm, 3))
        ms = arm.train_step_seconds(t_steps, t_warm) * 1e3
        out["train_step"] = {
            "ms_per_step": ms, "steps": t_steps, "warmup": t_warm, "utterances": TRAIN_BATCH,
            "audio_s_per_s": TRAIN_BATCH * (T_4S / SR) / (ms / 1e3),
            "cpu_baseline": {"value": ms, "unit": "ms/step", "cores": arm.cores, "kind": arm.kind,
                             "sample": "BASELINE configs[1] whole: %d x 4 s, fwd + cal_loss + bwd + clip(5) + Adam "
                                       "(solver.py:178-186), %d steps after %d warm-up; %s" % (TRAIN_BATCH, t_steps, t_warm, what)}}
    print(json.dumps(out), flush=True)


def cpu_baseline_subprocess(sample, steps, warmup, timeout=600):
    """The cpu_baseline leg of the b200 arm: the reference arm in a child process (the reference's module names
    -- conv_tasnet, utils, pit_criterion -- are the drop-in's, so the two cannot share an interpreter)."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", str(steps), "--warmup",
           str(warmup), "--sample", str(sample)]
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
    try:
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=timeout, env=env)
        line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
        return json.loads(line)
    except Exception as e:      # reported, never silently dropped
        return {"error": "%s: %s" % (type(e).__name__, e)}


# ------------------------------------------------------------------ GPU arm
def _roofs(acc, pk, traffic, names, how):
    """acc: kind -> [flops, bytes, ms, launches] summed over the timed repeats.  One roofline entry per kernel
    kind: tensor-bound kinds against the sustained bf16 peak (algorithmic FLOPs = ONE product; the hi/lo split
    executes three MMAs per product, reported as executed_*), HBM-bound kinds against the measured copy bandwidth."""
    tot_ms = sum(a[2] for a in acc.values())
    roofs = {}
    for k, a in acc.items():
        n = a[3]
        if a[2] <= 0:
            continue
        tensor = "gemm" in k or "wgrad" in k
        # c1::updc1_kernel moves 5 KB per frame (x, y in; x, a out) for one 256 x 512 product: its limiter is the
        # memory system (profiles/r2), so it is held against the HBM roofline; its tensor figures ride along
        if k.startswith("gemm_update") or k.startswith("fwd_gemm_update"):
            ach = a[1] / a[2] / 1e6
            fl = a[0] / a[2] / 1e9
            r = {"bound": "hbm", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"],
                 "peak_source": pk["source"] + " (copy bandwidth)", "algorithmic_tflops": fl,
                 "executed_tflops": 3.0 * fl, "executed_tensor_frac": 3.0 * fl / pk["tf_sustained"]}
        elif tensor and "encoder" not in k:
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
        r.update({"kernel": names.get(k, k), "avg_launch_ms": a[2] / n, "launches_per_pass": n,
                  "algorithmic_bytes_per_launch": a[1] / n, "share_of_step": a[2] / tot_ms, "traffic": traffic.get(k)})
        roofs[k] = r
    dom = max(roofs, key=lambda k: roofs[k]["share_of_step"])
    roof = dict(roofs[dom])
    roof["how"] = how
    roof["all_kernels"] = roofs
    return roof


def _traffic(section, mb):
    """DRAM bytes per launch from the committed ncu --set full capture of THESE kernel sources at THIS
    micro-batch (profiles/r2/ncu_traffic.json); anything else reports null rather than a stale number."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2", "ncu_traffic.json")) as f:
            tj = json.load(f)
        sec = tj[section]
        if tj.get("kernel_source_hash") == kernel_source_hash() and sec.get("batch") == mb:
            return {k: v["traffic_bytes"] for k, v in sec["kernels"].items()}
    except Exception:
        pass
    return {}


FWD_NAMES = {"gemm_conv1x1": "gemm_tc_kernel<256,pair,plain> (conv1x1 + PReLU statistics)",
             "gemm_pointwise": "gemm_tc_kernel<256,pair,affine> (pointwise + folded gLN + residual)",
             "gemm_dwpw_fused": "fb::dwpw_kernel (depthwise producers + pointwise GEMM)",
             "gemm_update_conv1x1": "c1::updc1_kernel (residual-update producers + conv1x1 GEMM + PReLU statistics)",
             "gemm_update_mask": "c1::updc1_kernel (residual update + mask conv)",
             "gemm_bottleneck": "gemm_tc_kernel (bottleneck)", "gemm_mask": "gemm_tc_kernel (mask conv)",
             "dwconv": "dwconv_march_kernel", "resid_update": "fb::resid_update_kernel",
             "encoder": "encoder_fwd_kernel", "decoder": "decoder_fwd_blocked_kernel"}


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
    sys.path.insert(0, PKG)
    import conv_tasnet

    cfg = PAPER
    T = T_4S
    n_local = GLOBAL_BATCH // world
    torch.manual_seed(WEIGHT_SEED)          # the drop-in's own seeded constructor: the reference arm's weights
    model = conv_tasnet.ConvTasNet(**cfg).to(dev).eval()
    eng = model.engine()
    eng.micro_batch = args.micro_batch
    eng.fuse_max_dilation = args.fuse
    eng.fuse_update = bool(args.fuse_update)
    eng.use_graph = bool(args.graph)
    eng.n_streams = args.streams

    g = torch.Generator().manual_seed(2025 + rank)
    host_mix = (0.1 * torch.randn(n_local, T, generator=g)).pin_memory()
    host_est = [torch.empty(n_local, cfg["C"], T).pin_memory() for _ in range(2)]
    mix = host_mix.to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps, finish=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        if finish is not None:
            finish()
        e1.record()
        torch.cuda.synchronize(dev)
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        barrier()
        return float(ms.item()) / 1e3

    def step_resident(i):
        with torch.no_grad():
            model(mix)

    # the public host-to-host call: pinned mixtures in, pinned estimates out EVERY step.  Steps are issued without
    # waiting for each other (ConvTasNet.separate(wait=False): two output buffers in turn, a step's buffer is
    # reused only after its copy-out event), so one step's copy-out runs under the next step's kernels; the
    # timed region ends when the last estimate has arrived on the host.
    pending = [None, None]

    def step_e2e(i):
        b = i % 2
        if pending[b] is not None:
            pending[b].synchronize()
        _, pending[b] = model.separate(host_mix, out=host_est[b], wait=False)

    def finish_e2e():
        cur = torch.cuda.current_stream(dev)
        for ev in pending:
            if ev is not None:
                cur.wait_event(ev)      # the closing event e1 is recorded behind the last copy-out

    for _ in range(args.warmup):
        step_resident(0)
    sampler = ClockSampler(physical_gpu_index(local))
    sampler.start()
    l0 = eng.launches
    sec = timed(step_resident, args.steps)
    launches = eng.launches - l0
    clocks = sampler.finish()
    audio_s = GLOBAL_BATCH * (T / SR) * args.steps
    value = audio_s / sec

    for i in range(max(2, args.warmup // 2)):
        step_e2e(i)
    finish_e2e()
    sec_e2e = timed(step_e2e, args.steps, finish_e2e)
    e2e = {"value": audio_s / sec_e2e, "unit": UNIT, "h2d_bytes_per_step": GLOBAL_BATCH * T * 4,
           "d2h_bytes_per_step": GLOBAL_BATCH * cfg["C"] * T * 4, "ms_per_step": sec_e2e / args.steps * 1e3,
           "api": "ConvTasNet.separate(host_mix, out=pinned, wait=False), every step copies its inputs in and its "
                  "estimates out; consecutive steps pipeline"}

    # BASELINE configs[0]: ONE 4 s mixture host-to-host, blocking (the separate.py:60-68 pattern)
    one_in = host_mix[:1].clone().pin_memory()
    one_out = torch.empty(1, cfg["C"], T).pin_memory()
    for _ in range(5):
        model.separate(one_in, out=one_out)
    t0 = time.perf_counter()
    n_lat = 50
    for _ in range(n_lat):
        model.separate(one_in, out=one_out)
    single = {"ms": (time.perf_counter() - t0) / n_lat * 1e3, "what": "BASELINE configs[0]: one 4 s mixture, pinned host "
              "in -> pinned host out, blocking call (wall clock, %d calls)" % n_lat}
    single["x_realtime"] = (T / SR) / (single["ms"] / 1e3)

    # per-kernel times of one micro-batch pass, CUDA events around every launch (rank 0)
    roof, breakdown = None, None
    pk = peaks()
    if rank == 0:
        mb = min(args.micro_batch, n_local)
        plan = eng._plan({k: p.detach() for k, p in model._params().items()}, mb, T, dev)
        plan.mix.copy_(mix[:mb])
        plan.launch()
        torch.cuda.synchronize(dev)
        acc = {}
        reps = 3
        for _ in range(reps):
            for kind, fl, by, ms in plan.launch_timed():
                a = acc.setdefault(kind, [0.0, 0.0, 0.0, 0])
                a[0] += fl
                a[1] += by
                a[2] += ms
                a[3] += 1
        tot_ms = sum(a[2] for a in acc.values())
        breakdown = {k: {"launches": a[3] // reps, "ms": a[2] / reps, "share": a[2] / tot_ms,
                         "tflops": a[0] / a[2] / 1e9, "gbs": a[1] / a[2] / 1e6} for k, a in acc.items()}
        for a in acc.values():          # per pass of one micro-batch
            a[0], a[1], a[2], a[3] = a[0] / reps, a[1] / reps, a[2] / reps, a[3] // reps
        roof = _roofs(acc, pk, _traffic("forward", mb), FWD_NAMES,
                      "dominant kernel by share of the step; CUDA events around every launch of one %d-utterance "
                      "micro-batch pass, %d repeats; traffic from profiles/r2/ncu_traffic.json when it was captured "
                      "from these kernel sources" % (mb, reps))

    train = None
    if not args.no_train and _precision() == "bf16x3":     # the fp16 inference precision has no training path
        train = train_step(args, world, rank, dev, timed, pk)
        train["ms_per_step_torch_optimizer"] = train_step(args, world, rank, dev, timed, pk, fused_optim=False)["ms_per_step"]

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        ref = cpu_baseline_subprocess(sample=8, steps=3, warmup=1)
        if "cpu_baseline" in ref:
            cpu = ref["cpu_baseline"]
            if train is not None and "train_step" in ref:
                train["cpu_baseline"] = ref["train_step"]["cpu_baseline"]
        else:
            cpu = ref

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None,
            "dtype": ("bf16 hi+lo split operands (3 MMAs), fp32 accumulate; fp32 elsewhere" if _precision() == "bf16x3"
                      else "fp16 operands (1 MMA), fp32 accumulate; fp32 elsewhere (CTN_PRECISION=fp16, opt-in)"),
            "data": "synthetic", "config": dict(CONFIG),
            "run": {"per_gpu": n_local, "micro_batch": args.micro_batch, "cuda_graph": bool(args.graph),
                    "streams": args.streams, "parallelism": "utterance-sharded x%d, no collective" % world,
                    "x_realtime_per_gpu": value / world, "kernel_source_hash": kernel_source_hash()},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "algorithmic_tflops": forward_flops(cfg, T) * GLOBAL_BATCH * args.steps / sec / 1e12,
            "roofline": roof, "kernel_breakdown": breakdown, "single_utterance": single,
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


TRAIN_NAMES = {"fwd_gemm_conv1x1": "gemm_tc_kernel (conv1x1 + PReLU statistics)",
               "fwd_gemm_pointwise": "gemm_tc_kernel (pointwise + folded gLN + residual)",
               "fwd_dwconv": "dwconv_march_kernel",
               "fwd_gemm_dwpw_fused": "fb::dwpw_kernel<false, true> (depthwise producers + pointwise GEMM, saves d)",
               "fwd_gemm_update_conv1x1": "c1::updc1_kernel (out-of-place residual update + conv1x1, a before the PReLU)",
               "fwd_gemm_update_mask": "c1::updc1_kernel (residual update + mask conv)", "bwd_wgrad": "wgrad_tc_kernel (MN-major tcgen05 weight gradient)",
               "bwd_gemm_dgrad_pointwise": "gemm_tc_kernel (dgrad of pointwise + norm-backward sums)",
               "bwd_gemm_dgrad_conv1x1": "gemm_tc_kernel (dgrad of conv1x1 + residual)",
               "bwd_dwconv": "bwd_dw_march_kernel", "bwd_prelu_norm": "bwd_prelu_norm_kernel",
               "bwd_norm_split": "norm_split_kernel", "bwd_reduce_partials": "reduce_partials_kernel"}


def train_step(args, world, rank, dev, timed, pk, fused_optim=True):
    """BASELINE configs[1]: paper config, batch 3 x 4 s per GPU (weak scaling), forward +
    cal_loss + backward (ONE NCCL all-reduce of the flat gradient when world > 1) + clip(5) +
    Adam(1e-3), as solver.py:178-186.  Host mixtures/sources are copied in every step and the
    loss is read back (solver.py:188 does loss.item())."""
    import conv_tasnet
    import pit_criterion
    from ctn_b200 import _dist
    cfg, T, M = PAPER, T_4S, TRAIN_BATCH
    torch.manual_seed(WEIGHT_SEED)
    model = conv_tasnet.ConvTasNet(**cfg).to(dev).train()
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
    losses = []

    def step(i=0):
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
        torch.cuda.synchronize(dev)
        losses.append(float(h_loss))
    l0 = eng.launches
    steps = max(args.steps, 5)
    sec = timed(step, steps)
    out = {"ms_per_step": sec / steps * 1e3, "steps": steps, "utterances_per_gpu": M, "global_batch": M * world,
           "scaling": "weak", "audio_s_per_s": M * world * (T / SR) * steps / sec,
           "algorithmic_tflops": 3.0 * forward_flops(cfg, T) * M * world * steps / sec / 1e12,
           "tensor_frac_of_sustained": 3.0 * forward_flops(cfg, T) * M * steps / sec / 1e12 / pk["tf_sustained"],
           # every step re-prepares the 1x1 weights (2 launches) because the optimizer changed them
           "gpu_launches_per_step": (eng.launches - l0) // steps + (4 if fused_optim else 0),
           "loss": float(h_loss), "loss_first_steps": losses[:3],
           "optimizer": "ctn_b200.optim.FusedAdam (clip + Adam, 2 launches)" if fused_optim
                        else "torch clip_grad_norm_ + torch.optim.Adam",
           "config": "BASELINE configs[1]: paper config, 3 x 4 s per GPU, fwd + PIT SI-SNR loss + bwd + "
                     "grad all-reduce + clip(5) + Adam, host inputs copied every step",
           "h2d_bytes_per_step": M * T * 4 * (1 + cfg["C"]), "d2h_bytes_per_step": 4}
    if fused_optim and rank == 0:
        plan = next(iter(eng.plans.values()))
        plan.launch_timed()
        acc = {}
        reps = 3
        for _ in range(reps):
            for kind, fl, by, ms in plan.launch_timed():
                a = acc.setdefault(kind, [0.0, 0.0, 0.0, 0])
                a[0] += fl
                a[1] += by
                a[2] += ms
                a[3] += 1
        for a in acc.values():
            a[3] //= reps
            a[0], a[1], a[2] = a[0] / reps, a[1] / reps, a[2] / reps
        tot = sum(a[2] for a in acc.values())
        out["kernel_breakdown"] = {k: {"launches": a[3], "ms": a[2], "share": a[2] / tot,
                                       "tflops": a[0] / a[2] / 1e9 if a[2] > 0 else 0.0,
                                       "gbs": a[1] / a[2] / 1e6 if a[2] > 0 else 0.0} for k, a in acc.items()}
        out["serial_kernel_ms"] = tot
        out["roofline"] = _roofs(acc, pk, _traffic("train", M), TRAIN_NAMES,
                                 "dominant kernel kind of one forward + backward at batch %d; CUDA events around every "
                                 "launch, all on one stream (the timed step overlaps the weight-gradient GEMMs on a "
                                 "side stream), %d repeats" % (M, reps))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--micro-batch", type=int, default=128)
    ap.add_argument("--fuse", type=int, default=1 << 30,
                    help="blocks with dilation <= this use the fused depthwise+pointwise kernel (0 = off)")
    ap.add_argument("--fuse-update", type=int, default=1,
                    help="fold the deferred residual update into the next conv1x1 (ctn_updc1_fwd)")
    ap.add_argument("--graph", type=int, default=1)
    ap.add_argument("--streams", type=int, default=1,
                    help="micro-batches in flight on separate streams (engine.n_streams)")
    ap.add_argument("--sample", type=int, default=16,
                    help="reference arm: 4 s mixtures separated per step (bounded sample of the 512-mixture workload)")
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
