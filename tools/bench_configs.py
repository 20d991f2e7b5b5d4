"""Times the BASELINE.json configurations that bench.py's headline line does not cover, on one GPU:

  cfg1  paper config, ONE 4 s mixture, forward (separate.py path): latency
  cfg3  causal cLN paper config on ONE 60 s stream: latency
  cfg5  N512 L16 B128 H512 X8 R3 C3 softmax, training step at batch 8 x 4 s

One JSON line per configuration.  CUDA events, warm-up, host inputs copied in / result read back
inside the timed region (same conventions as bench.py's e2e leg).
usage: python tools/bench_configs.py [cfg1] [cfg3] [cfg5]
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "conv-tasnet_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

SR = 8000


def fwd_flops(c, T):
    hop = c["L"] // 2
    K = (T - c["L"]) // hop + 1
    return 2.0 * K * (c["N"] * c["L"] + c["N"] * c["B"] + c["R"] * c["X"] * (2 * c["B"] * c["H"] + c["H"] * c["P"])
                      + c["B"] * c["C"] * c["N"] + c["C"] * c["N"] * c["L"])


def timed(fn, steps, warm):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def build(cfg, train=False):
    import conv_tasnet
    from oracle import convtasnet_oracle as O
    params = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=1234)
    model = conv_tasnet.ConvTasNet(**cfg)
    model.load_state_dict(params)
    model = model.cuda()
    return model.train() if train else model.eval()


def latency(name, cfg, T, M=1, steps=20, warm=5, graph=True):
    model = build(cfg)
    eng = model.engine()
    eng.use_graph = graph
    eng.micro_batch = max(1, M)
    g = torch.Generator().manual_seed(2025)
    h_mix = (0.1 * torch.randn(M, T, generator=g)).pin_memory()
    h_est = torch.empty(M, cfg["C"], T).pin_memory()

    def step():
        with torch.no_grad():
            est = model(h_mix.cuda(non_blocking=True))
            h_est.copy_(est, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    ms = timed(step, steps, warm)
    audio = M * T / SR
    return {"config": name, "ms": ms, "audio_s": audio, "x_realtime": audio / (ms / 1e3),
            "algorithmic_tflops": fwd_flops(cfg, T) * M / (ms / 1e3) / 1e12, "M": M, "T": T, "cuda_graph": graph,
            "launches": eng.num_launches(), "timed": "host mixture -> device, forward, estimate -> host, %d steps" % steps}


def train(name, cfg, M, T, steps=10, warm=4):
    import pit_criterion
    from ctn_b200.optim import FusedAdam
    model = build(cfg, train=True)
    opt = FusedAdam(model.parameters(), lr=1e-3, max_norm=5.0)
    g = torch.Generator().manual_seed(77)
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

    ms = timed(step, steps, warm)
    return {"config": name, "ms_per_step": ms, "batch": M, "T": T, "audio_s_per_s": M * T / SR / (ms / 1e3),
            "algorithmic_tflops": 3.0 * fwd_flops(cfg, T) * M / (ms / 1e3) / 1e12,
            "loss_first": losses[0], "loss_last": losses[-1],
            "timed": "host inputs -> device, fwd + PIT SI-SNR loss + bwd + clip(5) + Adam, loss -> host"}


def stream_latency(name, cfg, chunk, seconds=60, warm_chunks=8):
    """Chunked causal separation (model.streaming): per-chunk latency = host chunk -> device, one
    steady-state chunk (CUDA graph), separated chunk -> host."""
    model = build(cfg)
    hop = cfg["L"] // 2
    n = chunk * hop
    sep = model.streaming(chunk_frames=chunk)
    g = torch.Generator().manual_seed(2025)
    total = seconds * SR // n
    h_in = (0.1 * torch.randn(total + warm_chunks + 2, n, generator=g)).pin_memory()
    h_out = torch.empty(cfg["C"], n).pin_memory()
    state = {"i": 0}
    sep.feed(h_in[0, :hop].cuda())      # the one extra hop the first chunk needs

    def step():
        out = sep.feed(h_in[state["i"]].cuda(non_blocking=True))
        h_out.copy_(out, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        state["i"] += 1

    ms = timed(step, total, warm_chunks)
    audio = n / SR
    return {"config": name, "chunk_frames": chunk, "chunk_audio_ms": audio * 1e3, "ms_per_chunk": ms,
            "algorithmic_latency_ms": (n + hop) / SR * 1e3, "x_realtime_one_stream": audio / (ms / 1e3),
            "launches_per_chunk": len(sep.body_steady) + 3, "chunks_timed": total,
            "timed": "host chunk -> device, steady-state chunk (CUDA graph), separated chunk -> host"}


PAPER = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
CAUSAL = dict(PAPER, norm_type="cLN", causal=True)
BEST = dict(N=512, L=16, B=128, H=512, P=3, X=8, R=3, C=3, norm_type="gLN", causal=False, mask_nonlinear="softmax")


def main():
    which = sys.argv[1:] or ["cfg1", "cfg3", "cfg5"]
    if "cfg1" in which:
        print(json.dumps(latency("cfg1: paper gLN, 1 x 4 s forward", PAPER, 32000)), flush=True)
        print(json.dumps(latency("cfg1 (no CUDA graph)", PAPER, 32000, graph=False)), flush=True)
    if "cfg3" in which:
        print(json.dumps(latency("cfg3: causal cLN, 1 x 60 s stream", CAUSAL, 480000, steps=10, warm=3)), flush=True)
        for chunk in (16, 64, 256, 1024):
            print(json.dumps(stream_latency("cfg3 streaming: causal cLN, chunks of %d frames" % chunk, CAUSAL, chunk,
                                            seconds=20)), flush=True)
    if "cfg5" in which:
        print(json.dumps(train("cfg5: N512 L16 B128 H512 X8 R3 C3 softmax, train batch 8 x 4 s", BEST, 8, 32000)),
              flush=True)
        print(json.dumps(train("cfg2: paper gLN, train batch 3 x 4 s", PAPER, 3, 32000)), flush=True)


if __name__ == "__main__":
    main()
