"""What would a cheaper operand format cost in accuracy?  (planning aid, CPU only)

The tensor-core GEMMs form every 1x1 convolution as three bf16 MMAs on (hi, lo) split operands
(a_hi w_hi + a_lo w_hi + a_hi w_lo).  This script emulates that and cheaper alternatives on the oracle's
fp64 forward by rounding the operands of every 1x1 convolution (accumulation stays exact, as fp32
accumulation of 256..512 terms is not the limiting error), and reports the waveform error against the
unrounded fp64 forward at the paper configuration.  Tolerance of the path: 1e-3 relative L2.

usage: python tools/precision_study.py [T=32000] [seed ...]
"""
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import convtasnet_oracle as O  # noqa: E402


def rnd(x, dtype):
    return x.to(dtype).to(torch.float64)


def tf32(x):
    # round-to-nearest-even to a 10-bit mantissa
    i = x.to(torch.float32).view(torch.int32)
    i = (i + 0x0FFF + ((i >> 13) & 1)) & ~0x1FFF
    return i.view(torch.float32).to(torch.float64)


def split(x, dtype):
    hi = rnd(x, dtype)
    return hi, rnd(x - hi, dtype)


SCHEMES = {
    # name: (MMAs per product at bf16-equivalent cost, function(a, w) -> list of (a_part, w_part))
    "bf16 x3 (shipped)": (3, lambda a, w: [(sum(split(a, torch.bfloat16)), split(w, torch.bfloat16)[0]),
                                           (split(a, torch.bfloat16)[0], split(w, torch.bfloat16)[1])]),
    "bf16 x2, act split": (2, lambda a, w: [(sum(split(a, torch.bfloat16)), rnd(w, torch.bfloat16))]),
    "bf16 x2, weight split": (2, lambda a, w: [(rnd(a, torch.bfloat16), sum(split(w, torch.bfloat16)))]),
    "bf16 x1": (1, lambda a, w: [(rnd(a, torch.bfloat16), rnd(w, torch.bfloat16))]),
    "fp16 x3": (3, lambda a, w: [(sum(split(a, torch.float16)), split(w, torch.float16)[0]),
                                 (split(a, torch.float16)[0], split(w, torch.float16)[1])]),
    "fp16 x2, act split": (2, lambda a, w: [(sum(split(a, torch.float16)), rnd(w, torch.float16))]),
    "fp16 x2, weight split": (2, lambda a, w: [(rnd(a, torch.float16), sum(split(w, torch.float16)))]),
    "fp16 x1": (1, lambda a, w: [(rnd(a, torch.float16), rnd(w, torch.float16))]),
    # corrections in fp8 (e4m3): a_lo, w_lo pre-scaled by 2^9 (they are ~2^-9 of the value) and the scale
    # folded back, which on the device needs a second accumulator or block-scaled MMAs; 2 fp8 MMAs = 1 bf16 MMA
    "bf16 hi.hi + e4m3 corrections": (2, lambda a, w: [
        (rnd(a, torch.bfloat16), rnd(w, torch.bfloat16)),
        (rnd((a - rnd(a, torch.bfloat16)) * 512.0, torch.float8_e4m3fn) / 512.0, rnd(rnd(w, torch.bfloat16), torch.float8_e4m3fn)),
        (rnd(rnd(a, torch.bfloat16), torch.float8_e4m3fn), rnd((w - rnd(w, torch.bfloat16)) * 512.0, torch.float8_e4m3fn) / 512.0)]),
    "tf32 x1 (= 2 bf16 MMAs of tensor time)": (2, lambda a, w: [(tf32(a), tf32(w))]),
}


def run(scheme, mixture, params, cfg):
    real = F.conv1d

    def conv1d(x, w, *args, **kw):
        if w.shape[-1] == 1 and not args and not kw and scheme is not None:
            return sum(real(a, ww) for a, ww in scheme(x, w))
        return real(x, w, *args, **kw)

    O.F.conv1d = conv1d
    try:
        return O.forward(mixture, params, cfg)
    finally:
        O.F.conv1d = real


def main():
    T = int(sys.argv[1]) if len(sys.argv) > 1 else 32000
    seeds = [int(s) for s in sys.argv[2:]] or [1234]
    cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
    torch.set_grad_enabled(False)
    print("| operand format | MMAs (bf16-equivalent) | waveform rel-L2 vs fp64 |")
    print("|---|---|---|")
    rows = {}
    for seed in seeds:
        params = {k: v.double() for k, v in O.random_params(*(cfg[k] for k in "NLBHPXRC"), False, seed=seed).items()}
        g = torch.Generator().manual_seed(2025 + seed)
        mixture = (0.1 * torch.randn(1, T, generator=g)).double()
        exact = run(None, mixture, params, cfg)
        for name, (cost, fn) in SCHEMES.items():
            est = run(fn, mixture, params, cfg)
            rows.setdefault(name, (cost, []))[1].append(float((est - exact).norm() / exact.norm()))
    for name, (cost, errs) in rows.items():
        print("| %s | %d | %s |" % (name, cost, ", ".join("%.1e" % e for e in errs)))


if __name__ == "__main__":
    main()
