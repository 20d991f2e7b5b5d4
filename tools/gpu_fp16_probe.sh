# Experiment build -DCTN_OPERAND_FP16=1 (fp16 operands, one MMA per product): forward accuracy against the fp64
# oracle and bench throughput.  The swap happens only in the GPU box's scratch copy of the repo.
set -x
export CTN_ALLOW_EXPERIMENT_BUILD=1   # the Python layer refuses non-product operand formats otherwise
mkdir -p gpurun_out
cp conv-tasnet_b200/libctn_b200_fp16.so conv-tasnet_b200/libctn_b200.so
timeout 60 python - > gpurun_out/fp16_error.log 2>&1 <<'PY'
import sys, torch
sys.path[:0] = [".", "conv-tasnet_b200"]
import conv_tasnet
from oracle import convtasnet_oracle as O
cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
p = O.random_params(*(cfg[k] for k in "NLBHPXRC"), False, seed=1234)
m = conv_tasnet.ConvTasNet(**cfg); m.load_state_dict(p); m = m.cuda().eval()
g = torch.Generator().manual_seed(2025)
mix = 0.1 * torch.randn(1, 32000, generator=g)
with torch.no_grad():
    est = m(mix.cuda()).double().cpu()
    ref = O.forward(mix.double(), {k: v.double() for k, v in p.items()}, cfg)
print("fp16 one-term build, waveform rel-L2 vs fp64 oracle: %.3e" % float((est - ref).norm() / ref.norm()))
PY
cat gpurun_out/fp16_error.log | tail -2
timeout 100 python bench.py --no-cpu --no-train --steps 3 --warmup 3 > gpurun_out/bench_fp16.json 2> gpurun_out/bench_fp16.err
cut -c1-220 gpurun_out/bench_fp16.json
