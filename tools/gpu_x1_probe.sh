# Speed probe: the forward with ONE MMA per product (build -DCTN_MMA_TERMS=1) against the shipped three.
# The probe library replaces the product library ONLY in the GPU box's scratch copy of the repo.
set -x
export CTN_ALLOW_EXPERIMENT_BUILD=1   # the Python layer refuses non-product operand formats otherwise
mkdir -p gpurun_out
timeout 300 python bench.py --no-cpu --no-train > gpurun_out/bench_x3.json 2> gpurun_out/bench_x3.err
cp conv-tasnet_b200/libctn_b200.so /tmp/libctn_b200_x3.so
cp conv-tasnet_b200/libctn_b200_x1.so conv-tasnet_b200/libctn_b200.so
touch conv-tasnet_b200/libctn_b200.so
timeout 300 python bench.py --no-cpu --no-train > gpurun_out/bench_x1.json 2> gpurun_out/bench_x1.err
timeout 120 python - > gpurun_out/x1_error.log 2>&1 <<'PY'
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
print("one-term build, waveform rel-L2 vs fp64 oracle: %.3e" % float((est - ref).norm() / ref.norm()))
PY
cp /tmp/libctn_b200_x3.so conv-tasnet_b200/libctn_b200.so
cut -c1-260 gpurun_out/bench_x3.json; cut -c1-260 gpurun_out/bench_x1.json; cat gpurun_out/x1_error.log | tail -3; tail -3 gpurun_out/bench_x1.err
