"""Quick check of the opt-in inference precision on a GPU box (a few seconds):
    CTN_PRECISION=fp16 python tools/fp16_mode_check.py
small causal cLN model against the fp64 oracle, chunked streaming against the whole forward, training refused.
The full check is tests/test_gpu_zz_precision.py."""
import sys, torch
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200")]
import conv_tasnet
from ctn_b200 import _native as nat
from oracle import convtasnet_oracle as O
assert nat.PRECISION == "fp16" and nat.lib().ctn_operand_format() == 101
cfg = dict(N=64, L=16, B=64, H=128, P=3, X=3, R=2, C=2, norm_type="cLN", causal=True, mask_nonlinear="relu")
p = O.random_params(*(cfg[k] for k in "NLBHPXRC"), True, seed=7)
m = conv_tasnet.ConvTasNet(**cfg); m.load_state_dict(p); m = m.cuda().eval()
mix = 0.1 * torch.randn(1, 8000, generator=torch.Generator().manual_seed(2025))
with torch.no_grad():
    whole = m(mix.cuda())
ref = O.forward(mix.double(), {k: v.double() for k, v in p.items()}, cfg)
print("small causal rel-L2 %.3e" % float((whole.double().cpu() - ref).norm() / ref.norm()))
sep = m.streaming(chunk_frames=32)
with torch.no_grad():
    x = mix[0].cuda()
    got = torch.cat([sep.feed(x[i:i + 1000]) for i in range(0, 8000, 1000)] + [sep.flush()], 1)
print("stream equal:", torch.equal(got, whole[0]), float((got - whole[0]).abs().max()))
m.train()
try:
    m(mix.cuda()); print("TRAINING NOT REFUSED")
except RuntimeError as ex:
    print("refused:", str(ex)[:60])
