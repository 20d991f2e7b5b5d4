"""CTN_PRECISION=fp16: the opt-in inference precision (fp16 GEMM operands, one MMA per product; the same sources
compiled with -DCTN_OPERAND_FP16=1 into libctn_b200_fp16.so).  The precision is a per-process choice, so the
checks run in a child process.  Bars: waveform within the path's 1e-3 rel-L2 tolerance of the fp64 oracle at the
paper configuration (measured 4.9e-4, predicted 5.6e-4 by tools/precision_study.py), a chunked causal stream equal to
the whole forward, training refused."""
import os
import subprocess
import sys

import pytest

from conftest import PKG, ROOT

pytestmark = pytest.mark.gpu

CHILD = r"""
import sys, torch
sys.path[:0] = [%(root)r, %(pkg)r]
import conv_tasnet, pit_criterion
from ctn_b200 import _native as nat
from oracle import convtasnet_oracle as O
assert nat.PRECISION == "fp16" and nat.lib().ctn_operand_format() == 101

def err(cfg, M, T, seed):
    p = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=seed)
    m = conv_tasnet.ConvTasNet(**cfg); m.load_state_dict(p); m = m.cuda().eval()
    g = torch.Generator().manual_seed(2025)
    mix = 0.1 * torch.randn(M, T, generator=g)
    with torch.no_grad():
        est = m(mix.cuda())
    ref = O.forward(mix.double(), {k: v.double() for k, v in p.items()}, cfg)
    return float((est.double().cpu() - ref).norm() / ref.norm()), m, mix

paper = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
e, m, mix = err(paper, 2, 32000, 1234)
print("paper rel-L2 %%.3e" %% e)
assert 5e-5 < e < 1e-3, e          # inside the tolerance, and visibly not the bf16 x 3 path (1.5e-5)

small = dict(N=64, L=16, B=64, H=128, P=3, X=3, R=2, C=2, norm_type="cLN", causal=True, mask_nonlinear="relu")
e2, ms, mix2 = err(small, 1, 8000, 7)
print("small causal rel-L2 %%.3e" %% e2)
assert e2 < 1e-3, e2
sep = ms.streaming(chunk_frames=32)
with torch.no_grad():
    whole = ms(mix2.cuda())[0]
    x = mix2[0].cuda()
    parts = [sep.feed(x[i:i + 1000]) for i in range(0, 8000, 1000)] + [sep.flush()]
got = torch.cat(parts, 1)
d = float((got.double() - whole.double()).norm() / whole.double().norm())
print("stream vs whole rel-L2 %%.3e%%s" %% (d, " (bit-identical)" if torch.equal(got, whole) else ""))
assert got.shape == whole.shape and d < 1e-5, d

m.train()
try:
    m(mix.cuda())
except RuntimeError as ex:
    assert "inference-only" in str(ex), ex
else:
    raise AssertionError("training must be refused in fp16 precision")
print("fp16 precision ok")
"""


def test_fp16_inference_precision_in_a_child_process():
    env = dict(os.environ, CTN_PRECISION="fp16")
    r = subprocess.run([sys.executable, "-c", CHILD % dict(root=ROOT, pkg=PKG)], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0 and "fp16 precision ok" in r.stdout, (r.stdout[-800:], r.stderr[-1500:])
