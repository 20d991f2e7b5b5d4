"""Small fixed workload for ncu: a few forward passes of one micro-batch at the paper
config (no CUDA graph, so every kernel is a plain launch).  Usage:
    python tools/profile_forward.py [M] [passes]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200")]
import torch  # noqa: E402
import conv_tasnet  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 16
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
torch.manual_seed(0)
model = conv_tasnet.ConvTasNet(**cfg).cuda().eval()
model.engine().micro_batch = M
model.engine().fuse_update = os.environ.get("CTN_FUSE_UPDATE", "1") == "1"
mix = 0.1 * torch.randn(M, 32000, device="cuda")
with torch.no_grad():
    for _ in range(passes):
        est = model(mix)
torch.cuda.synchronize()
print("ok", tuple(est.shape), float(est.abs().mean()))
