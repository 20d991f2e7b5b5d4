"""A few training steps for ncu (paper config): python tools/profile_train_once.py [M] [steps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200")]
import torch
import conv_tasnet, pit_criterion
M = int(sys.argv[1]) if len(sys.argv) > 1 else 3
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
if os.environ.get("CTN_CFG") == "cfg5":   # BASELINE configs[4]
    cfg = dict(N=512, L=16, B=128, H=512, P=3, X=8, R=3, C=3, norm_type="gLN", causal=False, mask_nonlinear="softmax")
torch.manual_seed(0)
model = conv_tasnet.ConvTasNet(**cfg).cuda().train()
model.train_engine().use_graph = False      # plain launches: every kernel is its own ncu record
src = 0.1 * torch.randn(M, cfg["C"], 32000, device="cuda"); mix = src.sum(1)
lengths = torch.full((M,), 32000, device="cuda")
for _ in range(steps):
    model.zero_grad()
    est = model(mix)
    loss, *_ = pit_criterion.cal_loss(src, est, lengths)
    loss.backward()
torch.cuda.synchronize()
print("ok", float(loss))
