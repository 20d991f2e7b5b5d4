"""Host-side profile (cProfile) of training steps at the paper configuration: python tools/profile_train_host.py [M]"""
import cProfile, os, pstats, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200")]
import torch
import conv_tasnet, pit_criterion
from ctn_b200.optim import FusedAdam
M = int(sys.argv[1]) if len(sys.argv) > 1 else 3
cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
torch.manual_seed(0)
model = conv_tasnet.ConvTasNet(**cfg).cuda().train()
opt = FusedAdam(model.parameters(), lr=1e-3, max_norm=5.0)
src = 0.1 * torch.randn(M, 2, 32000, device="cuda"); mix = src.sum(1)
lengths = torch.full((M,), 32000, device="cuda")
def step():
    est = model(mix)
    loss, *_ = pit_criterion.cal_loss(src, est, lengths)
    opt.zero_grad()
    loss.backward()
    opt.step()
for _ in range(5): step()
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(20): step()
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print("host issue per step %.2f ms, with final sync %.2f ms" % ((t1 - t0) / 20 * 1e3, (t2 - t0) / 20 * 1e3))
pr = cProfile.Profile(); pr.enable()
for _ in range(20): step()
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
