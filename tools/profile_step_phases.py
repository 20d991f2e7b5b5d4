"""Device-time phases of the training step bench.py times (paper config, FusedAdam): python tools/profile_step_phases.py [M]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200")]
import torch
import conv_tasnet, pit_criterion
from ctn_b200.optim import FusedAdam
M = int(sys.argv[1]) if len(sys.argv) > 1 else 3
T = 32000
cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
torch.manual_seed(0)
dev = torch.device("cuda")
model = conv_tasnet.ConvTasNet(**cfg).to(dev).train()
model.train_engine().overlap_wgrad = bool(int(os.environ.get("CTN_OVERLAP", "1")))
opt = FusedAdam(model.parameters(), lr=1e-3, max_norm=5.0)
h_src = (0.1 * torch.randn(M, 2, T)).pin_memory(); h_mix = h_src.sum(1).pin_memory()
lengths = torch.full((M,), T, dtype=torch.int64, device=dev)
h_loss = torch.empty((), dtype=torch.float32).pin_memory()
names = ["h2d", "fwd (prep + graph)", "loss", "zero_grad", "bwd", "clip+adam", "d2h"]
def step(rec=None):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)]
    ev[0].record()
    mix = h_mix.to(dev, non_blocking=True); src = h_src.to(dev, non_blocking=True); ev[1].record()
    est = model(mix); ev[2].record()
    loss, *_ = pit_criterion.cal_loss(src, est, lengths); ev[3].record()
    opt.zero_grad(); ev[4].record()
    loss.backward(); ev[5].record()
    opt.step(); ev[6].record()
    h_loss.copy_(loss.detach(), non_blocking=True); ev[7].record()
    torch.cuda.synchronize()
    if rec is not None:
        rec.append([ev[i].elapsed_time(ev[i + 1]) for i in range(len(names))])
for _ in range(5): step()
rec = []
for _ in range(10): step(rec)
g = [sum(r[i] for r in rec) / len(rec) for i in range(len(names))]
print("M=%d  total %.3f ms" % (M, sum(g)))
for n, a in zip(names, g): print("  %-20s %.3f ms" % (n, a))
