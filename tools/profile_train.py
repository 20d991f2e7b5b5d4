"""Phase breakdown of one training step (paper config, M utterances): python tools/profile_train.py [M]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200")]
import torch
import conv_tasnet, pit_criterion
M = int(sys.argv[1]) if len(sys.argv) > 1 else 3
cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
torch.manual_seed(0)
model = conv_tasnet.ConvTasNet(**cfg).cuda().train()
model.train_engine().overlap_wgrad = bool(int(os.environ.get("CTN_OVERLAP", "1")))
opt = torch.optim.Adam(model.parameters(), lr=1e-3)
src = 0.1 * torch.randn(M, 2, 32000, device="cuda"); mix = src.sum(1)
lengths = torch.full((M,), 32000, device="cuda")
names = ["fwd", "loss", "zero", "bwd", "clip", "adam"]
def step(rec=None):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(7)]; t = [time.perf_counter()]
    ev[0].record()
    est = model(mix); ev[1].record(); t.append(time.perf_counter())
    loss, *_ = pit_criterion.cal_loss(src, est, lengths); ev[2].record(); t.append(time.perf_counter())
    opt.zero_grad(); ev[3].record(); t.append(time.perf_counter())
    loss.backward(); ev[4].record(); t.append(time.perf_counter())
    torch.nn.utils.clip_grad_norm_(model.parameters(), 5.0); ev[5].record(); t.append(time.perf_counter())
    opt.step(); ev[6].record(); t.append(time.perf_counter())
    torch.cuda.synchronize()
    if rec is not None:
        rec.append(([ev[i].elapsed_time(ev[i + 1]) for i in range(6)], [(t[i + 1] - t[i]) * 1e3 for i in range(6)]))
for _ in range(4): step()
rec = []
for _ in range(5): step(rec)
g = [sum(r[0][i] for r in rec) / len(rec) for i in range(6)]; c = [sum(r[1][i] for r in rec) / len(rec) for i in range(6)]
print("M=%d  total gpu %.2f ms, host %.2f ms" % (M, sum(g), sum(c)))
for n, a, b in zip(names, g, c): print("  %-5s gpu %.3f ms   host-issue %.3f ms" % (n, a, b))
