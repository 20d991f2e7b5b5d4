"""Per-tensor gradient errors of a small tensor-core-sized model against autograd over the fp64 oracle.
    python tools/diag_bn_grads.py BN|gLN|cLN causal(0|1) [seed] [brief]
env: BN_EVAL=1 (eval-mode BatchNorm), NO_KINK=1 (all PReLU slopes 1, softmax mask: no branch in the network)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200"), os.path.join(ROOT, "tests")]
import torch
import conv_tasnet, pit_criterion
from oracle import convtasnet_oracle as O
norm, causal = sys.argv[1], bool(int(sys.argv[2]))
seed = int(sys.argv[3]) if len(sys.argv) > 3 else 9
brief = len(sys.argv) > 4
EVAL = os.environ.get("BN_EVAL") == "1"
NOKINK = os.environ.get("NO_KINK") == "1"
DEV = "cuda"
cfg = dict(N=64, L=16, B=64, H=128, P=3, X=3, R=2, C=2, norm_type=norm, causal=causal,
           mask_nonlinear="softmax" if NOKINK else "relu")
p32 = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=seed, norm_type=norm)
if NOKINK:
    p32 = {k: (torch.ones_like(v) if v.shape == (1,) else v) for k, v in p32.items()}
model = conv_tasnet.ConvTasNet(**cfg); model.load_state_dict(p32); model = model.to(DEV).train()
if EVAL: model.eval()
g = torch.Generator().manual_seed(seed + 1)
M, T = 3, 3000
src = (0.1 * torch.randn(M, 2, T, generator=g)).to(DEV); mix = src.sum(1)
lengths = torch.tensor([T, 2500, T], device=DEV)
est = model(mix); est0 = est.detach().clone()
loss, *_ = pit_criterion.cal_loss(src, est, lengths); loss.backward()
p64 = {k: (v.to(DEV).double().requires_grad_(True) if v.is_floating_point() and "running_" not in k
           else (v.to(DEV).double() if v.is_floating_point() else v.to(DEV))) for k, v in p32.items()}
ns = {} if (O.is_bn(norm) and not EVAL) else None
est_ref = O.forward(mix.double(), p64, cfg, new_stats=ns)
loss_ref, *_ = O.cal_loss(src.double(), est_ref, lengths); loss_ref.backward()
def rel(a, b): return float((a.double() - b.double()).norm() / b.double().norm())
fl = torch.cat([p.grad.double().flatten() for k, p in model.named_parameters()])
fr = torch.cat([p64[k].grad.flatten() for k, p in model.named_parameters()])
print(norm, causal, seed, "eval" if EVAL else "train", "nokink" if NOKINK else "", "fwd err %.2e flat grad err %.3e" % (rel(est0, est_ref), rel(fl, fr)))
if brief: sys.exit(0)
for k, p in model.named_parameters():
    print("%-50s n=%6d |g|=%9.3e err=%9.3e" % (k, p.numel(), float(p64[k].grad.norm()), rel(p.grad, p64[k].grad)))
