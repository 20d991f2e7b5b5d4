"""GPU diagnostic: worst per-parameter gradient error for a sweep of tiny configurations."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200"), os.path.join(ROOT, "tests")]
import torch
from conftest import rel_l2
from oracle import convtasnet_oracle as O
import conv_tasnet, pit_criterion
DEV = "cuda"

def run(cfg, M, T, seed=0, verbose=False):
    w = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=seed)
    g = torch.Generator().manual_seed(seed + 1)
    src = 0.1 * torch.randn(M, cfg["C"], T, generator=g)
    mix = src.sum(1)
    lengths = torch.full((M,), T)
    model = conv_tasnet.ConvTasNet(**cfg); model.load_state_dict(w); model = model.to(DEV).train()
    est = model(mix.to(DEV))
    loss, *_ = pit_criterion.cal_loss(src.to(DEV), est, lengths.to(DEV))
    loss.backward()
    p = {k: v.to(DEV).double().requires_grad_(True) for k, v in w.items()}
    e = O.forward(mix.to(DEV).double(), p, cfg)
    l, *_ = O.cal_loss(src.to(DEV).double(), e, lengths.to(DEV))
    l.backward()
    rows = sorted(((rel_l2(pp.grad, p[k].grad), k) for k, pp in model.named_parameters()), reverse=True)
    print("N%d L%d B%d H%d X%d R%d M%d T%d seed%d: loss %.5f/%.5f worst %.2e %s | 2nd %.2e %s" % (
        cfg["N"], cfg["L"], cfg["B"], cfg["H"], cfg["X"], cfg["R"], M, T, seed, float(loss.detach()), float(l.detach()),
        rows[0][0], rows[0][1], rows[1][0], rows[1][1]))
    if verbose:
        for r in rows: print("    %.2e %s" % r)

base = dict(P=3, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
for (N, L, B, H, X, R, M, T) in [
    (3, 4, 2, 3, 3, 2, 2, 12), (3, 4, 2, 3, 1, 1, 2, 12), (3, 4, 2, 3, 2, 1, 2, 12), (3, 4, 2, 3, 2, 1, 2, 120),
    (8, 4, 8, 8, 2, 1, 2, 12), (8, 4, 8, 8, 2, 1, 2, 120), (3, 4, 8, 8, 2, 1, 2, 120), (8, 4, 2, 8, 2, 1, 2, 120),
    (8, 4, 8, 3, 2, 1, 2, 120), (5, 4, 6, 7, 3, 1, 2, 400), (12, 4, 12, 20, 3, 1, 2, 400), (16, 8, 8, 16, 2, 2, 3, 100)]:
    for seed in (0, 1):
        run(dict(base, N=N, L=L, B=B, H=H, X=X, R=R), M, T, seed)
