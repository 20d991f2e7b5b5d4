"""GPU diagnostic: per-parameter gradient errors vs the fp64 oracle, with conditioning info."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200"), os.path.join(ROOT, "tests")]
import torch
from conftest import load_golden, rel_l2
from oracle import convtasnet_oracle as O
import conv_tasnet, pit_criterion
DEV = "cuda"

def run(cfg, w, mix, src, lengths, tag, topn=12):
    model = conv_tasnet.ConvTasNet(**cfg); model.load_state_dict(w); model = model.to(DEV).train()
    est = model(mix.to(DEV))
    loss, *_ = pit_criterion.cal_loss(src.to(DEV), est, lengths.to(DEV))
    loss.backward()
    res = {}
    for dt in (torch.float32, torch.float64):
        p = {k: v.to(DEV).to(dt).clone().requires_grad_(True) for k, v in w.items()}
        taps = {}
        e = O.forward(mix.to(DEV).to(dt), p, cfg, taps)
        l, *_ = O.cal_loss(src.to(DEV).to(dt), e, lengths.to(DEV))
        l.backward()
        res[dt] = {k: v.grad.double() for k, v in p.items()}
    rows = []
    for k, p in model.named_parameters():
        r64 = res[torch.float64][k]
        rows.append((rel_l2(p.grad, r64), rel_l2(res[torch.float32][k], r64), k, float(r64.norm()), p.numel()))
    rows.sort(reverse=True)
    print("==", tag, "loss", float(loss), "oracle64", float(l))
    for r in rows[:topn]:
        print("  ours %.3e  torch32 %.3e  %-48s |g|=%.4g n=%d" % r)
    if tag == "demo_gln":
        k = "separator.network.0.beta"
        print("  beta0 ours", dict(model.named_parameters())[k].grad.flatten().tolist())
        print("  beta0 ref ", res[torch.float64][k].flatten().tolist())
        plan = list(model.train_engine().plans.values())[0]
        print("  w (encoder out) rows:", plan.w.view(-1, cfg["N"]).tolist())
        print("  dn0 ours:", plan.dn0.view(-1, cfg["N"]).tolist())
        print("  rowstat0:", plan.rowstat0.tolist())

for name in ["demo_gln", "softmax_c3", "tc_small"]:
    cfg, z = load_golden(name)
    w = {k[2:]: v for k, v in z.items() if k.startswith("w:")}
    run(cfg, w, z["mixture"], z["sources"], z["lengths"], name)

cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
p32 = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=1234)
g = torch.Generator().manual_seed(2025)
src = 0.1 * torch.randn(3, 2, 32000, generator=g)
run(cfg, p32, src.sum(1), src, torch.tensor([32000, 24000, 16001]), "paper", topn=25)
