"""GPU diagnostic: compare backward workspaces of block 0 with autograd intermediates (demo_gln)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200"), os.path.join(ROOT, "tests")]
import torch
import torch.nn.functional as F
from conftest import load_golden, rel_l2
from oracle import convtasnet_oracle as O
import conv_tasnet, pit_criterion
DEV = "cuda"
name = sys.argv[1] if len(sys.argv) > 1 else "demo_gln"
cfg, z = load_golden(name)
w = {k[2:]: v for k, v in z.items() if k.startswith("w:")}
model = conv_tasnet.ConvTasNet(**cfg); model.load_state_dict(w); model = model.to(DEV).train()
mix, src, lengths = z["mixture"].to(DEV), z["sources"].to(DEV), z["lengths"].to(DEV)
est = model(mix)
loss, *_ = pit_criterion.cal_loss(src, est, lengths)
plan = list(model.train_engine().plans.values())[0]
STOP = int(sys.argv[2]) if len(sys.argv) > 2 else -1   # block whose workspaces to inspect
if STOP >= 0:
    NBLK = cfg["R"] * cfg["X"]
    ncalls = 4 + 12 * (NBLK - 1 - STOP) + 6
    orig = plan.bwd_calls
    plan.bwd_calls = orig[:ncalls]
loss.backward()
BLK = STOP if STOP >= 0 else 0

p = {k: v.to(DEV).double().requires_grad_(True) for k, v in w.items()}
keep = {}
def tap(n, t):
    t.retain_grad(); keep[n] = t; return t
wenc = tap("w", O.encoder(mix.double(), p["encoder.conv1d_U.weight"]))
y = tap("n0", O.cln(wenc, p["separator.network.0.gamma"], p["separator.network.0.beta"]))
y = tap("x0", F.conv1d(y, p["separator.network.1.weight"]))
bi = 0
for r in range(cfg["R"]):
    for x in range(cfg["X"]):
        k = O.block_keys(r, x, cfg["causal"])
        a = tap("a%d" % bi, F.conv1d(y, p[k["w1"]]))
        pa = F.prelu(a, p[k["a1"]])
        n1 = tap("n1_%d" % bi, O.gln(pa, p[k["g1"]], p[k["b1"]]))
        d = tap("d%d" % bi, O.depthwise(n1, p[k["wd"]], 2 ** x, cfg["causal"]))
        q = F.prelu(d, p[k["a2"]])
        u = tap("u%d" % bi, O.gln(q, p[k["g2"]], p[k["b2"]]))
        y = tap("x%d" % (bi + 1), F.conv1d(u, p[k["w2"]]) + y)
        bi += 1
M, N, K = wenc.shape
score = tap("score", F.conv1d(y, p["separator.network.3.weight"]))
mask = F.relu(score.view(M, cfg["C"], N, K)) if cfg["mask_nonlinear"] == "relu" else F.softmax(score.view(M, cfg["C"], N, K), 1)
e = O.decoder(wenc, mask, p["decoder.basis_signals.weight"])
e = F.pad(e, (0, mix.shape[-1] - e.shape[-1]))
l, *_ = O.cal_loss(src.double(), e, lengths)
l.backward()
def fm(t):  # [M,C,K] -> frames-major [M*K, C]
    return t.permute(0, 2, 1).reshape(-1, t.shape[1])
def hl(pair):
    return pair[0].double() + pair[1].double()
print("loss", float(loss), float(l))
print("dscore", rel_l2(hl(plan.ds_hl), fm(keep["score"].grad)))
print("g (d x_blk+1 if STOP else x0)", rel_l2(plan.g, fm(keep["x%d" % (BLK + 1 if STOP >= 0 else 0)].grad)))
print("du block0", rel_l2(plan.du, fm(keep["u%d" % BLK].grad)))
print("dn1 block0", rel_l2(plan.dn1, fm(keep["n1_%d" % BLK].grad)))
print("da block0 (hl_H)", rel_l2(hl(plan.hl_H[(len(plan.a) - 1 - BLK) & 1]), fm(keep["a%d" % BLK].grad)))
print("dn0", rel_l2(plan.dn0, fm(keep["n0"].grad)))
print("fwd a", rel_l2(plan.a[BLK], fm(keep["a%d" % BLK])), "d", rel_l2(plan.d[BLK], fm(keep["d%d" % BLK])))
print("du ours", plan.du.tolist()); print("du ref", fm(keep["u%d" % BLK].grad).tolist())
print("dn1 ours", plan.dn1.tolist()); print("dn1 ref", fm(keep["n1_%d" % BLK].grad).tolist())
print("d", plan.d[BLK].tolist()); print("a", plan.a[BLK].tolist())
print("bstat", plan.bstat.tolist(), "stat2", plan.stat2[BLK].tolist(), "stat1", plan.stat1[BLK].tolist())
for i in range(bi + 1):
    print("fwd x%d" % i, rel_l2(hl(plan.x_hl[i]), fm(keep["x%d" % i])))
for k_, pp in model.named_parameters():
    print("  %-48s %.3e" % (k_, rel_l2(pp.grad, p[k_].grad)))
