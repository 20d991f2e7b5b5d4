"""CPU oracle for the Conv-TasNet separation path.  TEST INFRASTRUCTURE ONLY.

This file is a functional restatement of the reference's algorithm
(`/root/reference/src/conv_tasnet.py`, `pit_criterion.py`, `utils.py`) written
against plain tensors and a flat ``{state_dict key: tensor}`` dictionary.  It
exists so that the CUDA product path can be checked on a machine where the
reference itself is not present (the GPU box).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it; the product path
(``conv-tasnet_b200/``) never does and fails loudly without its CUDA library.

Parity pin: the reference ships no golden vectors or runnable tests
(SURVEY.md §8c), so the oracle is pinned against OUTPUTS OF THE REFERENCE
ITSELF, generated in the build container by ``oracle/make_golden.py`` (which
imports the unmodified reference modules) and committed under
``tests/golden/``.  ``tests/test_oracle_golden.py`` holds the oracle to those
fixtures, forward, loss and every parameter gradient.

Arithmetic is torch CPU ops in the dtype of the inputs (float32 = the
reference's own arithmetic; float64 = ground truth for error attribution).
Backward comes from autograd over these functions, exactly as the reference's
backward comes from autograd over its modules.
"""
from itertools import permutations
import math

import torch
import torch.nn.functional as F

EPS = 1e-8  # conv_tasnet.py:10, pit_criterion.py:9


# ----------------------------------------------------------------------------
# state_dict key helpers (conv_tasnet.py:167-185, 217-263)
# ----------------------------------------------------------------------------
def is_bn(norm_type):
    """chose_norm (conv_tasnet.py:292-303): every norm_type but "gLN" / "cLN" is nn.BatchNorm1d."""
    return norm_type not in ("gLN", "cLN")


def block_keys(r, x, causal, norm_type="gLN"):
    """Key names of one TemporalBlock.  Causal models have a Chomp1d at
    dsconv index 1, which shifts the later indices by one
    (conv_tasnet.py:258-263).  BatchNorm1d blocks (:300-303) name their affine
    weight / bias and add running_mean / running_var / num_batches_tracked buffers."""
    base = "separator.network.2.%d.%d.net." % (r, x)
    o = 1 if causal else 0
    if is_bn(norm_type):
        n1, n2 = base + "2.", base + "3.net.%d." % (2 + o)
        return {
            "w1": base + "0.weight", "a1": base + "1.weight",
            "g1": n1 + "weight", "b1": n1 + "bias", "rm1": n1 + "running_mean", "rv1": n1 + "running_var",
            "nb1": n1 + "num_batches_tracked",
            "wd": base + "3.net.0.weight", "a2": base + "3.net.%d.weight" % (1 + o),
            "g2": n2 + "weight", "b2": n2 + "bias", "rm2": n2 + "running_mean", "rv2": n2 + "running_var",
            "nb2": n2 + "num_batches_tracked",
            "w2": base + "3.net.%d.weight" % (3 + o),
        }
    return {
        "w1": base + "0.weight",              # [H,B,1]
        "a1": base + "1.weight",              # [1]
        "g1": base + "2.gamma", "b1": base + "2.beta",   # [1,H,1]
        "wd": base + "3.net.0.weight",        # [H,1,P]
        "a2": base + "3.net.%d.weight" % (1 + o),
        "g2": base + "3.net.%d.gamma" % (2 + o), "b2": base + "3.net.%d.beta" % (2 + o),
        "w2": base + "3.net.%d.weight" % (3 + o),  # [B,H,1]
    }


def param_shapes(N, L, B, H, P, X, R, C, causal=False, norm_type="gLN"):
    """Ordered {key: shape} in the reference's state_dict order (BatchNorm1d buffers included)."""
    shapes = {"encoder.conv1d_U.weight": (N, 1, L),
              "separator.network.0.gamma": (1, N, 1),
              "separator.network.0.beta": (1, N, 1),
              "separator.network.1.weight": (B, N, 1)}
    for r in range(R):
        for x in range(X):
            k = block_keys(r, x, causal, norm_type)
            bn = is_bn(norm_type)
            nshape = (H,) if bn else (1, H, 1)
            shapes[k["w1"]] = (H, B, 1)
            shapes[k["a1"]] = (1,)
            shapes[k["g1"]] = nshape
            shapes[k["b1"]] = nshape
            if bn:
                shapes[k["rm1"]], shapes[k["rv1"]], shapes[k["nb1"]] = (H,), (H,), ()
            shapes[k["wd"]] = (H, 1, P)
            shapes[k["a2"]] = (1,)
            shapes[k["g2"]] = nshape
            shapes[k["b2"]] = nshape
            if bn:
                shapes[k["rm2"]], shapes[k["rv2"]], shapes[k["nb2"]] = (H,), (H,), ()
            shapes[k["w2"]] = (B, H, 1)
    shapes["separator.network.3.weight"] = (C * N, B, 1)
    shapes["decoder.basis_signals.weight"] = (L, N)
    return shapes


def random_params(N, L, B, H, P, X, R, C, causal=False, seed=0, dtype=torch.float32,
                  scale_norm=True, norm_type="gLN"):
    """Seeded synthetic weights with the reference's init *distribution*
    (xavier-normal for dim>1 incl. gamma/beta, PReLU 0.25; conv_tasnet.py:41-43).
    Not bit-identical to a same-seed reference construction; golden fixtures
    carry the reference's actual weights."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for k, shp in param_shapes(N, L, B, H, P, X, R, C, causal, norm_type).items():
        if is_bn(norm_type) and k.split(".")[-1] in ("bias", "running_mean", "running_var", "num_batches_tracked") \
                or (is_bn(norm_type) and shp == (H,)):
            # BatchNorm1d tensors: non-trivial affine and running statistics (as after some training)
            leaf = k.split(".")[-1]
            if leaf == "num_batches_tracked":
                out[k] = torch.tensor(100, dtype=torch.long)
            elif leaf == "running_var":
                out[k] = (0.5 + torch.rand(shp, generator=g, dtype=torch.float64)).to(dtype)
            elif leaf == "weight":
                out[k] = (1.0 + 0.2 * torch.randn(shp, generator=g, dtype=torch.float64)).to(dtype)
            else:
                out[k] = (0.1 * torch.randn(shp, generator=g, dtype=torch.float64)).to(dtype)
            continue
        if len(shp) == 1:
            out[k] = torch.full(shp, 0.25, dtype=dtype)
            continue
        rf = 1
        for s in shp[2:]:
            rf *= s
        fan_in, fan_out = shp[1] * rf, shp[0] * rf
        std = math.sqrt(2.0 / (fan_in + fan_out))
        out[k] = (torch.randn(shp, generator=g, dtype=torch.float64) * std).to(dtype)
    return out


# ----------------------------------------------------------------------------
# model pieces
# ----------------------------------------------------------------------------
def encoder(mixture, U):
    """conv_tasnet.py:108-117: relu(Conv1d(1,N,L,stride=L//2)).  [M,T] -> [M,N,K]."""
    L = U.shape[-1]
    return F.relu(F.conv1d(mixture.unsqueeze(1), U, stride=L // 2))


def cln(y, gamma, beta):
    """conv_tasnet.py:326-328: per (m,k) mean / biased var over channels."""
    mean = y.mean(dim=1, keepdim=True)
    var = y.var(dim=1, keepdim=True, unbiased=False)
    return gamma * (y - mean) / torch.pow(var + EPS, 0.5) + beta


def gln(y, gamma, beta):
    """conv_tasnet.py:352-354: per utterance mean / two-pass var over (C,K)."""
    mean = y.mean(dim=(1, 2), keepdim=True)
    var = ((y - mean) ** 2).mean(dim=(1, 2), keepdim=True)
    return gamma * (y - mean) / torch.pow(var + EPS, 0.5) + beta


BN_EPS = 1e-5  # nn.BatchNorm1d default; conv_tasnet.py:303 passes none


def bn_eval(y, weight, bias, running_mean, running_var):
    """nn.BatchNorm1d in eval mode (conv_tasnet.py:300-303): running statistics per channel."""
    inv = 1.0 / torch.sqrt(running_var + BN_EPS)
    return (y - running_mean[None, :, None]) * inv[None, :, None] * weight[None, :, None] + bias[None, :, None]


BN_MOMENTUM = 0.1  # nn.BatchNorm1d default


def bn_train(y, weight, bias, running_mean, running_var, num_batches_tracked):
    """nn.BatchNorm1d in training mode (conv_tasnet.py:300-303 builds it with defaults): per channel, mean and
    BIASED variance over the batch and time dims normalise the output; the running statistics move by
    momentum 0.1 towards the batch mean and the UNBIASED variance; num_batches_tracked counts the call.
    Returns (out, (new_running_mean, new_running_var, new_num_batches_tracked)).  Utterances of a batch
    are coupled through the statistics, so an utterance-sharded run needs them summed over ranks."""
    n = y.shape[0] * y.shape[2]
    mean = y.mean(dim=(0, 2))
    var = ((y - mean[None, :, None]) ** 2).mean(dim=(0, 2))
    out = (y - mean[None, :, None]) / torch.sqrt(var[None, :, None] + BN_EPS) * weight[None, :, None] + bias[None, :, None]
    unbiased = var * (n / max(n - 1, 1))
    new = ((1 - BN_MOMENTUM) * running_mean + BN_MOMENTUM * mean.detach(),
           (1 - BN_MOMENTUM) * running_var + BN_MOMENTUM * unbiased.detach(), num_batches_tracked + 1)
    return out, new


def _norm(norm_type, y, gamma, beta, stats=None):
    if norm_type == "gLN":
        return gln(y, gamma, beta)
    if norm_type == "cLN":
        return cln(y, gamma, beta)
    if stats is None:
        raise NotImplementedError("BatchNorm1d needs its running statistics")
    return bn_eval(y, gamma, beta, stats[0], stats[1])


def depthwise(x, wd, dilation, causal):
    """conv_tasnet.py:175-176,247-252,289: zero-padded dilated depthwise
    cross-correlation; causal = pad (P-1)d both sides then chomp the right."""
    H, _, P = wd.shape
    pad = (P - 1) * dilation if causal else (P - 1) * dilation // 2
    y = F.conv1d(x, wd, stride=1, padding=pad, dilation=dilation, groups=H)
    if causal:
        y = y[:, :, :-pad] if pad > 0 else y
    return y


def temporal_block(x, p, k, dilation, norm_type, causal, taps=None, new_stats=None):
    """conv_tasnet.py:227-237 + 259-272.  `taps` (optional dict) collects
    intermediates for per-kernel parity tests.  `new_stats` (a dict) switches BatchNorm1d blocks to
    training mode: batch statistics normalise, the updated buffers are stored in it under their keys."""
    a = F.conv1d(x, p[k["w1"]])
    pa = F.prelu(a, p[k["a1"]])
    bn = is_bn(norm_type)

    def norm(y, i):
        g, b = p[k["g%d" % i]], p[k["b%d" % i]]
        if not bn:
            return _norm(norm_type, y, g, b)
        if new_stats is None:
            return _norm(norm_type, y, g, b, (p[k["rm%d" % i]], p[k["rv%d" % i]]))
        out, (rm, rv, nb) = bn_train(y, g, b, p[k["rm%d" % i]], p[k["rv%d" % i]], p[k["nb%d" % i]])
        new_stats[k["rm%d" % i]], new_stats[k["rv%d" % i]], new_stats[k["nb%d" % i]] = rm, rv, nb
        return out

    n1 = norm(pa, 1)
    d = depthwise(n1, p[k["wd"]], dilation, causal)
    q = F.prelu(d, p[k["a2"]])
    n2 = norm(q, 2)
    out = F.conv1d(n2, p[k["w2"]]) + x
    if taps is not None:
        taps.update(a=a, pa=pa, n1=n1, d=d, q=q, n2=n2, out=out)
    return out


def separator(w, p, cfg, taps=None, new_stats=None):
    """conv_tasnet.py:192-209.  [M,N,K] -> mask [M,C,N,K]."""
    M, N, K = w.shape
    y = cln(w, p["separator.network.0.gamma"], p["separator.network.0.beta"])
    y = F.conv1d(y, p["separator.network.1.weight"])
    if taps is not None:
        taps["bottleneck"] = y
    for r in range(cfg["R"]):
        for x in range(cfg["X"]):
            k = block_keys(r, x, cfg["causal"], cfg["norm_type"])
            y = temporal_block(y, p, k, 2 ** x, cfg["norm_type"], cfg["causal"], new_stats=new_stats)
            if taps is not None:
                taps["block_%d_%d" % (r, x)] = y
    score = F.conv1d(y, p["separator.network.3.weight"]).view(M, cfg["C"], N, K)
    if cfg["mask_nonlinear"] == "softmax":
        return F.softmax(score, dim=1)
    if cfg["mask_nonlinear"] == "relu":
        return F.relu(score)
    raise ValueError("Unsupported mask non-linear function")  # conv_tasnet.py:208


def overlap_and_add(frames, frame_step):
    """utils.py:9-46 restated as the closed-form two-(or more-)term gather:
    out[t] = sum over frames k with 0 <= t - k*step < L of frames[k, t - k*step].
    Implemented with F.fold (a deterministic column-to-signal sum)."""
    *outer, K, L = frames.shape
    T_out = frame_step * (K - 1) + L
    x = frames.reshape(-1, K, L).transpose(1, 2)          # [*, L, K]
    y = F.fold(x, output_size=(1, T_out), kernel_size=(1, L), stride=(1, frame_step))
    return y.reshape(*outer, T_out)


def decoder(w, mask, V):
    """conv_tasnet.py:137-142."""
    L = V.shape[0]
    src = (w.unsqueeze(1) * mask).transpose(2, 3)        # [M,C,K,N]
    frames = src @ V.t()                                  # [M,C,K,L]
    return overlap_and_add(frames, L // 2)


def forward(mixture, p, cfg, taps=None, new_stats=None):
    """ConvTasNet.forward (conv_tasnet.py:45-60).  [M,T] -> [M,C,T].  new_stats: see temporal_block
    (model.train() with norm_type="BN")."""
    w = encoder(mixture, p["encoder.conv1d_U.weight"])
    if taps is not None:
        taps["mixture_w"] = w
    mask = separator(w, p, cfg, taps, new_stats)
    est = decoder(w, mask, p["decoder.basis_signals.weight"])
    T = mixture.shape[-1]
    return F.pad(est, (0, T - est.shape[-1]))


# ----------------------------------------------------------------------------
# PIT SI-SNR loss (pit_criterion.py)
# ----------------------------------------------------------------------------
def length_mask(source, lengths):
    """pit_criterion.py:101-113."""
    B, _, T = source.shape
    t = torch.arange(T, device=source.device).view(1, 1, T)
    return (t < lengths.view(B, 1, 1).to(t.device)).to(source.dtype)


def pairwise_si_snr(source, estimate, lengths):
    """pit_criterion.py:37-62.  Returns [B, C_est, C_tgt] in dB.  Does not
    mutate its inputs (the in-place masking side effect is the caller's)."""
    mask = length_mask(source, lengths)
    est = estimate * mask
    n = lengths.view(-1, 1, 1).to(source.dtype)
    zt = (source - source.sum(2, keepdim=True) / n) * mask
    ze = (est - est.sum(2, keepdim=True) / n) * mask
    st = zt.unsqueeze(1)          # [B,1,C,T]
    se = ze.unsqueeze(2)          # [B,C,1,T]
    dot = (se * st).sum(3, keepdim=True)
    energy = (st ** 2).sum(3, keepdim=True) + EPS
    proj = dot * st / energy
    noise = se - proj
    ratio = (proj ** 2).sum(3) / ((noise ** 2).sum(3) + EPS)
    return 10 * torch.log10(ratio + EPS)


def cal_si_snr_with_pit(source, estimate, lengths):
    """pit_criterion.py:27-76 -> (max_snr [B,1], perms [C!,C] int64, idx [B])."""
    C = source.shape[1]
    snr = pairwise_si_snr(source, estimate, lengths)
    perms = torch.tensor(list(permutations(range(C))), dtype=torch.long, device=source.device)
    # snr_set[b,p] = sum_i snr[b, i, perms[p,i]]  (pit_criterion.py:68-71)
    rows = torch.arange(C, device=source.device)
    snr_set = snr[:, rows.unsqueeze(0).expand_as(perms), perms].sum(-1)
    idx = torch.argmax(snr_set, dim=1)
    max_snr = snr_set.max(dim=1, keepdim=True).values / C
    return max_snr, perms, idx


def reorder_source(source, perms, idx):
    """pit_criterion.py:79-98: out[b,c] = source[b, perms[idx[b]][c]]."""
    sel = perms[idx]                                   # [B,C]
    return torch.gather(source, 1, sel.unsqueeze(-1).expand_as(source))


def cal_loss(source, estimate, lengths):
    """pit_criterion.py:12-24, functional: returns the masked estimate as a
    new tensor instead of mutating `estimate`."""
    max_snr, perms, idx = cal_si_snr_with_pit(source, estimate, lengths)
    loss = 0 - max_snr.mean()
    masked = estimate * length_mask(source, lengths)
    return loss, max_snr, masked, reorder_source(masked, perms, idx), idx


# ----------------------------------------------------------------------------
# shape arithmetic (bit-exact integer outputs; SURVEY.md §8 notation)
# ----------------------------------------------------------------------------
def cal_SISNR(ref_sig, out_sig, eps=1e-8):
    """SI-SNR of out_sig against ref_sig in dB (evaluate.py:128-144); 1-D numpy arrays of equal
    length, arithmetic in the arrays' dtype like the reference."""
    import numpy as np
    assert len(ref_sig) == len(out_sig)
    ref_zm = ref_sig - np.mean(ref_sig)                     # :134
    out_zm = out_sig - np.mean(out_sig)                     # :135
    energy = np.sum(ref_zm ** 2) + eps                      # :136
    proj = np.sum(ref_zm * out_zm) * ref_zm / energy        # :137
    noise = out_zm - proj                                   # :138
    ratio = np.sum(proj ** 2) / (np.sum(noise ** 2) + eps)  # :139
    return 10 * np.log(ratio + eps) / np.log(10.0)          # :140


def cal_SISNRi(src_ref, src_est, mix):
    """Average SI-SNR improvement over the sources (evaluate.py:108-125; the reference hard-codes
    sources 0 and 1, this averages over all rows, identical for C = 2)."""
    C = src_ref.shape[0]
    return sum(cal_SISNR(src_ref[c], src_est[c]) - cal_SISNR(src_ref[c], mix) for c in range(C)) / C


def frame_geometry(T, L):
    hop = L // 2
    K = (T - L) // hop + 1
    T_conv = (K - 1) * hop + L
    return hop, K, T_conv


def forward_flops(cfg, T):
    """Algorithmic FLOPs per utterance forward."""
    N, L, B, H, P, X, R, C = (cfg[k] for k in "NLBHPXRC")
    _, K, _ = frame_geometry(T, L)
    return 2 * K * (N * L + N * B + R * X * (2 * B * H + H * P) + B * C * N + C * N * L)
