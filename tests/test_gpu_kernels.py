"""Per-kernel parity of the CUDA path (through the C ABI) against the oracle.

Tolerances: the GEMM operands carry 16 mantissa bits each (bf16 hi + lo), fp32
accumulate, so matrix products are compared at 3e-5 rel-L2 against an fp64
product of the SAME fp32 inputs; everything else is fp32 arithmetic in a
different summation order and is held to 1e-5.  Integer outputs (frame counts,
permutation indices) are bit-exact.
"""
import pytest
import torch
import torch.nn.functional as F

from conftest import load_golden, rel_l2
from oracle import convtasnet_oracle as O

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import ctypes as C

    import kernels as Kn
    from ctn_b200 import _native as nat

DEV = "cuda"


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV)


# ----------------------------------------------------------------- encoder
@pytest.mark.parametrize("M,T,N,L", [(2, 12, 3, 4), (3, 1207, 64, 16), (2, 32007, 256, 20), (1, 4000, 512, 16),
                                     (1, 57, 5, 5), (2, 997, 30, 20), (1, 16, 6, 16), (2, 5000, 1024, 20)])
def test_encoder(M, T, N, L):
    mix = _rand(M, T, seed=1, scale=0.1)
    U = _rand(N, 1, L, seed=2, scale=0.3)
    w, hi, lo, st, K = Kn.encoder(mix, U)
    assert K == O.frame_geometry(T, L)[1]
    ref = O.encoder(mix.double(), U.double())                  # [M,N,K]
    assert ref.shape[-1] == K
    ref_rows = Kn.to_rows(ref)
    assert rel_l2(w, ref_rows) < 1e-6
    assert rel_l2(hi.float() + lo.float(), ref_rows) < 2e-5
    mean = ref_rows.mean(1)
    rstd = 1.0 / torch.sqrt(ref_rows.var(1, unbiased=False) + 1e-8)
    assert rel_l2(st[:, 0], mean) < 1e-5
    assert rel_l2(st[:, 1], rstd) < 1e-4


# ------------------------------------------------------------- weight prep
@pytest.mark.parametrize("rows,cols", [(3, 2), (256, 512), (1536, 128)])
def test_weight_prep(rows, cols):
    W = _rand(rows, cols, seed=3, scale=0.1)
    gamma, beta = _rand(cols, seed=4), _rand(cols, seed=5)
    hi, lo, g, b = Kn.weight_prep(W, gamma, beta)
    Wg = W.double() * gamma.double()
    assert rel_l2(hi.float() + lo.float(), Wg) < 1e-5
    assert rel_l2(g, (hi.double() + lo.double()).sum(1)) < 1e-5
    assert rel_l2(b, W.double() @ beta.double()) < 1e-5
    hi_t, lo_t, _, _ = Kn.weight_prep(W, None, None, transpose=True)
    assert rel_l2(hi_t.float() + lo_t.float(), W.double().t()) < 1e-5


# -------------------------------------------------------------------- GEMM
def _gemm_case(M, K, Cin, Cout, simt, seed=0):
    R = M * K
    A = _rand(R, Cin, seed=seed + 10)
    W = _rand(Cout, Cin, seed=seed + 11, scale=0.1)
    a_hi, a_lo = Kn.split_hl(A)
    w_hi, w_lo = Kn.split_hl(W)
    Ad = a_hi.double() + a_lo.double()
    Wd = w_hi.double() + w_lo.double()
    return A, W, a_hi, a_lo, w_hi, w_lo, Ad @ Wd.t()


SHAPES = [(1, 5, 2, 3, True), (2, 149, 64, 128, False), (2, 149, 64, 128, True), (1, 300, 256, 512, False),
          (3, 3199, 256, 512, False), (3, 3199, 512, 256, False), (1, 3999, 128, 1536, False),
          (2, 130, 24, 40, False), (16, 3199, 256, 512, False)]


@pytest.mark.parametrize("M,K,Cin,Cout,simt", SHAPES)
def test_gemm_plain(M, K, Cin, Cout, simt):
    A, W, a_hi, a_lo, w_hi, w_lo, ref = _gemm_case(M, K, Cin, Cout, simt)
    res = Kn.gemm(M, K, a_hi, a_lo, w_hi, w_lo, out_f32=True, out_hl=True, force_simt=simt)
    assert rel_l2(res["out"], ref) < 3e-5
    assert rel_l2(res["hi"].float() + res["lo"].float(), ref) < 5e-5
    # fp32 product of the unsplit inputs: the split itself costs < 2e-5
    assert rel_l2(res["out"], A.double() @ W.double().t()) < 5e-5


@pytest.mark.parametrize("M,K,Cin,Cout,simt", [(2, 5, 3, 2, True), (2, 149, 128, 64, False), (3, 3199, 512, 256, False),
                                               (1, 300, 128, 256, True)])
@pytest.mark.parametrize("norm", ["gLN", "cLN"])
def test_gemm_folded_norm_residual(M, K, Cin, Cout, simt, norm):
    """pointwise conv: W.(gamma*(q-mu)*rstd+beta) + x  (conv_tasnet.py:254-256,237)"""
    R = M * K
    q = _rand(R, Cin, seed=20) * 0.7 + 0.3
    W = _rand(Cout, Cin, seed=21, scale=0.1)
    gamma, beta = _rand(Cin, seed=22), _rand(Cin, seed=23)
    x = _rand(R, Cout, seed=24)
    q_hi, q_lo = Kn.split_hl(q)
    w_hi, w_lo, col_g, col_bias = Kn.weight_prep(W, gamma, beta)
    qd = q.double().view(M, K, Cin)
    if norm == "gLN":
        # statistics arrive as (sum, sumsq) partials per utterance
        P = 7
        parts = torch.zeros(M, P, 2, device=DEV)
        chunks = torch.tensor_split(torch.arange(K), P)
        for m in range(M):
            for i, ch in enumerate(chunks):
                if len(ch):
                    parts[m, i, 0] = qd[m, ch].sum()
                    parts[m, i, 1] = (qd[m, ch] ** 2).sum()
        mean = qd.mean(dim=(1, 2), keepdim=True)
        var = ((qd - mean) ** 2).mean(dim=(1, 2), keepdim=True)
        kw = dict(in_norm=nat.NORM_GLN, in_stat=parts, in_P=P, in_count=float(K * Cin))
    else:
        mean = qd.mean(dim=2, keepdim=True)
        var = qd.var(dim=2, unbiased=False, keepdim=True)
        st = torch.stack([mean.flatten(), 1.0 / torch.sqrt(var.flatten() + 1e-8)], 1).float().contiguous()
        kw = dict(in_norm=nat.NORM_CLN, in_stat=st)
    normed = gamma.double() * (qd - mean) / torch.sqrt(var + 1e-8) + beta.double()
    ref = normed.view(R, Cin) @ W.double().t() + x.double()
    res = Kn.gemm(M, K, q_hi, q_lo, w_hi, w_lo, out_f32=True, out_hl=True, resid=x, col_g=col_g,
                  col_bias=col_bias, force_simt=simt, **kw)
    assert rel_l2(res["out"], ref) < 5e-5
    assert rel_l2(res["hi"].float() + res["lo"].float(), ref) < 7e-5


@pytest.mark.parametrize("M,K,Cin,Cout,simt", [(2, 5, 2, 3, True), (2, 149, 64, 128, False), (3, 3199, 256, 512, False),
                                               (2, 140, 64, 136, False), (1, 300, 64, 384, True)])
@pytest.mark.parametrize("norm", ["gLN", "cLN"])
def test_gemm_output_statistics(M, K, Cin, Cout, simt, norm):
    """conv1x1 + statistics of prelu(out) for the following norm (conv_tasnet.py:217-219)"""
    A, W, a_hi, a_lo, w_hi, w_lo, ref = _gemm_case(M, K, Cin, Cout, simt, seed=30)
    slope = torch.tensor([0.25], device=DEV)
    mode = nat.NORM_GLN if norm == "gLN" else nat.NORM_CLN
    res = Kn.gemm(M, K, a_hi, a_lo, w_hi, w_lo, out_stat=mode, stat_prelu=slope, force_simt=simt)
    assert rel_l2(res["out"], ref) < 3e-5
    p = F.prelu(ref, slope.double()).view(M, K, Cout)
    st = res["stat"].double()
    assert not torch.isnan(st).any()
    if norm == "gLN":
        assert rel_l2(st[..., 0].sum(1), p.sum(dim=(1, 2))) < 1e-4
        assert rel_l2(st[..., 1].sum(1), (p ** 2).sum(dim=(1, 2))) < 1e-4
    else:
        assert rel_l2(st[..., 0].sum(1), p.sum(dim=2).flatten()) < 1e-4
        assert rel_l2(st[..., 1].sum(1), (p ** 2).sum(dim=2).flatten()) < 1e-4
    # store_prelu: the stored output is activated too (inference), statistics unchanged
    res2 = Kn.gemm(M, K, a_hi, a_lo, w_hi, w_lo, out_stat=mode, stat_prelu=slope, force_simt=simt, store_prelu=True)
    assert rel_l2(res2["out"], p.reshape(M * K, Cout)) < 3e-5
    assert torch.equal(res2["stat"], res["stat"])


# --------------------------------------------------------------- depthwise
@pytest.mark.parametrize("H,K,M", [(3, 5, 2), (16, 24, 2), (512, 3199, 2)])
@pytest.mark.parametrize("dil", [1, 2, 8, 128])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("norm", ["gLN", "cLN"])
def test_dwconv(H, K, M, dil, causal, norm):
    if H < 512 and dil > 8:
        pytest.skip("large dilation only at the paper shape")
    P = 3
    a = _rand(M, H, K, seed=40)
    a1 = torch.tensor([0.25], device=DEV)
    a2 = torch.tensor([0.1], device=DEV)
    g1, b1 = _rand(H, seed=41), _rand(H, seed=42)
    wd = _rand(H, 1, P, seed=43, scale=0.5)
    ad = a.double()
    pa = F.prelu(ad, a1.double())
    nfun = O.gln if norm == "gLN" else O.cln
    n1 = nfun(pa, g1.double().view(1, H, 1), b1.double().view(1, H, 1))
    d_ref = O.depthwise(n1, wd.double(), dil, causal)
    q_ref = F.prelu(d_ref, a2.double())
    rows = Kn.to_rows(a)
    lib = nat.lib()
    if norm == "gLN":
        in_P = 5
        parts = torch.zeros(M, in_P, 2, device=DEV)
        parts[:, 0, 0] = pa.sum(dim=(1, 2))
        parts[:, 3, 1] = (pa ** 2).sum(dim=(1, 2))
        mode = nat.NORM_GLN
    else:
        in_P = 2
        parts = torch.zeros(M * K, in_P, 2, device=DEV)
        par = Kn.to_rows(pa)
        parts[:, 0, 0] = par[:, : H // 2].sum(1)
        parts[:, 1, 0] = par[:, H // 2:].sum(1)
        parts[:, 0, 1] = (par[:, : H // 2] ** 2).sum(1)
        parts[:, 1, 1] = (par[:, H // 2:] ** 2).sum(1)
        mode = nat.NORM_CLN
    pre, hi, lo, st = Kn.dwconv(M, K, rows, a1, g1, b1, mode, parts, in_P, wd, a2, dil, causal)
    assert rel_l2(pre, Kn.to_rows(d_ref)) < 2e-5
    qr = Kn.to_rows(q_ref)
    assert rel_l2(hi.float() + lo.float(), qr) < 3e-5
    assert not torch.isnan(st).any()
    if norm == "gLN":
        assert st.shape[1] == lib.ctn_dwconv_gln_partials(K, H, P, dil)
        assert rel_l2(st[..., 0].double().sum(1), q_ref.sum(dim=(1, 2))) < 1e-4
        assert rel_l2(st[..., 1].double().sum(1), (q_ref ** 2).sum(dim=(1, 2))) < 1e-4
    else:
        assert rel_l2(st[:, 0], qr.mean(1)) < 1e-4
        assert rel_l2(st[:, 1], 1.0 / torch.sqrt(qr.var(1, unbiased=False) + 1e-8)) < 1e-3


@pytest.mark.parametrize("H,K,M", [(64, 40, 3), (128, 700, 2), (512, 3199, 3)])
@pytest.mark.parametrize("dil", [1, 2, 4, 8, 16, 32, 64, 128])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("preact", [False, True])
def test_fused_depthwise_pointwise(H, K, M, dil, causal, preact):
    """ctn_dwpw_fwd + ctn_resid_update == PReLU, gLN, depthwise, PReLU, gLN, pointwise, residual
    (conv_tasnet.py:218-256, :237) of the oracle."""
    B, P = 256, 3
    a = _rand(M, H, K, seed=50)
    a1 = torch.tensor([0.25], device=DEV)
    a2 = torch.tensor([0.1], device=DEV)
    g1, b1 = 1.0 + 0.3 * _rand(H, seed=51), _rand(H, seed=52)
    g2, b2 = 1.0 + 0.3 * _rand(H, seed=53), _rand(H, seed=54)
    wd = _rand(H, 1, P, seed=55, scale=0.5)
    W2 = _rand(B, H, seed=56, scale=0.05)
    x = _rand(M, B, K, seed=57)
    ad = a.double()
    pa = F.prelu(ad, a1.double())
    n1 = O.gln(pa, g1.double().view(1, H, 1), b1.double().view(1, H, 1))
    q_ref = F.prelu(O.depthwise(n1, wd.double(), dil, causal), a2.double())
    n2 = O.gln(q_ref, g2.double().view(1, H, 1), b2.double().view(1, H, 1))
    xr_hi, xr_lo = Kn.split_hl(Kn.to_rows(x))
    x_in = (xr_hi.double() + xr_lo.double())
    out_ref = torch.einsum("bh,mhk->mbk", W2.double(), n2) + Kn.from_rows(x_in, M)
    in_P = 5
    parts = torch.zeros(M, in_P, 2, device=DEV)
    parts[:, 1, 0] = pa.sum(dim=(1, 2))
    parts[:, 4, 1] = (pa ** 2).sum(dim=(1, 2))
    w_hi, w_lo, col_g, col_bias = Kn.weight_prep(W2, g2, b2)
    if preact:   # the kernel is handed PReLU(a) (what ctn_gemm1x1 stores with store_prelu) and no slope
        y, st = Kn.dwpw(M, K, Kn.to_rows(pa.float()), None, g1, b1, parts, in_P, wd, a2, w_hi, w_lo, dil, causal)
    else:
        y, st = Kn.dwpw(M, K, Kn.to_rows(a), a1, g1, b1, parts, in_P, wd, a2, w_hi, w_lo, dil, causal)
    assert not torch.isnan(st).any()
    assert rel_l2(st[..., 0].double().sum(1), q_ref.sum(dim=(1, 2))) < 1e-4
    assert rel_l2(st[..., 1].double().sum(1), (q_ref ** 2).sum(dim=(1, 2))) < 1e-4
    assert not torch.isnan(y).any()                # every frame of every utterance was written
    y_ref = torch.einsum("bh,mhk->mbk", (w_hi.double() + w_lo.double()), q_ref)
    assert rel_l2(y, Kn.to_rows(y_ref)) < 2e-5
    Kn.resid_update(xr_hi, xr_lo, y, st, float(K) * H, col_g, col_bias, M, K)
    got = Kn.from_rows(xr_hi.double() + xr_lo.double(), M)
    assert rel_l2(got, out_ref) < 2e-5


@pytest.mark.parametrize("K,M,H", [(40, 3, 256), (700, 2, 512), (3199, 5, 512), (128, 1, 512), (129, 4, 256),
                                   (1000, 3, 768), (3199, 40, 1024), (3199, 64, 512)])
def test_fused_update_conv1x1(K, M, H):
    """ctn_updc1_fwd == ctn_resid_update followed by ctn_gemm1x1 (store_prelu, gLN statistics)."""
    B = 256
    x = _rand(M * K, B, seed=60)
    y = _rand(M * K, B, seed=61, scale=0.7)
    g2, b2 = 1.0 + 0.3 * _rand(512, seed=62), _rand(512, seed=63)
    W2 = _rand(B, 512, seed=64, scale=0.05)
    _, _, col_g, col_bias = Kn.weight_prep(W2, g2, b2)
    P = 7
    stat = torch.rand(M, P, 2, device=DEV) * 100 + 50
    count = float(K) * 512
    stat[:, :, 1] = stat[:, :, 0] ** 2 / (count / P) * 3.0 + 10.0
    W1 = _rand(H, B, seed=65, scale=0.06)
    w_hi, w_lo, _, _ = Kn.weight_prep(W1)
    slope = torch.tensor([0.25], device=DEV)
    x_hi, x_lo = Kn.split_hl(x)
    # reference path: the two unfused kernels
    r_hi, r_lo = x_hi.clone(), x_lo.clone()
    Kn.resid_update(r_hi, r_lo, y, stat, count, col_g, col_bias, M, K)
    ref = Kn.gemm(M, K, r_hi, r_lo, w_hi, w_lo, out_stat=nat.NORM_GLN, stat_prelu=slope, store_prelu=True)
    a, st = Kn.updc1(M, K, x_hi, x_lo, y, stat, count, col_g, col_bias, w_hi, w_lo, slope)
    assert torch.equal(x_hi, r_hi) and torch.equal(x_lo, r_lo)       # same arithmetic, same rounding
    assert not torch.isnan(a).any() and not torch.isnan(st).any()
    assert rel_l2(a, ref["out"]) < 1e-6
    assert rel_l2(st.double().sum(1), ref["stat"].double().sum(1)) < 1e-6
    # and against fp64
    s = stat.double().sum(1)
    mean = s[:, 0] / count
    r = 1.0 / torch.sqrt((s[:, 1] / count - mean * mean).clamp_min(0) + 1e-8)
    mm = torch.arange(M, device=DEV).repeat_interleave(K)
    xn = (x_hi.double() + x_lo.double())
    x_ref = (Kn.split_hl(x)[0].double() + Kn.split_hl(x)[1].double()) + r[mm, None] * y.double() \
        - (r * mean)[mm, None] * col_g.double()[None] + col_bias.double()[None]
    assert rel_l2(xn, x_ref) < 2e-5
    a_ref = F.prelu(xn @ (w_hi.double() + w_lo.double()).t(), slope.double())
    assert rel_l2(a, a_ref) < 2e-5


@pytest.mark.parametrize("K,M,H", [(700, 2, 512), (3199, 3, 512), (129, 4, 256)])
def test_fused_update_conv1x1_training_outputs(K, M, H):
    """The training forward's variant of ctn_updc1_fwd: the updated residual stream goes to a second pair (the input
    pair, which autograd saves for dW1, stays intact) and `a` leaves BEFORE the PReLU; statistics as before."""
    B = 256
    x = _rand(M * K, B, seed=70)
    y = _rand(M * K, B, seed=71, scale=0.7)
    g2, b2 = 1.0 + 0.3 * _rand(512, seed=72), _rand(512, seed=73)
    _, _, col_g, col_bias = Kn.weight_prep(_rand(B, 512, seed=74, scale=0.05), g2, b2)
    stat = torch.rand(M, 7, 2, device=DEV) * 100 + 50
    count = float(K) * 512
    stat[:, :, 1] = stat[:, :, 0] ** 2 / (count / 7) * 3.0 + 10.0
    w_hi, w_lo, _, _ = Kn.weight_prep(_rand(H, B, seed=75, scale=0.06))
    slope = torch.tensor([0.25], device=DEV)
    x_hi, x_lo = Kn.split_hl(x)
    r_hi, r_lo = x_hi.clone(), x_lo.clone()
    a_ref, st_ref = Kn.updc1(M, K, r_hi, r_lo, y, stat, count, col_g, col_bias, w_hi, w_lo, slope)     # in place, PReLU'd
    o_hi = torch.full_like(x_hi, float("nan"))
    o_lo = torch.full_like(x_lo, float("nan"))
    keep_hi, keep_lo = x_hi.clone(), x_lo.clone()
    a_raw, st = Kn.updc1(M, K, x_hi, x_lo, y, stat, count, col_g, col_bias, w_hi, w_lo, slope, x_out=(o_hi, o_lo),
                         store_raw=True)
    assert torch.equal(x_hi, keep_hi) and torch.equal(x_lo, keep_lo)       # the input pair is untouched
    assert torch.equal(o_hi, r_hi) and torch.equal(o_lo, r_lo)             # same update, other destination
    assert torch.equal(F.prelu(a_raw, slope), a_ref)                       # same product, PReLU not applied
    assert (a_raw < 0).any()
    assert torch.equal(st, st_ref)


@pytest.mark.parametrize("dil,causal", [(1, False), (8, True), (128, False)])
def test_fused_depthwise_pointwise_saves_the_depthwise_output(dil, causal):
    """d_out of ctn_dwpw_fwd (training) = the depthwise output before the second PReLU, as ctn_dwconv_fwd's out_pre."""
    H, K, M, B, P = 512, 700, 3, 256, 3
    a = Kn.to_rows(_rand(M, H, K, seed=80))
    a1 = torch.tensor([0.25], device=DEV)
    a2 = torch.tensor([0.1], device=DEV)
    g1, b1 = 1.0 + 0.3 * _rand(H, seed=81), _rand(H, seed=82)
    wd = _rand(H, 1, P, seed=83, scale=0.5)
    w_hi, w_lo, _, _ = Kn.weight_prep(_rand(B, H, seed=84, scale=0.05), 1.0 + 0.3 * _rand(H, seed=85), _rand(H, seed=86))
    pa = F.prelu(a, a1).view(M, K * H)
    parts = torch.zeros(M, 3, 2, device=DEV)
    parts[:, 0, 0] = pa.double().sum(1).float()
    parts[:, 2, 1] = (pa.double() ** 2).sum(1).float()
    y, st, d = Kn.dwpw(M, K, a, a1, g1, b1, parts, 3, wd, a2, w_hi, w_lo, dil, causal, want_d=True)
    y0, st0 = Kn.dwpw(M, K, a, a1, g1, b1, parts, 3, wd, a2, w_hi, w_lo, dil, causal)
    assert torch.equal(y, y0) and torch.equal(st, st0)                    # the extra store changes nothing else
    pre, _, _, _ = Kn.dwconv(M, K, a, a1, g1, b1, nat.NORM_GLN, parts, 3, wd, a2, dil, causal)
    assert not torch.isnan(d).any()
    assert rel_l2(d, pre) < 2e-6


def test_dwconv_rejects_length_changing_config():
    """P=2 non-causal changes the length; the reference crashes at the residual add"""
    H, K, M = 4, 10, 1
    z = torch.zeros(M * K, H, device=DEV)
    one = torch.ones(1, device=DEV)
    with pytest.raises(RuntimeError):
        Kn.dwconv(M, K, z, one, torch.ones(H, device=DEV), torch.zeros(H, device=DEV), nat.NORM_GLN,
                  torch.zeros(M, 1, 2, device=DEV), 1, torch.zeros(H, 1, 2, device=DEV), one, 1, False)


# ----------------------------------------------------------------- decoder
@pytest.mark.parametrize("M,K,N,L,C,extra", [(2, 5, 3, 4, 2, 0), (2, 149, 64, 16, 2, 7), (1, 3199, 256, 20, 2, 0),
                                             (2, 500, 512, 16, 3, 5), (1, 40, 6, 5, 2, 1)])
@pytest.mark.parametrize("mask", ["relu", "softmax"])
def test_decoder(M, K, N, L, C, extra, mask):
    hop = L // 2
    T = (K - 1) * hop + L + extra
    score = _rand(M, C * N, K, seed=50)
    w = _rand(M, N, K, seed=51).abs()
    V = _rand(L, N, seed=52, scale=0.2)
    sc = score.double().view(M, C, N, K)
    m = F.softmax(sc, dim=1) if mask == "softmax" else F.relu(sc)
    ref = F.pad(O.decoder(w.double(), m, V.double()), (0, extra))
    est = Kn.decoder(Kn.to_rows(score), Kn.to_rows(w), V, M, K, C, T,
                     nat.MASK_SOFTMAX if mask == "softmax" else nat.MASK_RELU)
    assert est.shape == ref.shape
    assert not torch.isnan(est).any()
    assert rel_l2(est, ref) < 1e-5
    if extra:
        assert torch.count_nonzero(est[..., -extra:]) == 0


def test_overlap_and_add_matches_reference_fixtures():
    import utils as U
    z = load_golden("ola_cases")
    for tag in ("demo", "l20", "l16", "odd"):
        out = U.overlap_and_add(z[tag + "_in"].to(DEV), int(z[tag + "_step"]))
        ref = z[tag + "_out"].to(DEV)
        assert out.shape == ref.shape
        if tag != "odd":
            assert torch.equal(out, ref)       # two commutative addends onto zero: bit-identical
        else:
            assert torch.allclose(out, ref, atol=1e-6)


# ------------------------------------------------- partial reduction, BatchNorm fold
def test_reduce_partials_wide_tall_and_strided_segments():
    """ctn_reduce_partials: one launch over segments of both layouts -- wide (37 partials x 70 001 outputs,
    the weight-gradient shape), tall (1 500 and 129 partials x 512 / 7 outputs, the channel sums of the
    depthwise backward), strided source / destination (the [H, 1, P] depthwise weight) -- and accumulation."""
    g = torch.Generator().manual_seed(60)
    wide = torch.randn(37, 70001, generator=g).to(DEV)
    tall = torch.randn(1500, 7 * 512, generator=g).to(DEV)
    tall2 = torch.randn(129, 40, generator=g).to(DEV)
    P = 3
    d_wide = torch.zeros(70001, device=DEV)
    d_t0 = torch.zeros(512, device=DEV)
    d_wd = torch.zeros(512, P, device=DEV)
    d_s = torch.zeros(7, device=DEV)
    r = nat.ReduceArgs()
    segs = [(wide, 37, 70001, 0, 70001, 1, d_wide.data_ptr(), 1),
            (tall, 1500, 7 * 512, 2 * 512, 512, 1, d_t0.data_ptr(), 1)]
    segs += [(tall, 1500, 7 * 512, (4 + j) * 512, 512, 1, d_wd.data_ptr() + 4 * j, P) for j in range(P)]
    segs.append((tall2, 129, 40, 3, 7, 5, d_s.data_ptr(), 1))          # source stride 5
    r.nseg = len(segs)
    for i, (part, n_part, ps, so, ln, ss, dp, ds) in enumerate(segs):
        r.seg_part[i], r.seg_n_part[i], r.seg_part_stride[i] = part.data_ptr(), n_part, ps
        r.src_off[i], r.len[i], r.src_stride[i], r.dst[i], r.dst_stride[i] = so, ln, ss, dp, ds
    for accumulate in (0, 1):
        r.accumulate = accumulate
        nat.check(nat.lib().ctn_reduce_partials(C.byref(r), nat.stream()))
        f = 1.0 + accumulate
        assert rel_l2(d_wide, f * wide.double().sum(0)) < 1e-6
        assert rel_l2(d_t0, f * tall.double().sum(0)[2 * 512:3 * 512]) < 1e-6
        for j in range(P):
            assert rel_l2(d_wd[:, j], f * tall.double().sum(0)[(4 + j) * 512:(5 + j) * 512]) < 1e-6
        assert rel_l2(d_s, f * tall2.double().sum(0)[3:3 + 7 * 5:5]) < 1e-6


def test_bn_fold():
    """ctn_bn_fold against nn.BatchNorm1d in eval mode (conv_tasnet.py:300-303)."""
    Cn = 515
    bn = torch.nn.BatchNorm1d(Cn).to(DEV)
    g = torch.Generator().manual_seed(61)
    with torch.no_grad():
        bn.weight.copy_(1 + 0.3 * torch.randn(Cn, generator=g))
        bn.bias.copy_(0.2 * torch.randn(Cn, generator=g))
        bn.running_mean.copy_(torch.randn(Cn, generator=g))
        bn.running_var.copy_(0.1 + torch.rand(Cn, generator=g))
    bn.eval()
    scale, shift = torch.empty(Cn, device=DEV), torch.empty(Cn, device=DEV)
    nat.check(nat.lib().ctn_bn_fold(bn.weight.data_ptr(), bn.bias.data_ptr(), bn.running_mean.data_ptr(),
                                    bn.running_var.data_ptr(), bn.eps, Cn, scale.data_ptr(), shift.data_ptr(),
                                    nat.stream()))
    x = torch.randn(3, Cn, 50, generator=g).to(DEV)
    with torch.no_grad():
        want = bn(x)
    got = x * scale[None, :, None] + shift[None, :, None]
    assert rel_l2(got, want) < 1e-6
