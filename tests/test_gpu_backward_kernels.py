"""Per-kernel parity of the non-GEMM backward kernels of one TemporalBlock
(csrc/backward.cu) against fp64 autograd over the oracle's functions
(conv_tasnet.py:212-272 differentiated by torch, as the reference does)."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

from conftest import rel_l2
from oracle import convtasnet_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _stat(t):
    """(sum, sumsq) per utterance as ONE partial: [M, 1, 2] fp32."""
    s = t.sum(dim=(1, 2))
    ss = (t * t).sum(dim=(1, 2))
    return torch.stack([s, ss], -1).view(-1, 1, 2).float().contiguous()


@pytest.mark.parametrize("H,K,M", [(3, 5, 2), (5, 37, 1), (16, 24, 2), (136, 333, 2), (512, 700, 2)])
@pytest.mark.parametrize("dil", [1, 2, 4, 128])
@pytest.mark.parametrize("causal", [False, True])
def test_block_backward_chain(H, K, M, dil, causal):
    import kernels as Kn
    from ctn_b200 import _native as nat
    lib = nat.lib()
    P = 3
    g = torch.Generator().manual_seed(H * 1000 + K + dil)
    rnd = lambda *s: torch.randn(*s, generator=g, dtype=torch.float64).to(DEV)
    a = rnd(M, H, K).requires_grad_(True)
    g1, b1, g2, b2 = (rnd(1, H, 1).requires_grad_(True) for _ in range(4))
    wd = (0.5 * rnd(H, 1, P)).requires_grad_(True)
    a1 = torch.tensor([0.25], dtype=torch.float64, device=DEV, requires_grad=True)
    a2 = torch.tensor([0.1], dtype=torch.float64, device=DEV, requires_grad=True)
    du = rnd(M, H, K)
    pa = F.prelu(a, a1)
    n1 = O.gln(pa, g1, b1)
    n1.retain_grad()
    d = O.depthwise(n1, wd, dil, causal)
    q = F.prelu(d, a2)
    u = O.gln(q, g2, b2)
    u.backward(du)

    f = lambda t: t.detach().float().contiguous()
    rows = lambda t: Kn.to_rows(t.detach().float())
    n = float(K * H)
    mu2 = q.mean(dim=(1, 2), keepdim=True)
    qh = (q - mu2) / torch.sqrt(((q - mu2) ** 2).mean(dim=(1, 2), keepdim=True) + O.EPS)
    gd = g2 * du
    bstat = torch.stack([gd.sum(dim=(1, 2)), (gd * qh).sum(dim=(1, 2))], -1).view(M, 1, 2).float().contiguous()
    stat1, stat2 = _stat(pa.detach()), _stat(q.detach())
    a_r, d_r, du_r = rows(a), rows(d), rows(du)
    tiles, rows = C.c_int(), C.c_int()
    lib.ctn_bwd_dw_layout(M, K, H, P, dil, nat.NORM_GLN, C.byref(tiles), C.byref(rows))
    tiles, rows = tiles.value, rows.value
    dn1 = torch.full((M * K, H), float("nan"), device=DEV)
    chan = torch.full((M * rows, (4 + P) * H), float("nan"), device=DEV)
    scal = torch.full((M * tiles, 4), float("nan"), device=DEV)
    prm = [f(t).view(-1) for t in (g1, b1, g2)]
    wd_f, a1_f, a2_f = f(wd).view(H, P), f(a1), f(a2)
    b = nat.BwdDwArgs()
    b.M, b.K, b.H, b.P, b.dilation, b.causal = M, K, H, P, dil, int(causal)
    b.P1, b.P2, b.Pb = 1, 1, 1
    b.du, b.d, b.a = du_r.data_ptr(), d_r.data_ptr(), a_r.data_ptr()
    b.stat1, b.stat2, b.bstat = stat1.data_ptr(), stat2.data_ptr(), bstat.data_ptr()
    b.g1, b.b1, b.g2 = (t.data_ptr() for t in prm)
    b.wd, b.slope1, b.slope2 = wd_f.data_ptr(), a1_f.data_ptr(), a2_f.data_ptr()
    b.dn1, b.chan_part, b.scal_part = dn1.data_ptr(), chan.data_ptr(), scal.data_ptr()
    # where the marching kernel runs it can also write u = gLN-2(prelu(d)) (hi, lo), the operand of dW2
    emits_u = bool(lib.ctn_bwd_dw_emits_u(H, P, nat.NORM_GLN))
    assert emits_u == (H % 4 == 0)
    if emits_u:
        b2_f = f(b2).view(-1)
        u_hi = torch.full((M * K, H), float("nan"), dtype=torch.bfloat16, device=DEV)
        u_lo = torch.full_like(u_hi, float("nan"))
        b.b2, b.u_hi, b.u_lo = b2_f.data_ptr(), u_hi.data_ptr(), u_lo.data_ptr()
    nat.check(lib.ctn_bwd_dwconv(C.byref(b), nat.stream()))
    assert not torch.isnan(dn1).any() and not torch.isnan(chan).any() and not torch.isnan(scal).any()
    if emits_u:
        assert rel_l2(u_hi.double() + u_lo.double(), Kn.to_rows(u.detach().float())) < 2e-5
    tol = 2e-4
    assert rel_l2(dn1, Kn.to_rows(n1.grad.float())) < tol
    cs = chan.double().sum(0).view(4 + P, H)
    assert rel_l2(cs[0], g2.grad.view(-1)) < tol
    assert rel_l2(cs[1], b2.grad.view(-1)) < tol
    assert rel_l2(cs[2], g1.grad.view(-1)) < tol
    assert rel_l2(cs[3], b1.grad.view(-1)) < tol
    assert rel_l2(cs[4:].t(), wd.grad.view(H, P)) < tol
    assert rel_l2(scal.double().sum(0)[2], a2.grad) < 5 * tol

    tiles_pn = lib.ctn_bwd_prelu_norm_tiles(K)
    da_hi = torch.zeros(M * K, H, dtype=torch.bfloat16, device=DEV)
    da_lo = torch.zeros_like(da_hi)
    alpha = torch.full((M * tiles_pn,), float("nan"), device=DEV)
    nat.check(lib.ctn_bwd_prelu_norm(dn1.data_ptr(), a_r.data_ptr(), stat1.data_ptr(), 1, scal.data_ptr(), tiles,
                                     prm[0].data_ptr(), a1_f.data_ptr(), M, K, H, da_hi.data_ptr(), da_lo.data_ptr(),
                                     alpha.data_ptr(), nat.stream()))
    assert rel_l2(da_hi.double() + da_lo.double(), Kn.to_rows(a.grad.float())) < tol
    assert rel_l2(alpha.double().sum(), a1.grad) < 5 * tol


def test_u_output_is_refused_where_the_kernel_cannot_write_it():
    """cLN blocks take the row-per-thread kernel, which has no u output: asking for one is an error, not a no-op."""
    from ctn_b200 import _native as nat
    lib = nat.lib()
    M, K, H, P = 1, 8, 8, 3
    assert lib.ctn_bwd_dw_emits_u(H, P, nat.NORM_CLN) == 0 and lib.ctn_bwd_dw_emits_u(H, 5, nat.NORM_GLN) == 0
    z = lambda *s: torch.zeros(*s, device=DEV)
    bufs = dict(du=z(M * K, H), d=z(M * K, H), a=z(M * K, H), stat1=z(M * K, 1, 2), stat2=z(M * K, 2), bstat=z(M * K, 1, 2),
                g1=z(H), b1=z(H), g2=z(H), wd=z(H, P), slope1=z(1), slope2=z(1), dn1=z(M * K, H))
    tiles, rows = C.c_int(), C.c_int()
    lib.ctn_bwd_dw_layout(M, K, H, P, 1, nat.NORM_CLN, C.byref(tiles), C.byref(rows))
    bufs["chan_part"], bufs["scal_part"] = z(M * rows.value, (4 + P) * H), z(M * tiles.value, 4)
    u = torch.zeros(M * K, H, dtype=torch.bfloat16, device=DEV)
    b = nat.BwdDwArgs()
    b.M, b.K, b.H, b.P, b.dilation, b.causal, b.norm = M, K, H, P, 1, 0, nat.NORM_CLN
    b.P1, b.P2, b.Pb = 1, 1, 1
    for k, v in bufs.items():
        setattr(b, k, v.data_ptr())
    assert lib.ctn_bwd_dwconv(C.byref(b), nat.stream()) == 0
    b.b2, b.u_hi, b.u_lo = bufs["g2"].data_ptr(), u.data_ptr(), u.data_ptr()
    assert lib.ctn_bwd_dwconv(C.byref(b), nat.stream()) != 0
    assert b"ctn_bwd_dw_emits_u" in lib.ctn_last_error()
