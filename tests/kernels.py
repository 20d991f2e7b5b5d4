"""Thin torch-tensor wrappers around the C ABI for the per-kernel parity tests."""
import ctypes as C

import torch

from ctn_b200 import _native as nat


def split_hl(x):
    hi = x.to(torch.bfloat16)
    lo = (x - hi.float()).to(torch.bfloat16)
    return hi.contiguous(), lo.contiguous()


def weight_prep(W, gamma=None, beta=None, transpose=False):
    rows, cols = W.shape
    shp = (cols, rows) if transpose else (rows, cols)
    hi = torch.empty(shp, dtype=torch.bfloat16, device=W.device)
    lo = torch.empty_like(hi)
    g = torch.empty(rows, device=W.device)
    b = torch.empty(rows, device=W.device)
    nat.check(nat.lib().ctn_weight_prep(W.data_ptr(), rows, cols, nat.ptr(gamma), nat.ptr(beta), int(transpose),
                                        hi.data_ptr(), lo.data_ptr(), g.data_ptr(), b.data_ptr(), nat.stream()))
    return hi, lo, g, b


def encoder(mix, U):
    M, T = mix.shape
    N, _, L = U.shape
    K = (T - L) // (L // 2) + 1
    dev = mix.device
    w = torch.empty(M * K, N, device=dev)
    hi = torch.empty(M * K, N, dtype=torch.bfloat16, device=dev)
    lo = torch.empty_like(hi)
    st = torch.empty(M * K, 2, device=dev)
    nat.check(nat.lib().ctn_encoder_fwd(mix.data_ptr(), U.data_ptr(), M, T, N, L, w.data_ptr(), hi.data_ptr(),
                                        lo.data_ptr(), st.data_ptr(), nat.stream()))
    return w, hi, lo, st, K


def gemm(M, K, a_hi, a_lo, w_hi, w_lo, out_f32=True, out_hl=False, resid=None, col_g=None, col_bias=None,
         in_norm=0, in_stat=None, in_P=0, in_count=0.0, out_stat=0, stat_prelu=None, force_simt=False, store_prelu=False):
    R, Cin = a_hi.shape
    Cout = w_hi.shape[0]
    dev = a_hi.device
    lib = nat.lib()
    g = nat.GemmArgs()
    g.M, g.K, g.Cin, g.Cout = M, K, Cin, Cout
    g.a_hi, g.a_lo, g.w_hi, g.w_lo = a_hi.data_ptr(), a_lo.data_ptr(), w_hi.data_ptr(), w_lo.data_ptr()
    res = {}
    if out_f32:
        res["out"] = torch.full((R, Cout), float("nan"), device=dev)
        g.out_f32 = res["out"].data_ptr()
    if out_hl:
        res["hi"] = torch.zeros(R, Cout, dtype=torch.bfloat16, device=dev)
        res["lo"] = torch.zeros_like(res["hi"])
        g.out_hi, g.out_lo = res["hi"].data_ptr(), res["lo"].data_ptr()
    g.resid, g.col_g, g.col_bias = nat.ptr(resid), nat.ptr(col_g), nat.ptr(col_bias)
    g.in_norm, g.in_stat, g.in_P, g.in_count = in_norm, nat.ptr(in_stat), in_P, in_count
    g.out_stat = out_stat
    if out_stat == nat.NORM_GLN:
        res["stat"] = torch.full((M, lib.ctn_gemm_gln_partials(K, Cout), 2), float("nan"), device=dev)
    elif out_stat == nat.NORM_CLN:
        res["stat"] = torch.full((R, lib.ctn_gemm_cln_groups(Cout), 2), float("nan"), device=dev)
    if out_stat:
        g.out_stat_buf = res["stat"].data_ptr()
    g.stat_prelu = nat.ptr(stat_prelu)
    g.store_prelu = int(store_prelu)
    g.force_simt = int(force_simt)
    nat.check(lib.ctn_gemm1x1(C.byref(g), nat.stream()))
    return res


def dwconv(M, K, a, a1, g1, b1, norm, in_stat, in_P, wd, a2, dilation, causal, want_pre=True):
    R, H = a.shape
    P = wd.shape[-1]
    dev = a.device
    lib = nat.lib()
    d = nat.DwconvArgs()
    d.M, d.K, d.H, d.P, d.dilation, d.causal = M, K, H, P, dilation, int(causal)
    d.inp, d.prelu_in, d.gamma_in, d.beta_in = a.data_ptr(), a1.data_ptr(), g1.data_ptr(), b1.data_ptr()
    d.norm, d.in_stat, d.in_P = norm, in_stat.data_ptr(), in_P
    d.wd, d.prelu_out = wd.data_ptr(), a2.data_ptr()
    pre = torch.full((R, H), float("nan"), device=dev) if want_pre else None
    hi = torch.zeros(R, H, dtype=torch.bfloat16, device=dev)
    lo = torch.zeros_like(hi)
    if norm == nat.NORM_GLN:
        st = torch.full((M, lib.ctn_dwconv_gln_partials(K, H, P, dilation), 2), float("nan"), device=dev)
    else:
        st = torch.full((R, 2), float("nan"), device=dev)
    d.out_pre, d.out_hi, d.out_lo, d.out_stat = nat.ptr(pre), hi.data_ptr(), lo.data_ptr(), st.data_ptr()
    nat.check(lib.ctn_dwconv_fwd(C.byref(d), nat.stream()))
    return pre, hi, lo, st


def dwpw(M, K, a, a1, g1, b1, in_stat, in_P, wd, a2, w_hi, w_lo, dilation, causal, want_d=False):
    """ctn_dwpw_fwd: returns (y [M*K, B] raw product, gLN-2 partials [M, P, 2]) (+ the depthwise output before the
    second PReLU with want_d, the training forward's extra output)"""
    R, H = a.shape
    B = w_hi.shape[0]
    dev = a.device
    lib = nat.lib()
    d = nat.DwpwArgs()
    d.M, d.K, d.H, d.B, d.dilation, d.causal = M, K, H, B, dilation, int(causal)
    d.a, d.prelu_in, d.gamma_in, d.beta_in = a.data_ptr(), nat.ptr(a1), g1.data_ptr(), b1.data_ptr()
    d.in_stat, d.in_P = in_stat.data_ptr(), in_P
    d.wd, d.prelu_out, d.w_hi, d.w_lo = wd.data_ptr(), a2.data_ptr(), w_hi.data_ptr(), w_lo.data_ptr()
    y = torch.full((R, B), float("nan"), device=dev)
    st = torch.full((M, lib.ctn_dwpw_partials(K, dilation), 2), float("nan"), device=dev)
    d.y, d.out_stat = y.data_ptr(), st.data_ptr()
    dout = torch.full((R, H), float("nan"), device=dev) if want_d else None
    d.d_out = nat.ptr(dout)
    nat.check(lib.ctn_dwpw_fwd(C.byref(d), nat.stream()))
    return (y, st, dout) if want_d else (y, st)


def resid_update(x_hi, x_lo, y, stat, count, col_g, col_bias, M, K):
    B = y.shape[1]
    nat.check(nat.lib().ctn_resid_update(x_hi.data_ptr(), x_lo.data_ptr(), y.data_ptr(), stat.data_ptr(), stat.shape[1],
                                         float(count), col_g.data_ptr(), col_bias.data_ptr(), M, K, B, nat.stream()))


def updc1(M, K, x_hi, x_lo, y, stat, count, col_g, col_bias, w_hi, w_lo, slope, x_out=None, store_raw=False):
    """ctn_updc1_fwd: updates (x_hi, x_lo) in place (or writes the pair x_out, leaving the input intact); returns
    (a [M*K, H] = PReLU(W1 x), or W1 x with store_raw, gLN-1 partials)"""
    B = y.shape[1]
    H = w_hi.shape[0]
    lib = nat.lib()
    u = nat.Updc1Args()
    u.M, u.K, u.B, u.H = M, K, B, H
    u.x_hi, u.x_lo, u.y = x_hi.data_ptr(), x_lo.data_ptr(), y.data_ptr()
    u.in_stat, u.in_P, u.in_count = stat.data_ptr(), stat.shape[1], float(count)
    u.col_g, u.col_bias, u.w_hi, u.w_lo = col_g.data_ptr(), col_bias.data_ptr(), w_hi.data_ptr(), w_lo.data_ptr()
    a = torch.full((M * K, H), float("nan"), device=y.device)
    st = torch.full((M, lib.ctn_gemm_gln_partials(K, H), 2), float("nan"), device=y.device)
    u.a, u.out_stat, u.slope = a.data_ptr(), st.data_ptr(), slope.data_ptr()
    if x_out is not None:
        u.x_out_hi, u.x_out_lo = x_out[0].data_ptr(), x_out[1].data_ptr()
    u.store_raw = int(store_raw)
    nat.check(lib.ctn_updc1_fwd(C.byref(u), nat.stream()))
    return a, st


def decoder(score, w, V, M, K, C_, T, mask):
    L, N = V.shape
    est = torch.full((M, C_, T), float("nan"), device=score.device)
    nat.check(nat.lib().ctn_decoder_fwd(score.data_ptr(), w.data_ptr(), V.data_ptr(), M, K, N, L, C_, T, mask,
                                        est.data_ptr(), nat.stream()))
    return est


def to_rows(x):
    """[M, C, K] -> frames-major [M*K, C]"""
    M, C_, K = x.shape
    return x.permute(0, 2, 1).reshape(M * K, C_).contiguous()


def from_rows(r, M):
    """[M*K, C] -> [M, C, K]"""
    R, C_ = r.shape
    return r.view(M, R // M, C_).permute(0, 2, 1).contiguous()
