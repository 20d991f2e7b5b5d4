"""GPU micro-benchmark of the fused depthwise + pointwise kernel (ctn_dwpw_fwd) and of
ctn_resid_update against the unfused pair (ctn_dwconv_fwd + ctn_gemm1x1), paper shapes.
    python tools/bench_fused.py [M]      env CTN_DBGS=0,1,2,4  CTN_DILS=1,8,128"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200"), os.path.join(ROOT, "tests")]
import torch  # noqa: E402
import kernels as Kn  # noqa: E402
from ctn_b200 import _native as nat  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 64
K, B, H = 3199, 256, 512
dev = "cuda"
lib = nat.lib()
torch.manual_seed(0)
R = M * K
NITER = int(os.environ.get("CTN_NITER", "20"))
DBGS = [int(x) for x in os.environ.get("CTN_DBGS", "0").split(",")]
DILS = [int(x) for x in os.environ.get("CTN_DILS", "1,8,128").split(",")]

a = torch.randn(R, H, device=dev)
a1 = torch.tensor([0.25], device=dev)
a2 = torch.tensor([0.1], device=dev)
g1, b1 = 1 + 0.3 * torch.randn(H, device=dev), torch.randn(H, device=dev)
g2, b2 = 1 + 0.3 * torch.randn(H, device=dev), torch.randn(H, device=dev)
wd = torch.randn(H, 1, 3, device=dev) * 0.5
W2 = torch.randn(B, H, device=dev) * 0.05
w_hi, w_lo, col_g, col_bias = Kn.weight_prep(W2, g2, b2)
pa = torch.where(a >= 0, a, 0.25 * a).view(M, K * H)
in_P = 4
parts = torch.zeros(M, in_P, 2, device=dev)
parts[:, 0, 0] = pa.double().sum(1).float()
parts[:, 0, 1] = (pa.double() ** 2).sum(1).float()
del pa
x_hi = torch.randn(R, B, device=dev).to(torch.bfloat16)
x_lo = (torch.randn(R, B, device=dev) * 1e-3).to(torch.bfloat16)
y = torch.zeros(R, B, device=dev)
s = nat.stream()


def timeit(fn):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(NITER):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / NITER * 1e3


for dil in DILS:
    P2 = lib.ctn_dwpw_partials(K, dil)
    st = torch.zeros(M, P2, 2, device=dev)
    d = nat.DwpwArgs()
    d.M, d.K, d.H, d.B, d.dilation, d.causal = M, K, H, B, dil, 0
    d.a, d.prelu_in, d.gamma_in, d.beta_in = a.data_ptr(), a1.data_ptr(), g1.data_ptr(), b1.data_ptr()
    d.in_stat, d.in_P = parts.data_ptr(), in_P
    d.wd, d.prelu_out, d.w_hi, d.w_lo = wd.data_ptr(), a2.data_ptr(), w_hi.data_ptr(), w_lo.data_ptr()
    d.y, d.out_stat = y.data_ptr(), st.data_ptr()
    for dbg in DBGS:
        os.environ["CTN_GEMM_DBG"] = str(dbg)
        us = timeit(lambda: nat.check(lib.ctn_dwpw_fwd(C.byref(d), s)))
        print("dwpw   dil %3d dbg %d M=%d: %7.1f us  (%.2f us/utt)  tiles/utt %d" % (dil, dbg, M, us, us / M, P2 // 8), flush=True)
    os.environ["CTN_GEMM_DBG"] = "0"
    us = timeit(lambda: nat.check(lib.ctn_resid_update(x_hi.data_ptr(), x_lo.data_ptr(), y.data_ptr(), st.data_ptr(), P2,
                                                       float(K) * H, col_g.data_ptr(), col_bias.data_ptr(), M, K, B, s)))
    print("update dil %3d       M=%d: %7.1f us  (%.2f us/utt)  %.0f GB/s" % (dil, M, us, us / M, R * B * 12 / us / 1e3), flush=True)
    # unfused pair
    q_hi = torch.zeros(R, H, dtype=torch.bfloat16, device=dev)
    q_lo = torch.zeros_like(q_hi)
    Pd = lib.ctn_dwconv_gln_partials(K, H, 3, dil)
    st2 = torch.zeros(M, Pd, 2, device=dev)
    dd = nat.DwconvArgs()
    dd.M, dd.K, dd.H, dd.P, dd.dilation, dd.causal = M, K, H, 3, dil, 0
    dd.inp, dd.prelu_in, dd.gamma_in, dd.beta_in = a.data_ptr(), a1.data_ptr(), g1.data_ptr(), b1.data_ptr()
    dd.norm, dd.in_stat, dd.in_P = nat.NORM_GLN, parts.data_ptr(), in_P
    dd.wd, dd.prelu_out = wd.data_ptr(), a2.data_ptr()
    dd.out_pre, dd.out_hi, dd.out_lo, dd.out_stat = None, q_hi.data_ptr(), q_lo.data_ptr(), st2.data_ptr()
    us_d = timeit(lambda: nat.check(lib.ctn_dwconv_fwd(C.byref(dd), s)))
    g = nat.GemmArgs()
    g.M, g.K, g.Cin, g.Cout = M, K, H, B
    g.a_hi, g.a_lo, g.w_hi, g.w_lo = q_hi.data_ptr(), q_lo.data_ptr(), w_hi.data_ptr(), w_lo.data_ptr()
    g.col_g, g.col_bias = col_g.data_ptr(), col_bias.data_ptr()
    g.in_norm, g.in_stat, g.in_P, g.in_count = nat.NORM_GLN, st2.data_ptr(), Pd, float(K) * H
    g.resid_hi, g.resid_lo, g.out_hi, g.out_lo = x_hi.data_ptr(), x_lo.data_ptr(), x_hi.data_ptr(), x_lo.data_ptr()
    us_g = timeit(lambda: nat.check(lib.ctn_gemm1x1(C.byref(g), s)))
    print("unfused dil %3d      M=%d: dwconv %7.1f + pointwise %7.1f = %7.1f us" % (dil, M, us_d, us_g, us_d + us_g), flush=True)
    del q_hi, q_lo
