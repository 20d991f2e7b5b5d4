"""GPU micro-benchmark + parity of ctn_gemm1x1 for the two block GEMM shapes (paper config).
    python tools/bench_gemm.py [M] [modes...]   mode 1 = single-CTA kernel, 2 = CTA-pair kernel"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200"), os.path.join(ROOT, "tests")]
import torch  # noqa: E402
import kernels as Kn  # noqa: E402
from ctn_b200 import _native as nat  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 16
modes = [int(x) for x in sys.argv[2:]] or [1, 2]
K, B, H = 3199, 256, 512
dev = "cuda"
lib = nat.lib()
torch.manual_seed(0)
R = M * K


def make(kind, mode):
    g = nat.GemmArgs()
    g.M, g.K = M, K
    keep = {}
    if kind == "conv1x1":
        Cin, Cout = B, H
    else:   # "pointwise" = training form (fp32 residual in, fp32 + hi/lo out); "pw_inf" = inference form
        Cin, Cout = H, B
    A = torch.randn(R, Cin, device=dev)
    W = torch.randn(Cout, Cin, device=dev) * 0.05
    a_hi, a_lo = Kn.split_hl(A)
    w_hi, w_lo = Kn.split_hl(W)
    keep.update(a_hi=a_hi, a_lo=a_lo, w_hi=w_hi, w_lo=w_lo)
    g.Cin, g.Cout = Cin, Cout
    g.a_hi, g.a_lo, g.w_hi, g.w_lo = a_hi.data_ptr(), a_lo.data_ptr(), w_hi.data_ptr(), w_lo.data_ptr()
    out = torch.zeros(R, Cout, device=dev)
    keep["out"] = out
    g.out_f32 = out.data_ptr()
    ref = (a_hi.double() + a_lo.double())[:min(4, M) * K] @ (w_hi.double() + w_lo.double()).t()
    if kind == "conv1x1":
        st = torch.zeros(M, lib.ctn_gemm_gln_partials(K, Cout), 2, device=dev)
        sl = torch.tensor([0.25], device=dev)
        keep.update(st=st, sl=sl)
        g.out_stat, g.out_stat_buf, g.stat_prelu = nat.NORM_GLN, st.data_ptr(), sl.data_ptr()
    else:
        resid = torch.randn(R, Cout, device=dev)
        hi = torch.zeros(R, Cout, dtype=torch.bfloat16, device=dev)
        lo = torch.zeros_like(hi)
        P = 7
        ist = torch.rand(M, P, 2, device=dev) * 100 + 50
        ist[:, :, 1] = ist[:, :, 0] ** 2 / (K * Cin / P) * 3.0 + 10.0
        cg = torch.randn(Cout, device=dev)
        cb = torch.randn(Cout, device=dev)
        keep.update(resid=resid, hi=hi, lo=lo, ist=ist, cg=cg, cb=cb)
        if kind == "pw_inf":   # residual read from the (hi, lo) pair the result overwrites
            r_hi, r_lo = Kn.split_hl(resid)
            resid = r_hi.float() + r_lo.float()
            hi.copy_(r_hi), lo.copy_(r_lo)
            keep.update(r_hi=r_hi, r_lo=r_lo)
            g.out_f32 = 0
            g.resid_hi, g.resid_lo = r_hi.data_ptr(), r_lo.data_ptr()
            g.out_hi, g.out_lo = hi.data_ptr(), lo.data_ptr()
        else:
            g.resid, g.out_hi, g.out_lo = resid.data_ptr(), hi.data_ptr(), lo.data_ptr()
        g.col_g, g.col_bias = cg.data_ptr(), cb.data_ptr()
        g.in_norm, g.in_stat, g.in_P, g.in_count = nat.NORM_GLN, ist.data_ptr(), P, float(K) * Cin
        s = ist.double().sum(1)
        mean = s[:, 0] / (K * Cin)
        var = (s[:, 1] / (K * Cin) - mean * mean).clamp_min(0)
        r = 1.0 / torch.sqrt(var + 1e-8)
        mm = torch.arange(min(4, M), device=dev).repeat_interleave(K)
        ref = r[mm, None] * ref - (r * mean)[mm, None] * cg.double()[None] + cb.double()[None] + resid[:min(4, M) * K].double()
    g.force_simt = mode << 1
    return g, keep, ref


def bench(kind, mode, g, s, dbg, err, extra):
    if True:
        for _ in range(5):
            lib.ctn_gemm1x1(C.byref(g), s)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = NITER
        e0.record()
        for _ in range(n):
            lib.ctn_gemm1x1(C.byref(g), s)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / n * 1e3
        fl = 2.0 * R * g.Cin * g.Cout
        print("%-9s mode %d dbg %2d M=%d: %.1f us  %.1f TFLOP/s (algorithmic)  rel err %.2e%s" % (kind, mode, dbg, M, us, fl / us / 1e6, err, extra))


KINDS = os.environ.get("CTN_KINDS", "conv1x1,pointwise").split(",")
NITER = int(os.environ.get("CTN_NITER", "30"))
DBGS = [int(x) for x in os.environ.get("CTN_DBGS", "0").split(",")]
for kind in KINDS:
    for mode in modes:
        g, keep, ref = make(kind, mode)
        s = nat.stream()
        nat.check(lib.ctn_gemm1x1(C.byref(g), s))
        torch.cuda.synchronize()
        out = keep["out"][:min(4, M) * K].double()
        err = float((out - ref).norm() / ref.norm()) if kind != "pw_inf" else float("nan")
        extra = ""
        if kind != "conv1x1":
            hl = keep["hi"][:min(4, M) * K].double() + keep["lo"][:min(4, M) * K].double()
            extra = " hl %.2e" % float((hl - ref).norm() / ref.norm())
        else:
            x = torch.where(ref >= 0, ref, 0.25 * ref)
            st = keep["st"][:min(4, M)].double().sum(1)
            extra = " stat %.2e %.2e" % (float(((st[:, 0] - x.view(min(4, M), -1).sum(1)) / x.view(min(4, M), -1).abs().sum(1)).abs().max()),
                                          float(((st[:, 1] - (x * x).view(min(4, M), -1).sum(1)) / (x * x).view(min(4, M), -1).sum(1)).abs().max()))
        for dbg in DBGS:
            os.environ["CTN_GEMM_DBG"] = str(dbg)
            bench(kind, mode, g, s, dbg, err, extra)
        os.environ["CTN_GEMM_DBG"] = "0"

