"""GPU micro-benchmark of ctn_updc1_fwd (residual update + conv1x1) against ctn_resid_update + ctn_gemm1x1, paper
shapes, with the kernel's ablation bits (CTN_C1_DBG).   python tools/bench_updc1.py [M]   env CTN_DBGS=0,1,2,4,8"""
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
x_hi = torch.randn(2 * R, B, device=dev).to(torch.bfloat16)
x_lo = (torch.randn(2 * R, B, device=dev) * 1e-3).to(torch.bfloat16)
y = torch.randn(R, B, device=dev) * 0.01
g2, b2 = 1 + 0.3 * torch.randn(H, device=dev), torch.randn(H, device=dev)
W2 = torch.randn(B, H, device=dev) * 0.05
_, _, col_g, col_bias = Kn.weight_prep(W2, g2, b2)
W1 = torch.randn(H, B, device=dev) * 0.05
w_hi, w_lo, _, _ = Kn.weight_prep(W1)
P = 300
stat = torch.rand(M, P, 2, device=dev) * 100 + 50
count = float(K) * H
stat[:, :, 1] = stat[:, :, 0] ** 2 / (count / P) * 3.0 + 10.0
slope = torch.tensor([0.25], device=dev)
a = torch.empty(R, H, device=dev)
st1 = torch.zeros(M, lib.ctn_gemm_gln_partials(K, H), 2, device=dev)
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


u = nat.Updc1Args()
u.M, u.K, u.B, u.H = M, K, B, H
u.x_hi, u.x_lo, u.y = x_hi.data_ptr(), x_lo.data_ptr(), y.data_ptr()
u.in_stat, u.in_P, u.in_count = stat.data_ptr(), P, count
u.col_g, u.col_bias, u.w_hi, u.w_lo = col_g.data_ptr(), col_bias.data_ptr(), w_hi.data_ptr(), w_lo.data_ptr()
u.a, u.out_stat, u.slope = a.data_ptr(), st1.data_ptr(), slope.data_ptr()
for dbg in DBGS:
    os.environ["CTN_C1_DBG"] = str(dbg)
    us = timeit(lambda: nat.check(lib.ctn_updc1_fwd(C.byref(u), s)))
    print("updc1  dbg %2d M=%d: %7.1f us  (%.0f GB/s of 20 B/elt)" % (dbg, M, us, R * B * 20 / us / 1e3), flush=True)
os.environ["CTN_C1_DBG"] = "0"
us_u = timeit(lambda: nat.check(lib.ctn_resid_update(x_hi.data_ptr(), x_lo.data_ptr(), y.data_ptr(), stat.data_ptr(), P, count,
                                                     col_g.data_ptr(), col_bias.data_ptr(), M, K, B, s)))
g = nat.GemmArgs()
g.M, g.K, g.Cin, g.Cout = M, K, B, H
g.a_hi, g.a_lo, g.w_hi, g.w_lo = x_hi.data_ptr(), x_lo.data_ptr(), w_hi.data_ptr(), w_lo.data_ptr()
g.out_f32, g.out_stat, g.out_stat_buf, g.stat_prelu, g.store_prelu = a.data_ptr(), nat.NORM_GLN, st1.data_ptr(), slope.data_ptr(), 1
us_g = timeit(lambda: nat.check(lib.ctn_gemm1x1(C.byref(g), s)))
print("unfused M=%d: update %7.1f + conv1x1 %7.1f = %7.1f us" % (M, us_u, us_g, us_u + us_g), flush=True)

if os.environ.get("CTN_TRACE") == "1":
    import numpy as np
    for dbg in DBGS:
        os.environ["CTN_C1_DBG"] = str(dbg)
        nat.check(lib.ctn_updc1_fwd(C.byref(u), s))
        torch.cuda.synchronize()
        buf = np.zeros(64 * 32, dtype=np.int64)
        lib.ctn_debug_trace_read.argtypes = [C.c_void_p, C.c_int]
        lib.ctn_debug_trace_read(buf.ctypes.data, buf.size)
        t = buf.reshape(64, 32)
        t0 = t[0][t[0] > 0].min()
        names = ["ld0", "ld1", "ld2", "ld3", "raw0", "raw1", "raw2", "raw3", "fix0", "fix1", "fix2", "fix3", "st0", "st1", "st2",
                 "st3", "m0", "m1", "m2", "m3", "m4", "m5", "m6", "m7", "d0s", "d0e", "d1s", "d1e", "entry", "exit"]
        print("trace dbg", dbg, "(cycles since the first event, CTA 0, row tiles 0-11)")
        for it in range(12):
            print(it, " ".join("%s=%d" % (n, (t[it][i] - t0) // 10 * 10) for i, n in enumerate(names) if t[it][i] > 0))
