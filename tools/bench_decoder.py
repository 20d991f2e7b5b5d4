"""GPU micro-benchmark of ctn_decoder_fwd (paper shapes): python tools/bench_decoder.py [M]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200")]
import torch  # noqa: E402
from ctn_b200 import _native as nat  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 64
lib = nat.lib()
for (N, L, C, T, mask) in ((256, 20, 2, 32000, 0), (512, 16, 3, 32000, 1), (256, 16, 2, 32000, 0)):
    hop = L // 2
    K = (T - L) // hop + 1
    score = torch.randn(M * K, C * N, device="cuda")
    w = torch.rand(M * K, N, device="cuda")
    V = torch.randn(L, N, device="cuda") * 0.1
    est = torch.empty(M, C, T, device="cuda")
    args = (score.data_ptr(), w.data_ptr(), V.data_ptr(), M, K, N, L, C, T, mask, est.data_ptr())
    for _ in range(3):
        nat.check(lib.ctn_decoder_fwd(*args, nat.stream()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        lib.ctn_decoder_fwd(*args, nat.stream())
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1e3
    nbytes = 4.0 * M * K * N * (C + 1) + 4.0 * M * C * T
    mk = torch.softmax(score[:K].view(K, C, N), 1) if mask else torch.relu(score[:K].view(K, C, N))
    fr = torch.einsum("kcn,ln->kcl", (mk * w[:K, None, :]).double(), V.double())
    ref = torch.zeros(C, T, dtype=torch.float64, device="cuda")
    for k in range(0, K):
        pass
    idx = (torch.arange(K, device="cuda")[:, None] * hop + torch.arange(L, device="cuda")[None, :]).flatten()
    for c in range(C):
        ref[c].index_add_(0, idx, fr[:, c, :].flatten())
    err = float((est[0].double() - ref).norm() / ref.norm())
    print("N=%d L=%d C=%d M=%d: %.1f us  %.0f GB/s (algorithmic)  rel err %.2e" % (N, L, C, M, us, nbytes / us / 1e3, err))
