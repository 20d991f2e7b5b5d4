"""Diagnostic (not a pytest): structured inputs through the tcgen05 GEMM to localise
descriptor / swizzle / pipeline mistakes.  Usage: python tests/gpu_diag_gemm.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200"), os.path.join(ROOT, "tests")]
import torch  # noqa: E402
import kernels as Kn  # noqa: E402


def report(tag, out, ref):
    err = (out.double() - ref.double())
    rel = float(err.norm() / ref.double().norm().clamp_min(1e-30))
    print("%-34s rel %.3e  max|err| %.3e  nan %d" % (tag, rel, float(err.abs().max()), int(torch.isnan(out).sum())))
    if rel > 1e-4:
        R, Cc = out.shape
        e = err.abs()
        rows = e.amax(1)
        cols = e.amax(0)
        bad_r = (rows > 1e-3).nonzero().flatten()
        bad_c = (cols > 1e-3).nonzero().flatten()
        print("   bad rows %d/%d first %s ; bad cols %d/%d first %s" % (
            len(bad_r), R, bad_r[:12].tolist(), len(bad_c), Cc, bad_c[:12].tolist()))
        print("   out[0,:8]", out[0, :8].tolist())
        print("   ref[0,:8]", ref[0, :8].tolist())


def main():
    dev = "cuda"
    torch.manual_seed(0)
    for (M, K, Cin, Cout) in [(1, 128, 64, 64), (1, 128, 64, 128), (1, 128, 128, 128), (1, 256, 256, 256),
                              (1, 300, 256, 512), (4, 3199, 256, 512), (40, 3199, 256, 512), (40, 3199, 512, 256)]:
        R = M * K
        A = torch.randn(R, Cin, device=dev)
        W = torch.randn(Cout, Cin, device=dev) * 0.1
        a_hi, a_lo = Kn.split_hl(A)
        w_hi, w_lo = Kn.split_hl(W)
        zero_a, zero_w = torch.zeros_like(a_lo), torch.zeros_like(w_lo)
        tag = "M%d K%d %dx%d" % (M, K, Cin, Cout)
        ref_hh = a_hi.double() @ w_hi.double().t()
        out = Kn.gemm(M, K, a_hi, zero_a, w_hi, zero_w)["out"]
        report(tag + " hi*hi only", out, ref_hh)
        ref = (a_hi.double() + a_lo.double()) @ (w_hi.double() + w_lo.double()).t()
        out = Kn.gemm(M, K, a_hi, a_lo, w_hi, w_lo)["out"]
        report(tag + " full split", out, ref)
        out = Kn.gemm(M, K, a_hi, a_lo, w_hi, w_lo, force_simt=True)["out"] if R * Cin * Cout < 2e9 else None
        if out is not None:
            report(tag + " simt", out, ref)
    torch.cuda.synchronize()
    print("diag done")


if __name__ == "__main__":
    main()
