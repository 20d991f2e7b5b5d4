"""Forward latency of small batches (CUDA graph) with the deferred update folded into the conv1x1 or not."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200")]
import torch
import conv_tasnet
cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
torch.manual_seed(0)
model = conv_tasnet.ConvTasNet(**cfg).cuda().eval()
eng = model.engine()
eng.use_graph = True
for M in (1, 2, 3, 4, 8, 16):
    mix = 0.1 * torch.randn(M, 32000, device="cuda")
    eng.micro_batch = M
    for fu in (0, 1):
        eng.fuse_update = bool(fu)
        with torch.no_grad():
            for _ in range(5):
                model(mix)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(50):
                model(mix)
            e1.record()
            torch.cuda.synchronize()
        print("M=%2d fuse_update=%d: %.3f ms per forward (%d launches)" % (M, fu, e0.elapsed_time(e1) / 50, eng.num_launches()), flush=True)
