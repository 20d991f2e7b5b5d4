"""Experiment: one training forward + backward of a batch of M utterances as ONE plan of M against M one-utterance
plans on M streams (independent utterances: gLN statistics are per utterance), paper configuration.
    python tools/exp_split_train.py [M]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "conv-tasnet_b200")]
import torch  # noqa: E402
import conv_tasnet  # noqa: E402
from ctn_b200._engine import aligned_params  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 3
T = 32000
cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
torch.manual_seed(0)
dev = torch.device("cuda:0")
model = conv_tasnet.ConvTasNet(**cfg).to(dev).train()
eng = model.train_engine()
params = {k: p.detach() for k, p in zip(model._keys, [dict(model.named_parameters())[k] for k in model._keys])}
params = aligned_params(eng._staged, params, dev)
mix = 0.1 * torch.randn(M, T, device=dev)
g = 1e-3 * torch.randn(M, cfg["C"], T, device=dev)


def timeit(fn, n=20):
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


planM = eng._plan(params, M, T, dev, slot=0)
planM.mix.copy_(mix)
planM.g_est.copy_(g)


def step_batch():
    planM.run_forward(True)
    planM.run_backward(True)


print("batch of %d as one plan: %.3f ms" % (M, timeit(step_batch)), flush=True)
ref = planM.flat_grad.clone()

plans = [eng._plan(params, 1, T, dev, slot=i) for i in range(M)]
for i, p in enumerate(plans):
    p.mix.copy_(mix[i:i + 1])
    p.g_est.copy_(g[i:i + 1])
streams = [torch.cuda.Stream(device=dev) for _ in range(M)]


def step_split():
    main = torch.cuda.current_stream(dev)
    ev = main.record_event()
    for st, p in zip(streams, plans):
        st.wait_event(ev)
        with torch.cuda.stream(st):
            p.run_forward(True)
            p.run_backward(True)
    for st in streams:
        main.wait_stream(st)


print("%d one-utterance plans on %d streams: %.3f ms" % (M, M, timeit(step_split)), flush=True)
tot = sum(p.flat_grad.double() for p in plans)
print("gradient rel-L2 split vs batch: %.2e" % float((tot - ref.double()).norm() / ref.double().norm()))
one = plans[0]
print("one utterance alone: %.3f ms" % timeit(lambda: (one.run_forward(True), one.run_backward(True))))
