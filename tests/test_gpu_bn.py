"""norm_type="BN": the reference's chose_norm falls through to nn.BatchNorm1d (conv_tasnet.py:300-303).
Inference (running statistics) runs through the same fused kernels as gLN with the statistics folded
into a per-channel affine (ctn_bn_fold); training mode adds the batch statistics, the buffer update and
the batch-mean terms of the backward (csrc/bn_train.cu)."""
import pytest
import torch

from conftest import load_golden, rel_l2
from oracle import convtasnet_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _build(cfg, weights):
    import conv_tasnet
    model = conv_tasnet.ConvTasNet(**cfg)
    model.load_state_dict(weights)
    return model.to(DEV).eval()


@pytest.mark.parametrize("name", ["tc_bn_eval", "bn_eval_causal"])
def test_bn_inference_matches_reference_fixture(name):
    import pit_criterion
    cfg, z = load_golden(name)
    w = {k[2:]: v for k, v in z.items() if k.startswith("w:")}
    model = _build(cfg, w)
    assert list(model.state_dict()) == list(w)          # same keys, same order, buffers included
    with torch.no_grad():
        est = model(z["mixture"].to(DEV))
    err = rel_l2(est.cpu(), z["est"])
    print(name, "rel-L2", err)
    assert err < 1e-3
    loss, max_snr, _, _ = pit_criterion.cal_loss(z["sources"].to(DEV), est, z["lengths"].to(DEV))
    assert abs(float(loss) - float(z["loss"])) < 0.05
    _, _, idx = pit_criterion.cal_si_snr_with_pit(z["sources"].to(DEV), est.clone(), z["lengths"].to(DEV))
    assert torch.equal(idx.cpu(), z["idx"])


def test_bn_paper_size_fused_and_unfused_match_oracle():
    cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="BN", causal=False, mask_nonlinear="relu")
    p32 = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=5, norm_type="BN")
    model = _build(cfg, p32)
    g = torch.Generator().manual_seed(6)
    mix = (0.1 * torch.randn(2, 32000, generator=g)).to(DEV)
    eng = model.engine()
    with torch.no_grad():
        fused = model(mix)
        assert all(eng._plan(model._params(), 2, 32000, mix.device).fused)
        eng.fuse_max_dilation = 0
        unfused = model(mix)
        ref = O.forward(mix.double(), {k: v.to(DEV).double() for k, v in p32.items()}, cfg)
    print("BN paper size: fused", rel_l2(fused, ref), "unfused", rel_l2(unfused, ref))
    assert rel_l2(fused, ref) < 1e-3 and rel_l2(unfused, ref) < 1e-3
    # updated running statistics are picked up (the engine re-folds when a buffer changes)
    with torch.no_grad():
        for k, b in model.named_buffers():
            if k.endswith("running_mean"):
                b.add_(0.05)
        shifted = model(mix)
        p2 = {k: v.to(DEV).double() for k, v in model.state_dict().items()}
        ref2 = O.forward(mix.double(), p2, cfg)
    assert rel_l2(shifted, ref2) < 1e-3 and rel_l2(shifted, unfused) > 1e-3


def test_bn_streaming_is_refused():
    cfg, z = load_golden("bn_eval_causal")
    model = _build(cfg, {k[2:]: v for k, v in z.items() if k.startswith("w:")})
    with pytest.raises(ValueError):
        model.streaming()


def _train_model(cfg, weights):
    import conv_tasnet
    model = conv_tasnet.ConvTasNet(**cfg)
    model.load_state_dict(weights)
    return model.to(DEV).train()


def test_bn_training_matches_reference_fixture():
    """nn.BatchNorm1d in train() mode (conv_tasnet.py:300-303): batch statistics in the forward, the running
    statistics / num_batches_tracked update, the loss and every parameter gradient against what the UNMODIFIED
    reference recorded (oracle/make_golden.py, case bn_train)."""
    import pit_criterion
    cfg, z = load_golden("bn_train")
    model = _train_model(cfg, {k[2:]: v for k, v in z.items() if k.startswith("w:")})
    est = model(z["mixture"].to(DEV))
    assert est.requires_grad
    assert rel_l2(est, z["est"].to(DEV)) < 1e-4       # fp32 reference; north_star tolerance 1e-3
    sd = model.state_dict()
    updated = {k[2:]: v for k, v in z.items() if k.startswith("u:")}
    assert len(updated) == 3 * 2 * cfg["R"] * cfg["X"]
    for k, v in updated.items():
        if v.is_floating_point():
            assert torch.allclose(sd[k].cpu(), v, rtol=2e-5, atol=2e-6), k
        else:
            assert int(sd[k]) == int(v), k
    loss, *_ = pit_criterion.cal_loss(z["sources"].to(DEV), est, z["lengths"].to(DEV))
    assert abs(float(loss.detach()) - float(z["loss"])) < 1e-3
    loss.backward()
    worst = ("", 0.0)
    sc, sc_ref = [], []
    for k, p in model.named_parameters():
        assert p.grad is not None, k
        ref = z["g:" + k].to(DEV)
        if p.numel() == 1:               # PReLU slopes: compared as one vector (tests/test_gpu_train.py)
            sc.append(p.grad.flatten()), sc_ref.append(ref.flatten())
            continue
        err = rel_l2(p.grad, ref)
        worst = max(worst, (k, err), key=lambda t: t[1])
        assert err < 2e-3, (k, err)
    assert rel_l2(torch.cat(sc), torch.cat(sc_ref)) < 2e-3
    print("bn_train worst grad rel-L2", worst)
    # the updated running statistics are what eval mode now uses
    with torch.no_grad():
        ev = model.eval()(z["mixture"].to(DEV))
        ref_ev = O.forward(z["mixture"].double(), {k: v.cpu().double() if v.is_floating_point() else v.cpu()
                                                   for k, v in model.state_dict().items()}, cfg)
    assert rel_l2(ev, ref_ev.to(DEV)) < 1e-4


@pytest.mark.parametrize("causal,kinks", [(False, False), (True, False), (False, True), (True, True)])
def test_bn_training_tensor_core_size_matches_oracle(causal, kinks):
    """A BatchNorm model on the tcgen05 path (channel counts that are multiples of 32), batch 3, ragged lengths:
    forward, buffers and gradients against autograd over the oracle's training-mode forward in fp64; a second
    step continues from the updated buffers; train() under no_grad updates the buffers too.
    kinks=False: every PReLU slope is 1 and the mask is a softmax, so the network has no branch point and the
    gradient is a smooth function of the activations: it must match to rounding (measured 1-2e-5), which pins
    the BatchNorm backward (batch-mean terms, d weight, d bias).  kinks=True (slope 0.25, ReLU mask): batch
    normalisation centres every channel ON the PReLU kink, so the elements within the forward's 1e-5 of zero
    (about one per layer here) take the other branch than in fp64 and each moves the gradient by
    0.75 |dy_i|: 1e-3 .. 7e-3 of the flat gradient over seeds (tools/diag_bn_grads.py; the same
    conditioning effect as test_gradients_demo_fixture_kink), held to 2e-2."""
    import pit_criterion
    cfg = dict(N=64, L=16, B=64, H=128, P=3, X=3, R=2, C=2, norm_type="BN", causal=causal,
               mask_nonlinear="relu" if kinks else "softmax")
    p32 = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=9, norm_type="BN")
    if not kinks:
        p32 = {k: (torch.ones_like(v) if v.shape == (1,) else v) for k, v in p32.items()}
    model = _train_model(cfg, p32)
    g = torch.Generator().manual_seed(10)
    M, T = 3, 3000
    src = (0.1 * torch.randn(M, 2, T, generator=g)).to(DEV)
    mix = src.sum(1)
    lengths = torch.tensor([T, 2500, T], device=DEV)
    state = {k: (v.to(DEV).double() if v.is_floating_point() else v.to(DEV)) for k, v in p32.items()}
    for step in range(2):
        model.zero_grad()
        est = model(mix)
        est0 = est.detach().clone()         # cal_loss masks its argument in place (pit_criterion.py:38)
        loss, *_ = pit_criterion.cal_loss(src, est, lengths)
        loss.backward()
        p64 = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running_" not in k else v)
               for k, v in state.items()}
        new_stats = {}
        est_ref = O.forward(mix.double(), p64, cfg, new_stats=new_stats)
        loss_ref, *_ = O.cal_loss(src.double(), est_ref, lengths)
        loss_ref.backward()
        assert rel_l2(est0, est_ref) < 1e-4, step
        assert abs(float(loss.detach()) - float(loss_ref.detach())) < 0.05
        sd = model.state_dict()
        for k, v in new_stats.items():
            if v.is_floating_point():
                assert torch.allclose(sd[k].double(), v.detach(), rtol=1e-4, atol=2e-6), (step, k)
            else:
                assert int(sd[k]) == int(v), (step, k)
        flat = torch.cat([p.grad.flatten().double() for _, p in model.named_parameters()])
        flat_ref = torch.cat([p64[k].grad.flatten() for k, _ in model.named_parameters()])
        print("bn train", causal, kinks, "step", step, "flat grad rel-L2", rel_l2(flat, flat_ref))
        assert rel_l2(flat, flat_ref) < (2e-2 if kinks else 2e-4), step
        if not kinks:
            for k, p in model.named_parameters():
                if p.numel() > 1:       # (PReLU slopes: sums with heavy cancellation, part of the flat comparison)
                    assert rel_l2(p.grad, p64[k].grad) < 2e-3, (step, k, rel_l2(p.grad, p64[k].grad))
        state.update({k: v.detach() for k, v in new_stats.items()})
    before = {k: v.clone() for k, v in model.state_dict().items() if "running_mean" in k}
    with torch.no_grad():
        out = model(mix)                    # train() mode without a graph: batch statistics, buffers move
    assert not out.requires_grad
    new_stats = {}
    ref = O.forward(mix.double(), {k: v.detach() for k, v in state.items()}, cfg, new_stats=new_stats)
    assert rel_l2(out, ref) < 1e-4
    assert all(not torch.equal(before[k], model.state_dict()[k]) for k in before)
    for k, v in new_stats.items():
        if v.is_floating_point():
            assert torch.allclose(model.state_dict()[k].double(), v, rtol=1e-4, atol=2e-6), k


def test_bn_eval_mode_gradients_use_running_statistics():
    """model.eval() with autograd: BatchNorm1d is a fixed per-channel affine of the running statistics (no batch
    terms in the backward, buffers untouched); gradients against autograd over the oracle's eval-mode forward."""
    import pit_criterion
    cfg = dict(N=64, L=16, B=64, H=128, P=3, X=2, R=2, C=2, norm_type="BN", causal=False, mask_nonlinear="relu")
    p32 = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=11, norm_type="BN")
    model = _train_model(cfg, p32).eval()
    g = torch.Generator().manual_seed(12)
    src = (0.1 * torch.randn(2, 2, 2000, generator=g)).to(DEV)
    mix = src.sum(1)
    lengths = torch.tensor([2000, 1500], device=DEV)
    before = {k: v.clone() for k, v in model.state_dict().items()}
    est = model(mix)
    est0 = est.detach().clone()             # cal_loss masks its argument in place (pit_criterion.py:38)
    loss, *_ = pit_criterion.cal_loss(src, est, lengths)
    loss.backward()
    after = model.state_dict()
    assert all(torch.equal(before[k], after[k]) for k in before)
    p64 = {k: (v.to(DEV).double().requires_grad_(True) if v.is_floating_point() and "running_" not in k
               else (v.to(DEV).double() if v.is_floating_point() else v.to(DEV))) for k, v in p32.items()}
    est_ref = O.forward(mix.double(), p64, cfg)
    loss_ref, *_ = O.cal_loss(src.double(), est_ref, lengths)
    loss_ref.backward()
    assert rel_l2(est0, est_ref) < 1e-4
    flat = torch.cat([p.grad.flatten().double() for _, p in model.named_parameters()])
    flat_ref = torch.cat([p64[k].grad.flatten() for k, _ in model.named_parameters()])
    assert rel_l2(flat, flat_ref) < 1e-3
