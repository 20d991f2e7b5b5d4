"""ctn_b200.optim.FusedAdam (ctn_clip_adam) against torch.nn.utils.clip_grad_norm_ +
torch.optim.Adam (solver.py:184-186, train.py:129-131) on the same parameters and gradients."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _params(seed, shapes):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(*s, generator=g).to(DEV)) for s in shapes]


SHAPES = [(256, 1, 20), (1, 256, 1), (512, 256, 1), (1,), (512, 1, 3), (3, 5), (8193,), (70000,)]


@pytest.mark.parametrize("max_norm,wd", [(5.0, 0.0), (0.05, 0.0), (5.0, 1e-3), (None, 0.0)])
def test_matches_torch_clip_and_adam(max_norm, wd):
    from ctn_b200.optim import FusedAdam
    ours = _params(1, SHAPES)
    ref = [torch.nn.Parameter(p.detach().clone()) for p in ours]
    o1 = FusedAdam(ours, lr=1e-3, weight_decay=wd, max_norm=max_norm)
    o2 = torch.optim.Adam(ref, lr=1e-3, weight_decay=wd)
    g = torch.Generator().manual_seed(2)
    for it in range(5):
        scale = [1.0, 30.0, 0.01, 1.0, 1.0][it]
        for a, b in zip(ours, ref):
            gr = (scale * torch.randn(*a.shape, generator=g)).to(DEV)
            a.grad = gr.clone()
            b.grad = gr.clone()
        norm_ref = torch.nn.utils.clip_grad_norm_(ref, max_norm) if max_norm is not None else None
        o2.step()
        o1.step()
        if max_norm is not None:
            assert abs(float(o1.total_norm) - float(norm_ref)) <= 1e-5 * float(norm_ref)
        for a, b in zip(ours, ref):
            assert torch.allclose(a, b, rtol=2e-6, atol=2e-7), (it, a.shape, float((a - b).abs().max()))
    for a, b in zip(ours, ref):
        # (sums of terms of mixed sign: absolute tolerance relative to the tensor's scale)
        m_ref, v_ref = o2.state[b]["exp_avg"], o2.state[b]["exp_avg_sq"]
        assert torch.allclose(o1.state[a]["exp_avg"], m_ref, rtol=1e-5, atol=1e-6 * float(m_ref.abs().max()))
        assert torch.allclose(o1.state[a]["exp_avg_sq"], v_ref, rtol=1e-5, atol=1e-6 * float(v_ref.abs().max()))
        assert int(o1.state[a]["step"]) == int(o2.state[b]["step"]) == 5


def test_already_clipped_gradients_and_written_back_clip():
    """solver.py clips before optimizer.step(): a second clip inside step() must be the identity."""
    from ctn_b200.optim import FusedAdam
    ours = _params(3, SHAPES)
    ref = [torch.nn.Parameter(p.detach().clone()) for p in ours]
    o1 = FusedAdam(ours, lr=1e-3, max_norm=5.0, write_clipped_grad=True)
    o2 = torch.optim.Adam(ref, lr=1e-3)
    g = torch.Generator().manual_seed(4)
    for a, b in zip(ours, ref):
        gr = (10.0 * torch.randn(*a.shape, generator=g)).to(DEV)
        a.grad, b.grad = gr.clone(), gr.clone()
    torch.nn.utils.clip_grad_norm_(ours, 5.0)
    torch.nn.utils.clip_grad_norm_(ref, 5.0)
    o1.step(), o2.step()
    for a, b in zip(ours, ref):
        assert torch.allclose(a, b, rtol=2e-6, atol=2e-7)
        assert torch.allclose(a.grad, b.grad, rtol=1e-5, atol=1e-9)


def test_state_dict_round_trip_with_torch_adam():
    """Checkpoints interoperate with torch.optim.Adam (solver.py:56, :116-119)."""
    from ctn_b200.optim import FusedAdam
    ours = _params(5, SHAPES)
    ref = [torch.nn.Parameter(p.detach().clone()) for p in ours]
    o1 = FusedAdam(ours, lr=1e-3, max_norm=5.0)
    o2 = torch.optim.Adam(ref, lr=1e-3)
    g = torch.Generator().manual_seed(6)

    def grads():
        for a, b in zip(ours, ref):
            gr = torch.randn(*a.shape, generator=g).to(DEV)
            a.grad, b.grad = gr.clone(), gr.clone()
    for _ in range(2):
        grads()
        torch.nn.utils.clip_grad_norm_(ref, 5.0)
        o1.step(), o2.step()
    # torch -> ours and ours -> torch, with the learning-rate halving of solver.py:116-119
    sd2 = copy.deepcopy(o2.state_dict())
    sd2["param_groups"][0]["lr"] /= 2.0
    sd1 = copy.deepcopy(o1.state_dict())
    sd1["param_groups"][0]["lr"] /= 2.0
    n1 = FusedAdam(ours, lr=1e-3, max_norm=5.0)
    n1.load_state_dict(sd2)
    n2 = torch.optim.Adam(ref, lr=1e-3)
    n2.load_state_dict({"state": sd1["state"], "param_groups": sd2["param_groups"]})
    assert n1.param_groups[0]["lr"] == 5e-4
    grads()
    torch.nn.utils.clip_grad_norm_(ref, 5.0)
    n1.step(), n2.step()
    for a, b in zip(ours, ref):
        assert torch.allclose(a, b, rtol=3e-6, atol=3e-7)
    assert int(n1.state[ours[0]]["step"]) == 3


def test_training_loop_with_fused_adam_reduces_loss():
    import conv_tasnet
    import pit_criterion
    from ctn_b200.optim import FusedAdam
    torch.manual_seed(0)
    model = conv_tasnet.ConvTasNet(64, 16, 64, 128, 3, 3, 2, 2).to(DEV).train()
    opt = FusedAdam(model.parameters(), lr=1e-3, max_norm=5.0)
    g = torch.Generator().manual_seed(1)
    src = (0.1 * torch.randn(3, 2, 4000, generator=g)).to(DEV)
    mix = src.sum(1)
    lens = torch.full((3,), 4000, dtype=torch.int64, device=DEV)
    losses = []
    for _ in range(12):
        est = model(mix)
        loss, *_ = pit_criterion.cal_loss(src, est, lens)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0] - 0.5
    assert float(opt.total_norm) > 0


def test_gradients_as_views_of_a_moving_flat_buffer():
    """The training engine hands back all gradients as views of one flat buffer that lives at a new
    address every step: the optimizer then shifts the gradient column of its chunk table on the device
    instead of rebuilding it.  Same result as torch's clip + Adam; a step where the layout itself
    changes (separate tensors again) falls back to a rebuild."""
    from ctn_b200.optim import FusedAdam
    ours = _params(5, SHAPES)
    ref = [torch.nn.Parameter(p.detach().clone()) for p in ours]
    o1 = FusedAdam(ours, lr=1e-3, max_norm=5.0)
    o2 = torch.optim.Adam(ref, lr=1e-3)
    g = torch.Generator().manual_seed(6)
    total = sum(p.numel() for p in ours)
    keep = []
    for it in range(6):
        flat = torch.randn(total + 64 * it, generator=g).to(DEV)
        keep.append(flat)            # keep earlier buffers alive so that every step sees a new address
        off = 0
        for a, b in zip(ours, ref):
            k = a.numel()
            view = flat[off:off + k].view_as(a)
            a.grad = view if it != 4 else view.clone()
            b.grad = view.clone()
            off += k
        torch.nn.utils.clip_grad_norm_(ref, 5.0)
        o2.step()
        o1.step()
        for a, b in zip(ours, ref):
            assert torch.allclose(a, b, rtol=2e-6, atol=2e-7), (it, a.shape, float((a - b).abs().max()))


def test_fused_adam_trains_every_weight_like_torch_adam():
    """The optimizer writes the parameters through raw device pointers; the engines' prepared weight copies
    (bf16 splits with the norms folded in, transposed backward operands: ~95 % of the parameters) are keyed
    on the tensors' version counters, so step() must advance them.  Two identical models, one stepped by
    FusedAdam and one by clip_grad_norm_ + torch.optim.Adam (solver.py:183-186), must stay together on EVERY
    tensor -- the 1x1 convolutions included -- and on the loss trajectory."""
    import conv_tasnet
    import pit_criterion
    from ctn_b200.optim import FusedAdam
    torch.manual_seed(0)
    m1 = conv_tasnet.ConvTasNet(64, 16, 64, 128, 3, 3, 2, 2).to(DEV).train()
    m2 = copy.deepcopy(m1)
    o1 = FusedAdam(m1.parameters(), lr=1e-3, max_norm=5.0)
    o2 = torch.optim.Adam(m2.parameters(), lr=1e-3)
    g = torch.Generator().manual_seed(1)
    src = (0.1 * torch.randn(3, 2, 4000, generator=g)).to(DEV)
    mix = src.sum(1)
    lens = torch.full((3,), 4000, dtype=torch.int64, device=DEV)
    w0 = {k: v.detach().clone() for k, v in m1.state_dict().items()}
    l1, l2 = [], []
    for _ in range(4):
        for model, opt, losses, clip in ((m1, o1, l1, False), (m2, o2, l2, True)):
            est = model(mix)
            loss, *_ = pit_criterion.cal_loss(src, est, lens)
            opt.zero_grad()
            loss.backward()
            if clip:
                torch.nn.utils.clip_grad_norm_(model.parameters(), 5.0)
            opt.step()
            losses.append(float(loss))
    for a, b in zip(l1, l2):
        assert abs(a - b) < 2e-3, (l1, l2)
    # the losses move: a stale set of prepared weights would leave the 1x1 convolutions at their
    # initial values inside the kernels and the two trajectories would part from step 2 on
    assert abs(l1[-1] - l1[0]) > 1e-2
    sd1, sd2 = m1.state_dict(), m2.state_dict()
    for k in sd1:
        # (norms, not maxima: Adam's m / sqrt(v) amplifies rounding differences of near-zero gradients)
        step = float((sd2[k] - w0[k]).float().norm())
        diff = float((sd1[k] - sd2[k]).float().norm())
        assert diff <= 0.02 * step + 1e-7, (k, diff, step)
    # and the forward really uses the updated 1x1 weights: same output as a fresh model loaded from them
    m3 = conv_tasnet.ConvTasNet(64, 16, 64, 128, 3, 3, 2, 2).to(DEV).eval()
    m3.load_state_dict(sd1)
    with torch.no_grad():
        e3 = m3(mix)
        e1 = m1.eval()(mix)
    assert float((e1 - e3).norm() / e3.norm()) < 1e-5
