"""norm_type="BN": the reference's chose_norm falls through to nn.BatchNorm1d (conv_tasnet.py:300-303).
Inference (running statistics) runs through the same fused kernels as gLN with the statistics folded
into a per-channel affine (ctn_bn_fold); training-mode batch statistics are not implemented and raise."""
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


def test_bn_training_mode_raises():
    cfg, z = load_golden("bn_eval_causal")
    model = _build(cfg, {k[2:]: v for k, v in z.items() if k.startswith("w:")})
    mix = z["mixture"].to(DEV)
    with pytest.raises(NotImplementedError):
        model.train()(mix)                       # batch statistics
    with pytest.raises(NotImplementedError):
        model.eval()(mix)                        # eval mode but autograd requested
    with torch.no_grad():
        model.eval()(mix)
    with pytest.raises(ValueError):
        model.streaming()
