"""Size-independent properties of the path at BASELINE.json's full sizes, where the
oracle is too slow to be the checker: linearity of overlap-and-add (utils.py:9-46),
permutation / scale invariance of the PIT SI-SNR (pit_criterion.py:27-76), the length
mask (pit_criterion.py:101-113), and the homogeneity of the relu-mask network in the
mixture gain for gLN (every op between encoder and decoder is scale-equivariant or
scale-invariant once EPS is negligible).
"""
import itertools

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


@pytest.mark.parametrize("shape,step", [((64, 2, 3199, 20), 10), ((3, 3, 5999, 16), 8), ((1, 1, 7, 2), 1),
                                        ((2, 2, 100, 20), 20)])
def test_overlap_and_add_is_linear_and_conserves_mass(shape, step):
    import utils
    g = torch.Generator().manual_seed(11)
    a = torch.randn(*shape, generator=g).to(DEV)
    b = torch.randn(*shape, generator=g).to(DEV)
    oa, ob = utils.overlap_and_add(a, step), utils.overlap_and_add(b, step)
    K, L = shape[-2], shape[-1]
    assert oa.shape == shape[:-2] + ((K - 1) * step + L,)
    lin = utils.overlap_and_add(2.0 * a - 3.0 * b, step)
    assert _rel(lin, 2.0 * oa - 3.0 * ob) < 1e-6
    # every frame sample lands in exactly one output sample: totals agree
    assert torch.allclose(oa.double().sum(-1), a.double().sum((-1, -2)), rtol=1e-5, atol=1e-3)
    # a single non-zero frame lands at k * step, untouched
    z = torch.zeros(*shape, device=DEV)
    k = K // 2
    z[..., k, :] = a[..., k, :]
    oz = utils.overlap_and_add(z, step)
    assert torch.equal(oz[..., k * step:k * step + L], a[..., k, :])
    oz[..., k * step:k * step + L] = 0
    assert float(oz.abs().max()) == 0.0


@pytest.mark.parametrize("B,C,T", [(64, 2, 32000), (16, 3, 32000), (4, 4, 8001)])
def test_pit_is_invariant_to_speaker_order_and_estimate_scale(B, C, T):
    import pit_criterion
    g = torch.Generator().manual_seed(12)
    src = torch.randn(B, C, T, generator=g).to(DEV)
    est = src[:, torch.randperm(C, generator=g)] + 0.5 * torch.randn(B, C, T, generator=g).to(DEV)
    lens = torch.randint(T // 2, T + 1, (B,), generator=g).to(DEV)
    lens[0] = T
    base, perms, idx0 = pit_criterion.cal_si_snr_with_pit(src, est.clone(), lens)
    assert perms.tolist() == [list(p) for p in itertools.permutations(range(C))]
    for p in list(itertools.permutations(range(C)))[1:4]:
        m, _, idx = pit_criterion.cal_si_snr_with_pit(src, est[:, list(p)].clone(), lens)
        assert torch.allclose(m, base, atol=1e-4)
        # estimate i of the shuffled batch is estimate p[i]: its target is perm0[p[i]]
        for b in (0, B - 1):
            perm0 = perms[int(idx0[b])].tolist()
            assert perms[int(idx[b])].tolist() == [perm0[p[i]] for i in range(C)]
    # SI-SNR does not see a per-speaker gain on the estimate
    gains = (0.1 + 5.0 * torch.rand(1, C, 1, generator=g)).to(DEV)
    m, _, idx = pit_criterion.cal_si_snr_with_pit(src, (est * gains).clone(), lens)
    assert torch.allclose(m, base, atol=2e-4) and torch.equal(idx, idx0)
    # nor a common gain on the sources
    m, _, idx = pit_criterion.cal_si_snr_with_pit(7.0 * src, est.clone(), lens)
    assert torch.allclose(m, base, atol=2e-4) and torch.equal(idx, idx0)


def test_pit_ignores_samples_beyond_the_length_and_masks_them_in_place():
    import pit_criterion
    g = torch.Generator().manual_seed(13)
    B, C, T = 8, 2, 32000
    src = torch.randn(B, C, T, generator=g).to(DEV)
    est = src + 0.3 * torch.randn(B, C, T, generator=g).to(DEV)
    lens = torch.tensor([T, T - 1, 20000, 16001, 16000, 9, 3, 2], device=DEV)
    mask = pit_criterion.get_mask(src, lens)
    assert mask.shape == (B, 1, T) and mask.dtype == src.dtype
    assert torch.equal(mask[:, 0].sum(-1).long(), lens)
    e1 = est.clone()
    base, _, idx0 = pit_criterion.cal_si_snr_with_pit(src, e1, lens)
    assert torch.equal(e1, est * mask)                       # pit_criterion.py:38 side effect
    junk = est + (1 - mask) * 100.0 * torch.randn(B, C, T, generator=g).to(DEV)
    m, _, idx = pit_criterion.cal_si_snr_with_pit(src, junk.clone(), lens)
    assert torch.equal(m, base) and torch.equal(idx, idx0)
    # truncating to the longest length changes nothing either
    m, _, idx = pit_criterion.cal_si_snr_with_pit(src[:2], est[:2].clone(), lens[:2])
    assert torch.equal(m, base[:2])


def test_relu_mask_gln_network_is_homogeneous_in_the_mixture_gain():
    """Paper configuration, 4 s: separate(g * x) == g * separate(x) up to EPS effects."""
    import conv_tasnet
    torch.manual_seed(14)
    model = conv_tasnet.ConvTasNet(256, 20, 256, 512, 3, 8, 4, 2, norm_type="gLN", causal=False,
                                   mask_nonlinear="relu").to(DEV).eval()
    g = torch.Generator().manual_seed(15)
    mix = (0.1 * torch.randn(4, 32000, generator=g)).to(DEV)
    with torch.no_grad():
        y1 = model(mix)
        y4 = model(4.0 * mix)
        again = model(mix)
    assert torch.equal(y1, again)                            # run-to-run bit-reproducible
    assert _rel(y4, 4.0 * y1) < 1e-3


def test_ragged_input_length_is_zero_padded_back():
    """T not a multiple of the hop: output has the input's length (conv_tasnet.py:56-59)."""
    import conv_tasnet
    torch.manual_seed(16)
    model = conv_tasnet.ConvTasNet(64, 16, 64, 128, 3, 3, 2, 2).to(DEV).eval()
    g = torch.Generator().manual_seed(17)
    for T in (16, 17, 4001, 4007):
        mix = (0.1 * torch.randn(2, T, generator=g)).to(DEV)
        with torch.no_grad():
            y = model(mix)
        assert y.shape == (2, 2, T)
        K = (T - 16) // 8 + 1
        covered = (K - 1) * 8 + 16
        assert float(y[..., covered:].abs().sum()) == 0.0


def test_micro_batches_on_side_streams_are_bit_identical():
    """engine.n_streams > 1 issues consecutive micro-batches on separate streams (separate workspaces):
    the result must not depend on it."""
    import conv_tasnet
    from conftest import load_golden
    cfg, z = load_golden("tc_small")
    model = conv_tasnet.ConvTasNet(**cfg)
    model.load_state_dict({k[2:]: v for k, v in z.items() if k.startswith("w:")})
    model = model.to("cuda").eval()
    g = torch.Generator().manual_seed(4)
    mix = (0.1 * torch.randn(7, 1207, generator=g)).to("cuda")
    eng = model.engine()
    eng.micro_batch = 2
    outs = []
    with torch.no_grad():
        for n, graph in ((1, False), (2, False), (3, True), (2, True)):
            eng.n_streams, eng.use_graph = n, graph
            outs.append(model(mix).clone())
    for o in outs[1:]:
        assert torch.equal(o, outs[0])


def test_host_to_host_separation_equals_device_forward():
    """ConvTasNet.separate (pinned host in / out, double-buffered copies on side streams) must return
    exactly what forward() returns, for ragged micro-batch splits, with and without CUDA graphs."""
    import conv_tasnet
    from conftest import load_golden
    cfg, z = load_golden("tc_small")
    model = conv_tasnet.ConvTasNet(**cfg)
    model.load_state_dict({k[2:]: v for k, v in z.items() if k.startswith("w:")})
    model = model.to("cuda").eval()
    g = torch.Generator().manual_seed(8)
    mix = 0.1 * torch.randn(7, 1207, generator=g)
    eng = model.engine()
    with torch.no_grad():
        want = model(mix.cuda()).cpu()
    for mb, graph, pinned in ((2, False, True), (3, True, True), (16, False, False), (2, True, False)):
        eng.micro_batch, eng.use_graph = mb, graph
        src = mix.pin_memory() if pinned else mix.clone()
        got = model.separate(src)
        assert got.device.type == "cpu" and torch.equal(got, want)
        out = torch.empty(7, cfg["C"], 1207).pin_memory()
        assert model.separate(src, out=out) is out and torch.equal(out, want)
    with pytest.raises(RuntimeError):
        model.separate(mix.cuda())


def test_full_size_config4_bulk_separation_is_batch_independent():
    """BASELINE configs[3] at full size (512 x 4 s, paper configuration) through the host-to-host call:
    utterance m of the bulk result equals the same utterance separated alone, whatever micro-batch it
    fell into.  With the unfused kernels that is bit-exact.  The fused depthwise + pointwise kernel hands
    its k-blocks to producer groups round-robin ACROSS the tiles of a CTA pair, so from 6 utterances per
    micro-batch on (a pair owns more than one tile) the float partial sums of the gLN statistics are
    grouped differently than for a lone utterance: deterministic, and equal to rounding (~1e-5, the size
    of the error against the fp64 oracle; tolerance 1e-3)."""
    import conv_tasnet
    from oracle import convtasnet_oracle as O
    cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")
    model = conv_tasnet.ConvTasNet(**cfg)
    model.load_state_dict(O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=1234))
    model = model.to("cuda").eval()
    eng = model.engine()
    g = torch.Generator().manual_seed(2025)
    mix = (0.1 * torch.randn(512, 32000, generator=g)).pin_memory()
    for fuse, tol in ((1 << 30, 5e-5), (0, 0.0)):
        eng.fuse_max_dilation = fuse
        eng.micro_batch, eng.use_graph = 64, True
        est = model.separate(mix)
        assert tuple(est.shape) == (512, 2, 32000) and bool(torch.isfinite(est).all())
        again = model.separate(mix)
        assert torch.equal(est, again)                      # run-to-run: bit-identical either way
        eng.use_graph = False
        with torch.no_grad():
            for m in (0, 63, 64, 300, 511):
                alone = model(mix[m:m + 1].cuda()).cpu()[0]
                err = float((alone.double() - est[m].double()).norm() / alone.double().norm())
                assert err <= tol, (fuse, m, err)
