"""Chunked (streaming) causal separation against the whole-utterance forward of the same model
(SURVEY.md 8(f) rank 2): concatenated feed()/flush() outputs must reproduce ConvTasNet.forward of
the full stream -- same kernels, same per-frame arithmetic -- for any chunk size and any way the
stream is cut into pieces; and the whole-utterance forward itself is checked against the oracle."""
import pytest
import torch

from conftest import load_golden, rel_l2
from oracle import convtasnet_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 0.0       # same kernels, same per-frame arithmetic: chunked output is BIT-IDENTICAL to the whole forward


def _build(cfg, weights):
    import conv_tasnet
    model = conv_tasnet.ConvTasNet(**cfg)
    model.load_state_dict(weights)
    return model.to(DEV).eval()


def _stream(sep, mix, pieces):
    outs, pos = [], 0
    for n in pieces:
        outs.append(sep.feed(mix[pos:pos + n]))
        pos += n
    if pos < mix.numel():
        outs.append(sep.feed(mix[pos:]))
    outs.append(sep.flush())
    return torch.cat(outs, dim=1)


def _small():
    cfg, z = load_golden("tc_causal_cln")
    return cfg, _build(cfg, {k[2:]: v for k, v in z.items() if k.startswith("w:")})


@pytest.mark.parametrize("chunk", [1, 7, 64, 300])
@pytest.mark.parametrize("graph", [False, True])
def test_stream_equals_whole_forward_small(chunk, graph):
    cfg, model = _small()
    hop = cfg["L"] // 2
    g = torch.Generator().manual_seed(5)
    T = 40 * 64 * hop // 8 + 3
    mix = (0.1 * torch.randn(T, generator=g)).to(DEV)
    with torch.no_grad():
        whole = model(mix[None])[0]
    sep = model.streaming(chunk_frames=chunk, use_graph=graph)
    got = _stream(sep, mix, [1, 17, 500, 3, 1000])
    assert got.shape == whole.shape
    err = rel_l2(got, whole)
    print("chunk", chunk, "graph", graph, "rel-L2", err, "bit-exact", bool(torch.equal(got, whole)))
    assert err <= TOL and torch.equal(got, whole)
    # the object is reusable: a second stream gives the same answer
    got2 = _stream(sep, mix, [T])
    assert torch.equal(got2, got)


def test_stream_short_streams_and_errors():
    cfg, model = _small()
    L, hop = cfg["L"], cfg["L"] // 2
    sep = model.streaming(chunk_frames=32)
    g = torch.Generator().manual_seed(6)
    for T in (L, L + 1, 3 * hop + 5, 33 * hop, 33 * hop + 1, 34 * hop - 1, 65 * hop, 65 * hop + hop - 1):
        mix = (0.1 * torch.randn(T, generator=g)).to(DEV)
        with torch.no_grad():
            whole = model(mix[None])[0]
        got = _stream(sep, mix, [T // 3])
        assert got.shape == whole.shape, T
        assert torch.equal(got, whole), T
        # zero tail beyond the last whole frame (conv_tasnet.py:56-59)
        t_conv = ((T - L) // hop) * hop + L
        assert float(got[:, t_conv:].abs().sum()) == 0.0
    sep.feed(torch.zeros(L - 1, device=DEV))
    with pytest.raises(RuntimeError):
        sep.flush()
    import conv_tasnet
    with pytest.raises(ValueError):
        conv_tasnet.ConvTasNet(**dict(cfg, norm_type="gLN")).to(DEV).streaming()
    with pytest.raises(ValueError):
        conv_tasnet.ConvTasNet(**dict(cfg, causal=False)).to(DEV).streaming()
    with pytest.raises(RuntimeError):
        sep.feed(torch.zeros(10))


@pytest.mark.parametrize("chunk", [100, 256, 1000])
def test_stream_paper_causal_config(chunk):
    """BASELINE configs[2] model (causal cLN paper configuration): chunked = whole, and whole = oracle."""
    cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="cLN", causal=True, mask_nonlinear="relu")
    p32 = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=1234)
    model = _build(cfg, p32)
    g = torch.Generator().manual_seed(9)
    T = 32007
    mix = (0.1 * torch.randn(T, generator=g)).to(DEV)
    with torch.no_grad():
        whole = model(mix[None])[0]
    sep = model.streaming(chunk_frames=chunk)
    got = _stream(sep, mix, [4000, 123, 16000])
    err = rel_l2(got, whole)
    print("paper causal, chunk", chunk, "rel-L2", err, "bit-exact", bool(torch.equal(got, whole)))
    assert err <= TOL and torch.equal(got, whole)
    if chunk == 256:
        ref = O.forward(mix[None].double(), {k: v.to(DEV).double() for k, v in p32.items()}, cfg)[0]
        assert rel_l2(got, ref) < 1e-3


def test_stream_full_size_config3():
    """BASELINE configs[2] at full size: one 60 s stream (480 000 samples, 47 999 frames) separated in
    1.28 s chunks is bit-identical to the whole-stream forward."""
    cfg = dict(N=256, L=20, B=256, H=512, P=3, X=8, R=4, C=2, norm_type="cLN", causal=True, mask_nonlinear="relu")
    p32 = O.random_params(*(cfg[k] for k in "NLBHPXRC"), cfg["causal"], seed=1234)
    model = _build(cfg, p32)
    g = torch.Generator().manual_seed(10)
    mix = (0.1 * torch.randn(480000, generator=g)).to(DEV)
    with torch.no_grad():
        whole = model(mix[None])[0]
    sep = model.streaming(chunk_frames=1024)
    got = _stream(sep, mix, [100000, 200000, 7])
    assert got.shape == whole.shape and torch.equal(got, whole)
