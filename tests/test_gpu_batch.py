"""Device-side batch assembly (ctn_pad_batch) and remove_pad (ctn_remove_pad) against the reference's
padded tensors (fixtures recorded from data.py's collate functions) and against host slicing
(utils.py:49-66).  Pure copies: bit-exact."""
import json
import os
import sys

import numpy as np
import pytest
import torch

from conftest import GOLDEN, ROOT

sys.path.insert(0, os.path.join(ROOT, "oracle"))
import synth_audio as sa  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda"
ARR = np.load(os.path.join(GOLDEN, "data_cases.npz"))


@pytest.mark.parametrize("name", sorted(sa.DATA_CASES))
def test_assemble_matches_reference_collate(name, monkeypatch, tmp_path):
    import data
    from ctn_b200 import batch
    monkeypatch.setattr(data, "_load_wav", lambda path, sample_rate: sa.synth(path))
    case = sa.DATA_CASES[name]
    for kind, lst in zip(("mix", "s1", "s2"), sa.make_infos(case["lengths"], case["shuffle"])):
        with open(tmp_path / (kind + ".json"), "w") as f:
            json.dump(lst, f)
    ds = data.AudioDataset(str(tmp_path), **case["kw"])
    loader = data.AudioDataLoader(ds, batch_size=1, shuffle=False, packed=True, pin_memory=True)
    for b, (mix_p, ilens, src_p) in enumerate(loader):
        mix, lengths, src = batch.assemble(mix_p, ilens, src_p, DEV)
        assert np.array_equal(mix.cpu().numpy(), ARR["%s.%d.mix" % (name, b)])
        assert np.array_equal(src.cpu().numpy(), ARR["%s.%d.src" % (name, b)])
        assert lengths.dtype == torch.int64 and np.array_equal(lengths.cpu().numpy(), ARR["%s.%d.ilens" % (name, b)])
        # and back: remove_pad returns each utterance's valid samples
        back = batch.remove_pad(src, lengths)
        for i, r in enumerate(back):
            n = int(ilens[i])
            assert r.shape == (src.shape[1], n) and np.array_equal(r, ARR["%s.%d.src" % (name, b)][i, :, :n])


@pytest.mark.parametrize("B,C,lens,T", [(3, 2, [32000, 24000, 16001], None), (5, 3, [1, 0, 4099, 1024, 1023], 4200),
                                        (1, 1, [7], None), (64, 2, None, 32000)])
def test_pad_batch_and_remove_pad_round_trip(B, C, lens, T):
    from ctn_b200 import batch
    import utils
    rng = np.random.RandomState(B * 7 + C)
    if lens is None:
        lens = rng.randint(1, 32001, size=B).tolist()
    mixtures = [rng.randn(n).astype(np.float32) for n in lens]
    sources = [rng.randn(n, C).astype(np.float32) for n in lens]
    mix_p, ilens, src_p = batch.pack(mixtures, sources)
    assert mix_p.is_pinned() and src_p.is_pinned()
    mix, lengths, src = batch.assemble(mix_p, ilens, src_p, DEV, T=T, pad_value=0.0)
    Tm = T or max(lens)
    assert mix.shape == (B, Tm) and src.shape == (B, C, Tm) and lengths.tolist() == lens
    mh, sh = mix.cpu().numpy(), src.cpu().numpy()
    for i, n in enumerate(lens):
        assert np.array_equal(mh[i, :n], mixtures[i]) and not mh[i, n:].any()
        assert np.array_equal(sh[i, :, :n], sources[i].T) and not sh[i, :, n:].any()
    # non-zero pad value, mixtures only
    mix2, _, none = batch.assemble(mix_p, ilens, None, DEV, T=Tm + 3, pad_value=-2.0)
    assert none is None and all((mix2[i, n:] == -2.0).all() for i, n in enumerate(lens))
    # remove_pad: [B, C, T] and [B, T], host or device lengths, through the drop-in's utils too
    for got in (batch.remove_pad(src, lengths), utils.remove_pad(src, ilens)):
        assert all(np.array_equal(g, s.T) for g, s in zip(got, sources))
    got = utils.remove_pad(mix, lengths)
    assert all(g.ndim == 1 and np.array_equal(g, m) for g, m in zip(got, mixtures))


def test_assemble_feeds_the_model_like_host_padding():
    """Ragged batch through the packed route = the same batch padded on the host, bit for bit."""
    from conv_tasnet import ConvTasNet
    from ctn_b200 import batch
    torch.manual_seed(0)
    model = ConvTasNet(64, 16, 32, 64, 3, 3, 2, 2).to(DEV).eval()
    rng = np.random.RandomState(1)
    lens = [4000, 3111, 2500]
    mixtures = [0.1 * rng.randn(n).astype(np.float32) for n in lens]
    host = torch.zeros(3, 4000)
    for i, m in enumerate(mixtures):
        host[i, :lens[i]] = torch.from_numpy(m)
    mix_p, ilens, _ = batch.pack(mixtures)
    mix, lengths, _ = batch.assemble(mix_p, ilens, None, DEV)
    with torch.no_grad():
        a, b = model(mix), model(host.to(DEV))
    assert torch.equal(a, b)
    outs = batch.remove_pad(a, lengths)
    assert [o.shape for o in outs] == [(2, n) for n in lens]


def test_errors():
    from ctn_b200 import batch
    with pytest.raises(ValueError):
        batch.assemble(torch.zeros(5), torch.tensor([2, 2]), None, DEV)
    with pytest.raises(RuntimeError):
        batch.assemble(torch.zeros(4), torch.tensor([2, 2]), None, "cpu")
    with pytest.raises(RuntimeError):
        batch.remove_pad(torch.zeros(2, 3), torch.tensor([1, 1]))
    mix, lengths, _ = batch.assemble(torch.zeros(0), torch.zeros(0, dtype=torch.int64), None, DEV)
    assert mix.shape == (0, 0) and lengths.numel() == 0
    assert batch.remove_pad(torch.zeros(0, 2, 5, device=DEV), torch.zeros(0, dtype=torch.int64)) == []


def test_separate_script_end_to_end(tmp_path):
    """The drop-in separate.py on a directory of ragged wav files: every output file holds what the model
    gives for the batch padded on the host (separate.py:60-80), as PCM16."""
    import wave
    import data
    import separate as sep
    from conv_tasnet import ConvTasNet
    torch.manual_seed(3)
    model = ConvTasNet(64, 16, 32, 64, 3, 3, 2, 2).to(DEV).eval()
    torch.save(ConvTasNet.serialize(model, torch.optim.Adam(model.parameters()), 1), tmp_path / "model.pth.tar")
    rng = np.random.RandomState(4)
    mix_dir = tmp_path / "mix"
    mix_dir.mkdir()
    lens = {"u1.wav": 4000, "b.wav": 2777, "c.wav": 3500}
    for name, n in lens.items():
        sep.write_wav(0.2 * rng.randn(n), str(mix_dir / name), 8000)
    args = sep.parser.parse_args(["--model_path", str(tmp_path / "model.pth.tar"), "--mix_dir", str(mix_dir),
                                  "--out_dir", str(tmp_path / "out"), "--batch_size", "2"])
    sep.separate(args, num_workers=0)    # no forked loader workers inside the test process
    # expected: batches of 2 in length order (u1, c), (b); padded on the host
    def read(path):
        return data._load_wav(str(path), 8000)
    for group in (["u1.wav", "c.wav"], ["b.wav"]):
        waves = [read(mix_dir / n) for n in group]
        T = max(len(w) for w in waves)
        host = torch.zeros(len(group), T)
        for i, w in enumerate(waves):
            host[i, :len(w)] = torch.from_numpy(w)
        with torch.no_grad():
            est = model(host.to(DEV)).cpu().numpy()
        for i, n in enumerate(group):
            stem = n.strip('.wav')     # the reference's way of dropping the suffix (separate.py:73)
            assert np.array_equal(read(tmp_path / "out" / (stem + ".wav")),
                                  np.clip(np.rint(waves[i].astype(np.float64) * 32767), -32768, 32767).astype(np.float32) / 32768)
            for c in range(2):
                got = read(tmp_path / "out" / ("%s_s%d.wav" % (stem, c + 1)))
                want = np.clip(np.rint(est[i, c, :len(waves[i])].astype(np.float64) * 32767), -32768, 32767) / 32768
                assert got.shape == (len(waves[i]),) and np.array_equal(got, want.astype(np.float32))


def test_evaluate_script_matches_per_utterance_numpy_metric(tmp_path, capsys):
    """The drop-in evaluate.py on ragged wav triples = the reference's loop: forward on the padded batch,
    PIT reordering, remove_pad, the oracle's cal_SISNRi (pinned to the reference's evaluate.py) per
    utterance."""
    import json
    import data
    import evaluate as ev
    import separate as sep
    from conv_tasnet import ConvTasNet
    from pit_criterion import cal_loss
    from oracle import convtasnet_oracle as O
    torch.manual_seed(5)
    model = ConvTasNet(64, 16, 32, 64, 3, 3, 2, 2).to(DEV).eval()
    torch.save(ConvTasNet.serialize(model, torch.optim.Adam(model.parameters()), 1), tmp_path / "model.pth.tar")
    rng = np.random.RandomState(6)
    lens = [4000, 2777, 3500, 3100, 1200]
    infos = {"mix": [], "s1": [], "s2": []}
    for i, n in enumerate(lens):
        s = 0.2 * rng.randn(2, n)
        for kind, x in (("s1", s[0]), ("s2", s[1]), ("mix", s.sum(0))):
            (tmp_path / kind).mkdir(exist_ok=True)
            path = str(tmp_path / kind / ("u%d.wav" % i))
            sep.write_wav(x, path, 8000)
            infos[kind].append([path, n])
    for kind, lst in infos.items():
        json.dump(lst, open(tmp_path / (kind + ".json"), "w"))
    args = ev._parser().parse_args(["--model_path", str(tmp_path / "model.pth.tar"), "--data_dir", str(tmp_path),
                                    "--batch_size", "2"])
    avg = ev.evaluate(args, verbose=False, num_workers=0)
    assert "Average SISNR improvement" in capsys.readouterr().out
    # reference loop, host padding, numpy metric
    ds = data.AudioDataset(str(tmp_path), 2, sample_rate=8000, segment=-1)
    vals = []
    for mix, ilens, src in data.AudioDataLoader(ds, batch_size=1):
        with torch.no_grad():
            est = model(mix.to(DEV))
            _, _, _, reorder = cal_loss(src.to(DEV), est, ilens.to(DEV))
        for b, n in enumerate(ilens.tolist()):
            vals.append(O.cal_SISNRi(src[b, :, :n].numpy(), reorder[b, :, :n].cpu().numpy(), mix[b, :n].numpy()))
    assert len(vals) == len(lens) and abs(avg - float(np.mean(vals))) < 1e-3
