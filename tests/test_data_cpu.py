"""Batch planning, segmenting, padding and the collate functions of the drop-in data.py /
ctn_b200.batch against fixtures recorded from the UNMODIFIED reference data.py
(oracle/make_golden.py make_data_cases; data.py:32-232, :295-301).  Host logic only: runs without a GPU."""
import json
import os
import sys
import wave

import numpy as np
import pytest
import torch

from conftest import GOLDEN, ROOT

sys.path.insert(0, os.path.join(ROOT, "oracle"))
import synth_audio as sa  # noqa: E402  (test infrastructure: the waveforms the fixtures were recorded on)

PLAN = json.load(open(os.path.join(GOLDEN, "data_cases.json")))
ARR = np.load(os.path.join(GOLDEN, "data_cases.npz"))


@pytest.fixture()
def data_mod(monkeypatch):
    import data
    monkeypatch.setattr(data, "_load_wav", lambda path, sample_rate: sa.synth(path))
    return data


def _write_jsons(tmp_path, infos):
    for kind, lst in zip(("mix", "s1", "s2"), infos):
        with open(tmp_path / (kind + ".json"), "w") as f:
            json.dump(lst, f)


@pytest.mark.parametrize("name", sorted(sa.DATA_CASES))
def test_audio_dataset_minibatches_and_collate_match_reference(name, data_mod, tmp_path):
    case = sa.DATA_CASES[name]
    _write_jsons(tmp_path, sa.make_infos(case["lengths"], case["shuffle"]))
    ds = data_mod.AudioDataset(str(tmp_path), **case["kw"])
    got = [[[i[0] for i in part] for part in mb[:3]] + [mb[3], mb[4]] for mb in ds.minibatch]
    assert got == PLAN[name]
    assert len(ds) == len(PLAN[name])
    loader = data_mod.AudioDataLoader(ds, batch_size=1, shuffle=False, num_workers=0)
    packed = data_mod.AudioDataLoader(ds, batch_size=1, shuffle=False, num_workers=0, packed=True)
    for b, ((mix, ilens, src), (mix_p, ilens_p, src_p)) in enumerate(zip(loader, packed)):
        assert np.array_equal(mix.numpy(), ARR["%s.%d.mix" % (name, b)])
        assert np.array_equal(ilens.numpy(), ARR["%s.%d.ilens" % (name, b)])
        assert np.array_equal(src.numpy(), ARR["%s.%d.src" % (name, b)])
        assert mix.dtype == torch.float32 and src.dtype == torch.float32 and ilens.dtype == torch.int64
        # the packed route carries the same samples: padding it on the host gives the same tensors
        assert torch.equal(ilens_p, ilens)
        off = np.concatenate([[0], np.cumsum(ilens.numpy())])
        assert mix_p.shape == (off[-1],) and src_p.shape == (off[-1], src.shape[1])
        for i in range(len(ilens)):
            n = int(ilens[i])
            assert torch.equal(mix_p[off[i]:off[i] + n], mix[i, :n])
            assert torch.equal(src_p[off[i]:off[i] + n].t(), src[i, :, :n])
            assert not mix[i, n:].any() and not src[i, :, n:].any()


@pytest.mark.parametrize("name", sorted(sa.EVAL_CASES))
def test_eval_dataset_minibatches_and_collate_match_reference(name, data_mod, tmp_path):
    case = sa.EVAL_CASES[name]
    with open(tmp_path / "mix.json", "w") as f:
        json.dump(sa.make_infos(case["lengths"], case["shuffle"])[0], f)
    ds = data_mod.EvalDataset(None, str(tmp_path / "mix.json"), **case["kw"])
    assert [[[i[0] for i in mb[0]], mb[1]] for mb in ds.minibatch] == PLAN[name]
    for b, (mix, ilens, names) in enumerate(data_mod.EvalDataLoader(ds, batch_size=1)):
        assert names == PLAN[name][b][0]
        assert np.array_equal(mix.numpy(), ARR["%s.%d.mix" % (name, b)])
        assert np.array_equal(ilens.numpy(), ARR["%s.%d.ilens" % (name, b)])
    for b, (mix_p, ilens, names) in enumerate(data_mod.EvalDataLoader(ds, batch_size=1, packed=True)):
        ref = ARR["%s.%d.mix" % (name, b)]
        assert np.array_equal(mix_p.numpy(), np.concatenate([ref[i, :n] for i, n in enumerate(ilens.tolist())]))


def test_pad_list_matches_reference():
    import data
    xs = [torch.arange(6.).view(3, 2), torch.arange(10.).view(5, 2), torch.arange(2.).view(1, 2)]
    assert np.array_equal(data.pad_list(xs, -1.5).numpy(), ARR["pad_list.out"])


def test_segment_bounds_and_plans_edge_cases():
    from ctn_b200 import batch
    assert batch.segment_bounds(120, 40) == [(0, 40), (40, 80), (80, 120)]
    assert batch.segment_bounds(121, 40) == [(0, 40), (40, 80), (80, 120), (81, 121)]
    assert batch.segment_bounds(39, 40) == [(0, 39)]          # x[-40:] of a 39-sample array is all of it
    assert batch.segment_bounds(77, -1) == [(0, 77)]
    assert batch.plan_segment_minibatches([], 3, 40) == []
    assert batch.plan_segment_minibatches([10, 5], 3, 40) == []            # everything too short
    assert batch.plan_segment_minibatches([500, 40], 3, 40) == [[1]]      # 13 segments never fit: dropped
    assert batch.plan_segment_minibatches([80, 80, 40], 3, 40) == [[0], [1, 2]]
    assert batch.plan_full_minibatches([9, 8, 7], 2) == [[0, 1], [2]]
    assert batch.plan_full_minibatches([900, 8, 7], 2, 800) == [[2]]       # batch led by a long one is skipped whole
    with pytest.raises(IndexError):                                        # as the reference: nothing short enough
        batch.plan_full_minibatches([900, 850], 2, 800)


def test_wav_reader_and_directory_listing(tmp_path):
    import data
    rng = np.random.RandomState(0)
    pcm = (rng.randn(3, 321) * 8000).astype("<i2")
    for i, name in enumerate(["b.wav", "a.wav", "c.wav"]):
        with wave.open(str(tmp_path / name), "wb") as w:
            w.setnchannels(1), w.setsampwidth(2), w.setframerate(8000)
            w.writeframes(pcm[i, :321 - 50 * i].tobytes())
    (tmp_path / "notes.txt").write_text("not audio")
    x = data._load_wav(str(tmp_path / "a.wav"), 8000)
    assert x.dtype == np.float32 and np.array_equal(x, pcm[1, :271].astype(np.float32) / 32768.0)
    with pytest.raises(RuntimeError, match="resample"):
        data._load_wav(str(tmp_path / "a.wav"), 16000)
    with wave.open(str(tmp_path / "st.wav"), "wb") as w:      # stereo: down-mixed like librosa's mono=True
        w.setnchannels(2), w.setsampwidth(2), w.setframerate(8000)
        w.writeframes(np.stack([pcm[0, :100], pcm[1, :100]], 1).tobytes())
    st = data._load_wav(str(tmp_path / "st.wav"), 8000)
    assert np.allclose(st, (pcm[0, :100].astype(np.float32) + pcm[1, :100]) / 65536.0, atol=1e-7)
    os.remove(tmp_path / "st.wav")
    ds = data.EvalDataset(str(tmp_path), None, batch_size=2, sample_rate=8000)
    listed = json.load(open(tmp_path / "mix.json"))
    assert [os.path.basename(p) for p, _ in listed] == ["a.wav", "b.wav", "c.wav"]
    assert [n for _, n in listed] == [271, 321, 221]
    assert [[os.path.basename(i[0]) for i in mb[0]] for mb in ds.minibatch] == [["b.wav", "a.wav"], ["c.wav"]]
    mix, ilens, names = next(iter(data.EvalDataLoader(ds, batch_size=1)))
    assert mix.shape == (2, 321) and ilens.tolist() == [321, 271] and not mix[1, 271:].any()


def test_pack_rejects_mismatched_sources():
    from ctn_b200 import batch
    with pytest.raises(ValueError):
        batch.pack([np.zeros(5, np.float32)], [np.zeros((4, 2), np.float32)], pin=False)
    with pytest.raises(ValueError):
        batch.pack([np.zeros(5, np.float32)], [], pin=False)
    mix, ilens, src = batch.pack([], None, pin=False)
    assert mix.numel() == 0 and ilens.numel() == 0 and src is None


def test_preprocess_lists_the_directory_tree(tmp_path):
    """preprocess.py:12-36: one json per set and speaker with (absolute path, samples); the result feeds
    AudioDataset directly."""
    import types
    import data
    import preprocess
    rng = np.random.RandomState(1)
    lens = [300, 500, 410]
    for s in preprocess.SETS:
        for spk in preprocess.SPEAKERS:
            d = tmp_path / "wav" / s / spk
            d.mkdir(parents=True)
            for i, n in enumerate(lens):
                with wave.open(str(d / ("u%d.wav" % i)), "wb") as w:
                    w.setnchannels(1), w.setsampwidth(2), w.setframerate(8000)
                    w.writeframes((rng.randn(n) * 3000).astype("<i2").tobytes())
    preprocess.preprocess(types.SimpleNamespace(in_dir=str(tmp_path / "wav"), out_dir=str(tmp_path / "json"),
                                                sample_rate=8000))
    for s in preprocess.SETS:
        for spk in preprocess.SPEAKERS:
            infos = json.load(open(tmp_path / "json" / s / (spk + ".json")))
            assert [n for _, n in infos] == lens
            assert all(os.path.isabs(p) and p.endswith("u%d.wav" % i) and ("/%s/%s/" % (s, spk)) in p
                       for i, (p, _) in enumerate(infos))
    ds = data.AudioDataset(str(tmp_path / "json" / "tr"), batch_size=2, sample_rate=8000, segment=-1)
    mix, ilens, src = next(iter(data.AudioDataLoader(ds, batch_size=1)))
    assert ilens.tolist() == [500, 410] and mix.shape == (2, 500) and src.shape == (2, 2, 500)
