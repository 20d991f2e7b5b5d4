"""Drop-in for the reference's src/data.py: datasets whose items are whole minibatches, and the collate
functions that turn them into padded tensors.

Same classes, arguments and return values as the reference (data.py:32-228):

    AudioDataset(json_dir, batch_size, sample_rate=8000, segment=4.0, cv_maxlen=8.0)
    AudioDataLoader(dataset, batch_size=1, ...)   -> (mixtures_pad [B,T], ilens [B], sources_pad [B,C,T])
    EvalDataset(mix_dir, mix_json, batch_size, sample_rate=8000)
    EvalDataLoader(dataset, batch_size=1, ...)    -> (mixtures_pad [B,T], ilens [B], filenames)

plus the packed route that the device-side batch assembly takes (ctn_b200/batch.py):

    AudioDataLoader(dataset, batch_size=1, packed=True, pin_memory=True)
        -> (mix_packed [sum T_b], ilens [B], src_packed [sum T_b, C])  host tensors, valid samples only
    ctn_b200.batch.assemble(mix_packed, ilens, src_packed, device)
        -> padded_mixture, mixture_lengths, padded_source on the device (what solver.py:173-176 uploads)

Which utterances share a minibatch and how they are cut into segments is the reference's logic
(ctn_b200/batch.py restates it, tests pin it to fixtures recorded from the reference).  Audio files are
read with the standard library's `wave` (PCM wav, down-mixed to mono, scaled to [-1, 1) like
librosa.load); the reference's resampling of files stored at another rate is not reproduced - such a
file raises.
"""
import json
import os
import wave

import numpy as np
import torch
import torch.utils.data as data

from ctn_b200 import batch as _batch


def _by_length(infos):
    # longest first; Python's sort is stable, so equal lengths keep their json order in all three lists
    return sorted(infos, key=lambda info: int(info[1]), reverse=True)


def _read_json(path):
    with open(path, "r") as f:
        return json.load(f)


class AudioDataset(data.Dataset):
    """Items are minibatches [mix_infos, s1_infos, s2_infos, sample_rate, segment_len] (data.py:32-118);
    an info is (wav_path, n_samples)."""

    def __init__(self, json_dir, batch_size, sample_rate=8000, segment=4.0, cv_maxlen=8.0):
        super().__init__()
        mix, s1, s2 = (_by_length(_read_json(os.path.join(json_dir, name + ".json"))) for name in ("mix", "s1", "s2"))
        lengths = [int(info[1]) for info in mix]

        def take(idx):
            return [mix[i] for i in idx], [s1[i] for i in idx], [s2[i] for i in idx]

        if segment >= 0.0:
            segment_len = int(segment * sample_rate)
            short = [n for n in lengths if n < segment_len]
            print("Drop {} utts({:.2f} h) which is short than {} samples".format(
                len(short), sum(short) / sample_rate / 36000, segment_len))   # the reference's own scale
            plan = _batch.plan_segment_minibatches(lengths, batch_size, segment_len)
            self.minibatch = [[*take(idx), sample_rate, segment_len] for idx in plan]
        else:
            plan = _batch.plan_full_minibatches(lengths, batch_size, cv_maxlen * sample_rate)
            self.minibatch = [[*take(idx), sample_rate, segment] for idx in plan]

    def __getitem__(self, index):
        return self.minibatch[index]

    def __len__(self):
        return len(self.minibatch)


class AudioDataLoader(data.DataLoader):
    """batch_size=1: a dataset item already is a minibatch (data.py:121-128).  packed=True yields the
    valid samples only, for ctn_b200.batch.assemble."""

    def __init__(self, *args, packed=False, **kwargs):
        super().__init__(*args, **kwargs)
        self.collate_fn = _collate_fn_packed if packed else _collate_fn


def _collate_fn(batch):
    """[minibatch] -> mixtures_pad [B, T], ilens [B] int64, sources_pad [B, C, T] (data.py:131-155)."""
    assert len(batch) == 1
    mixtures, sources = load_mixtures_and_sources(batch[0])
    ilens = torch.from_numpy(np.array([mix.shape[0] for mix in mixtures]))
    mixtures_pad = pad_list([torch.from_numpy(mix).float() for mix in mixtures], 0)
    sources_pad = pad_list([torch.from_numpy(s).float() for s in sources], 0)
    return mixtures_pad, ilens, sources_pad.permute((0, 2, 1)).contiguous()


def _collate_fn_packed(batch):
    """[minibatch] -> mix_packed [sum T_b], ilens [B], src_packed [sum T_b, C]: no padding on the host."""
    assert len(batch) == 1
    mixtures, sources = load_mixtures_and_sources(batch[0])
    return _batch.pack(mixtures, sources, pin=False)   # the loader pins (pin_memory=True); workers cannot


class EvalDataset(data.Dataset):
    """Items are [mix_infos, sample_rate] (data.py:161-200).  mix_dir: a directory of wav files, listed
    (and written to mix_dir/mix.json, as preprocess_one_dir does, preprocess.py:12-27) on the spot."""

    def __init__(self, mix_dir, mix_json, batch_size, sample_rate=8000):
        super().__init__()
        assert mix_dir != None or mix_json != None
        if mix_dir is not None:
            mix_json = os.path.join(mix_dir, "mix.json")
            _list_wav_dir(mix_dir, mix_json, sample_rate)
        infos = _by_length(_read_json(mix_json))
        plan = _batch.plan_full_minibatches([int(i[1]) for i in infos], batch_size)
        self.minibatch = [[[infos[i] for i in idx], sample_rate] for idx in plan]

    def __getitem__(self, index):
        return self.minibatch[index]

    def __len__(self):
        return len(self.minibatch)


class EvalDataLoader(data.DataLoader):
    """batch_size=1 (data.py:203-210)."""

    def __init__(self, *args, packed=False, **kwargs):
        super().__init__(*args, **kwargs)
        self.collate_fn = _collate_fn_eval_packed if packed else _collate_fn_eval


def _collate_fn_eval(batch):
    """[minibatch] -> mixtures_pad [B, T], ilens [B], filenames (data.py:213-232)."""
    assert len(batch) == 1
    mixtures, filenames = load_mixtures(batch[0])
    ilens = torch.from_numpy(np.array([mix.shape[0] for mix in mixtures]))
    return pad_list([torch.from_numpy(mix).float() for mix in mixtures], 0), ilens, filenames


def _collate_fn_eval_packed(batch):
    assert len(batch) == 1
    mixtures, filenames = load_mixtures(batch[0])
    mix_packed, ilens, _ = _batch.pack(mixtures, None, pin=False)
    return mix_packed, ilens, filenames


# ------------------------------ utils ------------------------------------
def load_mixtures_and_sources(batch):
    """minibatch -> (mixtures: B arrays [T], sources: B arrays [T, C]) after cutting every utterance
    into segments (data.py:236-271); B counts segments."""
    mixtures, sources = [], []
    mix_infos, s1_infos, s2_infos, sample_rate, segment_len = batch
    for mix_info, s1_info, s2_info in zip(mix_infos, s1_infos, s2_infos):
        assert mix_info[1] == s1_info[1] and s1_info[1] == s2_info[1]
        mix = _load_wav(mix_info[0], sample_rate)
        s = np.stack((_load_wav(s1_info[0], sample_rate), _load_wav(s2_info[0], sample_rate)), axis=1)  # T x C
        for lo, hi in _batch.segment_bounds(mix.shape[-1], segment_len):
            mixtures.append(mix[lo:hi])
            sources.append(s[lo:hi])
    return mixtures, sources


def load_mixtures(batch):
    """minibatch -> (mixtures: B arrays [T], filenames) (data.py:274-292)."""
    mix_infos, sample_rate = batch
    paths = [info[0] for info in mix_infos]
    return [_load_wav(p, sample_rate) for p in paths], paths


def pad_list(xs, pad_value):
    """List of tensors [T_b, ...] -> [B, max T_b, ...] filled with pad_value behind each item
    (data.py:295-301).  Host tensors; the device route is ctn_b200.batch.assemble."""
    out = xs[0].new_full((len(xs), max(x.size(0) for x in xs), *xs[0].size()[1:]), pad_value)
    for row, x in zip(out, xs):
        row[:x.size(0)] = x
    return out


def _load_wav(path, sample_rate):
    """PCM wav -> float32 mono in [-1, 1) (what librosa.load(path, sr=sample_rate) returns for a file
    stored at sample_rate, data.py:255-257)."""
    with wave.open(path, "rb") as w:
        rate, width, channels, frames = w.getframerate(), w.getsampwidth(), w.getnchannels(), w.getnframes()
        raw = w.readframes(frames)
    if rate != sample_rate:
        raise RuntimeError("%s is stored at %d Hz, %d Hz wanted: resample it beforehand" % (path, rate, sample_rate))
    if width == 2:
        x = np.frombuffer(raw, dtype="<i2").astype(np.float32) / 32768.0
    elif width == 4:
        x = (np.frombuffer(raw, dtype="<i4").astype(np.float64) / 2147483648.0).astype(np.float32)
    elif width == 1:
        x = (np.frombuffer(raw, dtype=np.uint8).astype(np.float32) - 128.0) / 128.0
    elif width == 3:
        b = np.frombuffer(raw, dtype=np.uint8).reshape(-1, 3).astype(np.int32)
        v = b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16)
        x = ((v ^ 0x800000) - 0x800000).astype(np.float32) / 8388608.0
    else:
        raise RuntimeError("%s: unsupported sample width %d" % (path, width))
    if channels > 1:
        x = x.reshape(-1, channels).mean(axis=1, dtype=np.float32)
    return x


def _list_wav_dir(in_dir, out_json, sample_rate):
    """[(absolute wav path, n_samples)] of a directory, sorted by name, written as json
    (preprocess.py:12-27)."""
    infos = []
    root = os.path.abspath(in_dir)
    for name in sorted(os.listdir(root)):
        if not name.endswith(".wav"):
            continue
        path = os.path.join(root, name)
        infos.append((path, len(_load_wav(path, sample_rate))))
    with open(out_json, "w") as f:
        json.dump(infos, f, indent=4)
    return infos
