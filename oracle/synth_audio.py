"""Synthetic utterances for the data-pipeline fixtures (TEST INFRASTRUCTURE, like everything under
oracle/: only tests/ and oracle/make_golden.py import it).

A fake wav "path" such as `s1/u07_n0123.wav` names a deterministic float32 waveform of 123 samples; the
golden generator serves these through a stub `librosa.load` to the UNMODIFIED reference data.py, the tests
serve the same waveforms to the drop-in's loader, so that only the reference's batching / segmenting /
padding logic is under test."""
import re
import zlib

import numpy as np


def synth(path):
    n = int(re.search(r"_n(\d+)\.wav$", path).group(1))
    rng = np.random.RandomState(zlib.crc32(path.encode()) & 0x7FFFFFFF)
    return (0.1 * rng.randn(n)).astype(np.float32)


def make_infos(lengths, shuffle_seed=None):
    """(mix, s1, s2) info lists [(path, n_samples)] in the json order (optionally shuffled, the same way
    for the three lists)."""
    order = list(range(len(lengths)))
    if shuffle_seed is not None:
        np.random.RandomState(shuffle_seed).shuffle(order)
    out = []
    for kind in ("mix", "s1", "s2"):
        out.append([["%s/u%02d_n%04d.wav" % (kind, i, lengths[i]), lengths[i]] for i in order])
    return out


# dataset set-ups recorded in tests/golden/data_cases.*: name -> (utterance lengths, AudioDataset kwargs)
DATA_CASES = {
    # training: 40-sample segments, at most 3 segments per batch; shorter-than-segment utterances dropped,
    # a 5-segment utterance that can never fit, exact multiples and ragged tails
    "train_seg": dict(lengths=[200, 39, 120, 95, 81, 80, 40, 41, 119, 10, 79, 64], shuffle=3,
                      kw=dict(batch_size=3, sample_rate=100, segment=0.4)),
    "train_seg_b5": dict(lengths=[200, 161, 120, 95, 81, 80, 40, 41, 119, 44, 79, 64, 40, 40], shuffle=5,
                         kw=dict(batch_size=5, sample_rate=100, segment=0.4)),
    # validation: whole utterances, batches led by an utterance longer than cv_maxlen are skipped
    "cv_full": dict(lengths=[900, 850, 300, 120, 95, 81, 80, 40, 41, 119, 10], shuffle=7,
                    kw=dict(batch_size=2, sample_rate=100, segment=-1, cv_maxlen=8.0)),
    "cv_full_b4": dict(lengths=[120, 95, 81, 80, 40, 41, 119, 10, 33], shuffle=None,
                       kw=dict(batch_size=4, sample_rate=100, segment=-1, cv_maxlen=8.0)),
}
EVAL_CASES = {
    "eval_b3": dict(lengths=[120, 95, 81, 80, 40, 41, 119, 10], shuffle=9, kw=dict(batch_size=3, sample_rate=100)),
    "eval_b1": dict(lengths=[50, 70, 60], shuffle=None, kw=dict(batch_size=1, sample_rate=100)),
}
