#!/usr/bin/env python
"""Drop-in for the reference's src/preprocess.py: lists the wav files of the wsj0-mix directory tree with their
lengths in samples, one json per (tr | cv | tt) x (mix | s1 | s2), the input of data.AudioDataset
(preprocess.py:12-36).  Same functions and command line; files are opened with the standard library
(data._load_wav: PCM wav at the requested rate, no resampling) instead of librosa, and listed in name order
(the reference uses os.listdir order, which differs between file systems - and must agree between the
mix / s1 / s2 directories for the length-sorted lists of data.py to line up)."""
import argparse
import os

from data import _list_wav_dir

SETS = ("tr", "cv", "tt")
SPEAKERS = ("mix", "s1", "s2")


def preprocess_one_dir(in_dir, out_dir, out_filename, sample_rate=8000):
    """[(absolute wav path, samples), ...] of in_dir -> out_dir/out_filename.json"""
    os.makedirs(out_dir, exist_ok=True)
    return _list_wav_dir(in_dir, os.path.join(out_dir, out_filename + ".json"), sample_rate)


def preprocess(args):
    for data_type in SETS:
        for speaker in SPEAKERS:
            preprocess_one_dir(os.path.join(args.in_dir, data_type, speaker), os.path.join(args.out_dir, data_type),
                               speaker, sample_rate=args.sample_rate)


if __name__ == "__main__":
    parser = argparse.ArgumentParser("WSJ0 data preprocessing")
    for flag, kw in (("--in-dir", dict(type=str, default=None, help="wsj0-mix directory holding tr, cv and tt")),
                     ("--out-dir", dict(type=str, default=None, help="where the json lists go")),
                     ("--sample-rate", dict(type=int, default=8000, help="sample rate of the audio files"))):
        parser.add_argument(flag, **kw)
    args = parser.parse_args()
    print(args)
    preprocess(args)
