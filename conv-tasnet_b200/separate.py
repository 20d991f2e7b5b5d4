#!/usr/bin/env python
"""Drop-in for the reference's src/separate.py: separates every mixture of a directory / json list with a
trained model and writes `<name>.wav` (the mixture) and `<name>_s<c>.wav` per source (separate.py:34-80).

Same command line.  What differs underneath: batches reach the GPU packed (valid samples only, pinned,
one copy), are padded on the device (ctn_b200/batch.py), the padding is taken off the estimates on the
device and every batch comes back in one copy; the next batch is loaded by a background worker while the
current one is separated.  Audio files are PCM16 wav written with the standard library (the reference
uses soundfile's 'PCM_16'): samples are scaled by 32767, rounded and saturated.
"""
import argparse
import os
import wave

import numpy as np
import torch

from conv_tasnet import ConvTasNet
from ctn_b200 import batch as _batch
from data import EvalDataLoader, EvalDataset
from utils import remove_pad

# the reference's options (separate.py:18-32), kept name for name
_OPTIONS = [
    ("--model_path", dict(type=str, required=True, help="model package written by training (ConvTasNet.serialize)")),
    ("--mix_dir", dict(type=str, default=None, help="directory of mixture wav files (wins over --mix_json)")),
    ("--mix_json", dict(type=str, default=None, help="json list of [wav path, samples] pairs")),
    ("--out_dir", dict(type=str, default="exp/result", help="where the separated wav files go")),
    ("--use_cuda", dict(type=int, default=0, help="accepted for compatibility: the model always runs on the GPU, "
                                                  "as in this fork of the reference (separate.py:44-46)")),
    ("--sample_rate", dict(type=int, default=8000, help="sample rate of the files")),
    ("--batch_size", dict(type=int, default=1, help="mixtures per forward pass")),
]
parser = argparse.ArgumentParser("Separate speech using Conv-TasNet")
for _flag, _kw in _OPTIONS:
    parser.add_argument(_flag, **_kw)


def write_wav(samples, filename, sample_rate):
    """float [-1, 1] -> mono PCM16 wav."""
    pcm = np.clip(np.rint(np.asarray(samples, dtype=np.float64) * 32767.0), -32768, 32767).astype("<i2")
    with wave.open(filename, "wb") as w:
        w.setnchannels(1)
        w.setsampwidth(2)
        w.setframerate(sample_rate)
        w.writeframes(pcm.tobytes())


def separate_batches(model, loader, device="cuda"):
    """Generator over a packed EvalDataLoader: yields (filenames, mixtures, estimates) per minibatch,
    mixtures a list of [len_b] and estimates a list of [C, len_b] numpy arrays, padding removed
    (separate.py:60-71)."""
    with torch.no_grad():
        for mix_packed, ilens, filenames in loader:
            mixture, lengths, _ = _batch.assemble(mix_packed, ilens, None, device)
            estimate_source = model(mixture)                       # [B, C, T]
            flat_estimate = remove_pad(estimate_source, ilens)     # one device-to-host copy
            off = np.concatenate([[0], np.cumsum(ilens.numpy())])
            host = mix_packed.numpy()
            yield filenames, [host[off[i]:off[i + 1]] for i in range(len(filenames))], flat_estimate


def separate(args, num_workers=1):
    if args.mix_dir is None and args.mix_json is None:
        print("Must provide mix_dir or mix_json! When providing mix_dir, mix_json is ignored.")
    model = ConvTasNet.load_model(args.model_path)
    print(model)
    model.eval()
    model.cuda()
    eval_dataset = EvalDataset(args.mix_dir, args.mix_json, batch_size=args.batch_size, sample_rate=args.sample_rate)
    eval_loader = EvalDataLoader(eval_dataset, batch_size=1, packed=True, pin_memory=True, num_workers=num_workers)
    os.makedirs(args.out_dir, exist_ok=True)
    for filenames, mixtures, estimates in separate_batches(model, eval_loader):
        for filename, mixture, estimate in zip(filenames, mixtures, estimates):
            # str.strip('.wav') strips CHARACTERS, not the suffix (separate.py:73): names that begin or end
            # with any of ". w a v" lose those characters in the reference; kept, so the files get the same names
            stem = os.path.join(args.out_dir, os.path.basename(filename).strip('.wav'))
            write_wav(mixture, stem + '.wav', args.sample_rate)
            for c in range(estimate.shape[0]):
                write_wav(estimate[c], stem + '_s{}.wav'.format(c + 1), args.sample_rate)


if __name__ == '__main__':
    args = parser.parse_args()
    print(args)
    separate(args)
