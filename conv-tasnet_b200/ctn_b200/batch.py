"""Batch assembly either side of the separator: which utterances go into a minibatch (host logic of
data.py:52-112, :188-200), how ragged utterances become the padded tensors the model takes (pad_list,
data.py:295-301, as used by the collate functions :146-154 and :222-225) and how the padding is taken off
the results again (utils.py:49-66).

The reference pads on the host, uploads the padded tensors and downloads every separated utterance with
its own `.cpu()` call.  Here the valid samples travel packed - one host-to-device copy per batch, one
device-to-host copy of exactly the valid bytes - and `ctn_pad_batch` / `ctn_remove_pad` (csrc/batch.cu)
produce / undo the padded layout on the device.  The padded tensors are what the reference's model sees,
zeros included (gLN statistics count the padded frames, conv_tasnet.py:352-353), so results do not change.
"""
import math

import numpy as np
import torch

from . import _native as nat


# ------------------------------------------------------------------ host-side planning
def segment_bounds(utt_len, segment_len):
    """[start, end) of the training segments cut from one utterance (data.py:260-267): consecutive
    windows of segment_len samples, plus - when the length is not a multiple - the LAST segment_len
    samples (overlapping the previous window).  segment_len < 0: the whole utterance (data.py:268-270).
    Utterances shorter than a segment yield nothing from the loop but the trailing window still applies,
    exactly as in the reference (its dataset drops those utterances beforehand, data.py:77)."""
    if segment_len < 0:
        return [(0, utt_len)]
    bounds = [(s, s + segment_len) for s in range(0, utt_len - segment_len + 1, segment_len)]
    if utt_len % segment_len != 0:
        bounds.append((max(utt_len - segment_len, 0), utt_len))
    return bounds


def plan_segment_minibatches(lengths, batch_size, segment_len):
    """Groups utterances (given by their sample counts, sorted longest first as data.py:52-57 does) into
    minibatches of at most batch_size SEGMENTS (data.py:71-95).  Returns lists of positions into `lengths`.
    Reference behaviour kept: utterances shorter than one segment are skipped (:77); an utterance that
    would overflow the batch closes it and opens the next one (:80-84); an utterance with more segments
    than batch_size on its own is dropped (:82-83)."""
    batches = []
    n = len(lengths)
    pos = 0
    while True:
        first = pos
        segments = 0
        chosen = []
        while segments < batch_size and pos < n:
            utt_len = int(lengths[pos])
            if utt_len >= segment_len:
                segments += math.ceil(utt_len / segment_len)
                if segments > batch_size:
                    if pos == first:
                        pos += 1
                    break
                chosen.append(pos)
            pos += 1
        if chosen:
            batches.append(chosen)
        if pos == n:
            return batches


def plan_full_minibatches(lengths, batch_size, max_len=None):
    """Whole-utterance minibatches of batch_size consecutive utterances of the length-sorted list
    (data.py:96-112 for validation with max_len = cv_maxlen * sample_rate, :188-200 for evaluation with
    no limit).  A batch whose FIRST (longest) utterance exceeds max_len is skipped as a whole (:102-104).
    When every remaining utterance is too long the reference indexes past the end (IndexError); so does
    this."""
    batches = []
    n = len(lengths)
    start = 0
    while True:
        end = min(n, start + batch_size)
        if max_len is not None and int(lengths[start]) > max_len:   # IndexError when start == n, as the reference
            start = end
            continue
        batches.append(list(range(start, end)))
        if end == n:
            return batches
        start = end


# ------------------------------------------------------------------ packed batches
def pack(mixtures, sources=None, pin=True):
    """Packs a ragged batch for one upload.
    mixtures: list of B float arrays [T_b]; sources: list of B arrays [T_b, C] (as
    load_mixtures_and_sources returns them, data.py:246-271) or None.
    Returns (mix_packed [sum T_b], ilens int64 [B], src_packed [sum T_b, C] or None) as CPU tensors,
    in pinned memory when `pin`."""
    ilens = np.array([int(np.shape(m)[0]) for m in mixtures], dtype=np.int64)
    total = int(ilens.sum())

    def _buf(*shape):
        t = torch.empty(*shape, dtype=torch.float32)
        return t.pin_memory() if pin and torch.cuda.is_available() else t

    mix = _buf(total)
    mv = mix.numpy()
    off = 0
    for m, n in zip(mixtures, ilens):
        mv[off:off + n] = np.asarray(m, dtype=np.float32).reshape(-1)
        off += n
    src = None
    if sources is not None:
        if len(sources) != len(mixtures):
            raise ValueError("as many source arrays as mixtures are needed")
        C = int(np.shape(sources[0])[1]) if len(sources) else 0
        src = _buf(total, C)
        sv = src.numpy()
        off = 0
        for s, n in zip(sources, ilens):
            s = np.asarray(s, dtype=np.float32)
            if s.shape != (n, C):
                raise ValueError("sources must be [T_b, C] with T_b of the mixture; got %r for T_b=%d, C=%d"
                                 % (s.shape, n, C))
            sv[off:off + n] = s
            off += n
    return mix, torch.from_numpy(ilens), src


def _offsets(ilens, device):
    """int64 [B + 1] exclusive prefix sum on the device, from HOST lengths (one small upload)."""
    host = torch.zeros(ilens.numel() + 1, dtype=torch.int64)
    if ilens.numel():
        if int(ilens.min()) < 0:
            raise ValueError("negative utterance length")
        torch.cumsum(ilens.to(torch.int64).cpu(), 0, out=host[1:])
    return host, host.to(device, non_blocking=True)


def assemble(mix_packed, ilens, src_packed=None, device="cuda", T=None, pad_value=0.0):
    """Packed host batch -> the tensors the reference's collate + `.cuda()` produce
    (data.py:146-155, solver.py:173-176):
        padded_mixture [B, T] float32, mixture_lengths [B] int64, padded_source [B, C, T] float32 (or None)
    on `device`, T = the longest utterance unless given.  One upload of the valid samples per tensor and
    one launch; asynchronous on the current stream."""
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("assemble() places batches on a CUDA device: this implementation has no CPU path")
    if ilens.is_cuda:
        raise ValueError("ilens must be a host tensor (it sizes the batch)")
    B = int(ilens.numel())
    host_off, dev_off = _offsets(ilens, device)
    total = int(host_off[-1])
    if mix_packed is not None and mix_packed.numel() != total:
        raise ValueError("mix_packed holds %d samples, the lengths add up to %d" % (mix_packed.numel(), total))
    if src_packed is not None and (src_packed.dim() != 2 or src_packed.size(0) != total):
        raise ValueError("src_packed must be [sum of lengths, C]")
    if T is None:
        T = int(ilens.max()) if B else 0
    C = int(src_packed.size(1)) if src_packed is not None else 0
    mix_d = src_d = mix_pad = src_pad = None
    with torch.cuda.device(device):
        if mix_packed is not None:
            mix_d = mix_packed.to(device=device, dtype=torch.float32, non_blocking=True).contiguous()
            mix_pad = torch.empty(B, T, dtype=torch.float32, device=device)
        if src_packed is not None:
            src_d = src_packed.to(device=device, dtype=torch.float32, non_blocking=True).contiguous()
            src_pad = torch.empty(B, C, T, dtype=torch.float32, device=device)
        if mix_pad is None and src_pad is None:
            raise ValueError("nothing to assemble")
        if B == 0 or T == 0 or total == 0:    # empty tensors have no storage to point at
            for t in (mix_pad, src_pad):
                if t is not None:
                    t.fill_(pad_value)
            return mix_pad, dev_off[1:] - dev_off[:-1], src_pad
        nat.check(nat.lib().ctn_pad_batch(
            mix_d.data_ptr() if mix_d is not None else None, src_d.data_ptr() if src_d is not None else None,
            dev_off.data_ptr(), B, C, T, float(pad_value),
            mix_pad.data_ptr() if mix_pad is not None else None,
            src_pad.data_ptr() if src_pad is not None else None, nat.stream()))
    lengths = dev_off[1:] - dev_off[:-1]
    return mix_pad, lengths, src_pad


def remove_pad(inputs, inputs_lengths):
    """utils.py:49-66 on the device: inputs [B, C, T] or [B, T] CUDA float32, lengths [B] (host or
    device).  Returns a list of B numpy arrays [C, len_b] (or [len_b]).  The valid samples are compacted
    by one launch and cross to the host in ONE copy (the reference issues one per utterance and moves
    them through pageable memory); the arrays are views of that one host buffer."""
    nat.require_cuda(inputs, "inputs")
    dim = inputs.dim()
    if dim not in (2, 3):
        raise ValueError("inputs must be [B, C, T] or [B, T]")
    if inputs.dtype != torch.float32:
        raise RuntimeError("remove_pad expects float32")
    x = inputs.detach().contiguous()
    B, T = x.size(0), x.size(-1)
    C = x.size(1) if dim == 3 else 1
    ilens = torch.as_tensor(inputs_lengths).detach().to("cpu", torch.int64).clamp(max=T)
    if ilens.numel() != B:
        raise ValueError("one length per utterance is needed")
    host_off, dev_off = _offsets(ilens, x.device)
    total = int(host_off[-1])
    with torch.cuda.device(x.device):
        packed = torch.empty(C * total, dtype=torch.float32, device=x.device)
        if total > 0:
            nat.check(nat.lib().ctn_remove_pad(x.data_ptr(), dev_off.data_ptr(), B, C, T, packed.data_ptr(), nat.stream()))
        host = torch.empty(C * total, dtype=torch.float32).pin_memory()
        host.copy_(packed, non_blocking=True)
        torch.cuda.current_stream(x.device).synchronize()
    flat = host.numpy()
    results = []
    for b in range(B):
        lo, n = C * int(host_off[b]), int(ilens[b])
        block = flat[lo:lo + C * n]
        results.append(block.reshape(C, n) if dim == 3 else block)
    return results
