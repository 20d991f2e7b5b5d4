"""Drop-in for the reference's src/utils.py (overlap_and_add, remove_pad)."""
import torch

from ctn_b200 import _native as nat


def overlap_and_add(signal, frame_step):
    """Reconstructs a signal from a framed representation (utils.py:9-46).

    signal: [..., frames, frame_length] float32 CUDA tensor; frames are offset by
    `frame_step` (<= frame_length).  Returns [..., (frames-1)*frame_step + frame_length].
    The reference scatters gcd-sized subframes with index_add_; here each output sample
    is one thread's sum of the frames covering it (csrc/decoder.cu), so the result is
    deterministic and needs no index tensor.
    """
    nat.require_cuda(signal, "signal")
    if signal.dtype != torch.float32:
        raise RuntimeError("overlap_and_add expects float32")
    outer = signal.size()[:-2]
    frames, frame_length = signal.size()[-2:]
    if frame_step <= 0 or frame_step > frame_length:
        raise ValueError("frame_step must be in (0, frame_length]")
    output_size = frame_step * (frames - 1) + frame_length
    sig = signal.contiguous().view(-1, frames, frame_length)
    out = torch.empty(sig.size(0), output_size, dtype=signal.dtype, device=signal.device)
    with torch.cuda.device(signal.device):
        nat.check(nat.lib().ctn_overlap_add(sig.data_ptr(), sig.size(0), frames, frame_length, int(frame_step),
                                            out.data_ptr(), nat.stream()))
    return out.view(*outer, output_size)


def remove_pad(inputs, inputs_lengths):
    """
    Args:
        inputs: torch.Tensor, [B, C, T] or [B, T], B is batch size
        inputs_lengths: torch.Tensor, [B]
    Returns:
        results: a list containing B items, each item is [C, T] (or [T]) numpy, T varies
    utils.py:49-66.  CUDA inputs: the valid samples are compacted on the device (ctn_remove_pad) and
    cross to the host in one copy (ctn_b200/batch.py); the reference copies utterance by utterance.
    Host inputs are only sliced.
    """
    dim = inputs.dim()
    if dim not in (2, 3):
        raise ValueError("inputs must be [B, C, T] or [B, T]")
    if inputs.is_cuda and inputs.dtype == torch.float32:
        from ctn_b200 import batch
        return batch.remove_pad(inputs, inputs_lengths)
    host = inputs.detach().cpu().numpy()
    lengths = [int(n) for n in torch.as_tensor(inputs_lengths).detach().cpu().tolist()]
    results = []
    for row, length in zip(host, lengths):
        results.append(row[:, :length].copy() if dim == 3 else row[:length].copy())
    return results
