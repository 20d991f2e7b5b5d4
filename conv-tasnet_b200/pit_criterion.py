"""Drop-in for the reference's src/pit_criterion.py: utterance-level SI-SNR with
permutation-invariant training.  Same functions, arguments and return tuples
(pit_criterion.py:12-113); the arithmetic runs in libctn_b200.so as a one-pass
reduction (csrc/loss.cu) instead of [B, C, C, T] broadcast temporaries and
per-utterance host syncs.
"""
from itertools import permutations

import torch

from ctn_b200 import _native as nat

EPS = 1e-8
_PERMS = {}


def _perms(C, device):
    """[C!, C] int64 in itertools.permutations order (pit_criterion.py:66), cached per device."""
    key = (C, device)
    p = _PERMS.get(key)
    if p is None:
        p = _PERMS[key] = torch.tensor(list(permutations(range(C))), dtype=torch.long, device=device)
    return p


def _check_inputs(source, estimate_source, source_lengths):
    assert source.size() == estimate_source.size()  # pit_criterion.py:34
    nat.require_cuda(source, "source")
    nat.require_cuda(estimate_source, "estimate_source")
    if source.dtype != torch.float32 or estimate_source.dtype != torch.float32:
        raise RuntimeError("cal_loss expects float32 source / estimate_source")
    if source.dim() != 3:
        raise RuntimeError("source must be [B, C, T]")
    lengths = source_lengths.to(device=source.device, dtype=torch.long).contiguous()
    return lengths


class _PitSiSnr(torch.autograd.Function):
    """max_snr [B,1], loss [] and max_idx [B] from (source, estimate); masks `estimate` in place."""

    @staticmethod
    def forward(ctx, source, estimate, lengths):
        B, C, T = source.shape
        dev = source.device
        lib = nat.lib()
        if not estimate.is_contiguous():
            raise RuntimeError("estimate_source must be contiguous (it is masked in place)")
        src = source.contiguous()
        perms = _perms(C, dev)
        ws = torch.empty(lib.ctn_pit_workspace_doubles(B, C, T), dtype=torch.float64, device=dev)
        max_snr = torch.empty(B, 1, dtype=torch.float32, device=dev)
        idx = torch.empty(B, dtype=torch.long, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        coef = torch.empty(B, C, 4, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            nat.check(lib.ctn_pit_sisnr_fwd(src.data_ptr(), estimate.data_ptr(), lengths.data_ptr(), B, C, T,
                                            perms.data_ptr(), perms.shape[0], ws.data_ptr(), max_snr.data_ptr(),
                                            idx.data_ptr(), loss.data_ptr(), None, coef.data_ptr(), nat.stream()))
        ctx.mark_dirty(estimate)
        ctx.mark_non_differentiable(idx)
        # unused outputs (max_snr, the masked estimate) arrive as None in backward instead of as
        # zero tensors, so the usual loss.backward() neither allocates them nor has to look at them
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(src, estimate, lengths, perms, idx, coef)
        return loss, max_snr, idx, estimate

    @staticmethod
    def backward(ctx, g_loss, g_max_snr, _g_idx, g_est):
        src, est, lengths, perms, idx, coef = ctx.saved_tensors
        B, C, T = src.shape
        lib = nat.lib()
        grad = torch.empty_like(est)
        # d loss/d est; max_snr's own gradient (rarely used) folds into the same form:
        # loss = -mean(max_snr)  =>  d/d max_snr[b] = -1/B, so upstream_b = g_loss - B * g_max_snr[b]
        up = None
        if g_loss is not None:
            up = g_loss.reshape(1).to(torch.float32).contiguous()
        if g_max_snr is not None and bool((g_max_snr != 0).any()):
            raise RuntimeError("gradient through max_snr is not supported; differentiate the loss")
        if up is None:
            up = torch.zeros(1, dtype=torch.float32, device=src.device)
        with torch.cuda.device(src.device):
            nat.check(lib.ctn_pit_sisnr_bwd(src.data_ptr(), est.data_ptr(), lengths.data_ptr(), perms.data_ptr(),
                                            idx.data_ptr(), coef.data_ptr(), up.data_ptr(), B, C, T,
                                            grad.data_ptr(), nat.stream()))
        if g_est is not None:
            # the masked estimate is also returned to the caller; pass its gradient through the mask
            t = torch.arange(T, device=src.device).view(1, 1, T)
            grad = grad + g_est * (t < lengths.view(B, 1, 1)).to(grad.dtype)
        return None, grad, None


def cal_loss(source, estimate_source, source_lengths):
    """
    Args:
        source: [B, C, T], B is batch size
        estimate_source: [B, C, T]
        source_lengths: [B]
    Returns (pit_criterion.py:24):
        loss, max_snr [B,1], estimate_source (the same tensor, masked in place),
        reorder_estimate_source [B, C, T]
    """
    lengths = _check_inputs(source, estimate_source, source_lengths)
    loss, max_snr, idx, estimate_source = _PitSiSnr.apply(source, estimate_source, lengths)
    perms = _perms(source.size(1), source.device)
    reorder_estimate_source = reorder_source(estimate_source, perms, idx)
    return loss, max_snr, estimate_source, reorder_estimate_source


def cal_si_snr_with_pit(source, estimate_source, source_lengths):
    """Calculate SI-SNR with PIT training.
    Args:
        source: [B, C, T], B is batch size
        estimate_source: [B, C, T]  (masked in place, pit_criterion.py:38)
        source_lengths: [B], each item is between [0, T]
    Returns:
        max_snr [B,1], perms [C!, C], max_snr_idx [B]
    """
    lengths = _check_inputs(source, estimate_source, source_lengths)
    _, max_snr, idx, _ = _PitSiSnr.apply(source, estimate_source, lengths)
    return max_snr, _perms(source.size(1), source.device), idx


def reorder_source(source, perms, max_snr_idx):
    """
    Args:
        source: [B, C, T]
        perms: [C!, C], permutations
        max_snr_idx: [B], each item is between [0, C!)
    Returns:
        reorder_source: [B, C, T], out[b, c] = source[b, perms[max_snr_idx[b]][c]]
    """
    nat.require_cuda(source, "source")
    B, C, T = source.size()
    src = source.detach().contiguous()
    if src.dtype != torch.float32:
        raise RuntimeError("reorder_source expects float32")
    out = torch.empty_like(src)
    with torch.cuda.device(src.device):
        nat.check(nat.lib().ctn_pit_reorder(src.data_ptr(), perms.contiguous().data_ptr(),
                                            max_snr_idx.contiguous().data_ptr(), B, C, T, out.data_ptr(),
                                            nat.stream()))
    return out


def get_mask(source, source_lengths):
    """
    Args:
        source: [B, C, T]
        source_lengths: [B]
    Returns:
        mask: [B, 1, T], 1 for t < length else 0 (pit_criterion.py:101-113), built without host syncs
    """
    B, _, T = source.size()
    t = torch.arange(T, device=source.device).view(1, 1, T)
    return (t < source_lengths.to(source.device).view(B, 1, 1)).to(source.dtype)
