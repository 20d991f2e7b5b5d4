"""Device-side evaluation metrics: the SI-SNR improvement that the reference's src/evaluate.py
computes per utterance in numpy on the host (evaluate.py:66-80, 108-144), here for a whole padded
batch in one kernel, plus the evaluation step that feeds it.

Same function names and argument meaning as the reference module for the SI-SNR functions.
`cal_SDRi` (evaluate.py:88-105) is mir_eval's bss_eval_sources (a third-party host routine the
reference calls as a black box): out of scope, it raises.
"""
import numpy as np
import torch

from ctn_b200 import _native as nat


def batch_SISNRi(padded_source, reorder_estimate_source, padded_mixture, mixture_lengths=None):
    """SI-SNRi of every utterance of a padded batch, on the device, no host synchronisation.

    padded_source, reorder_estimate_source: float32 CUDA [B, C, T]; padded_mixture: [B, T];
    mixture_lengths: int [B] or None (= T).  Samples at t >= length are ignored, exactly as if the
    utterances had gone through remove_pad (evaluate.py:66-71).
    Returns (sisnri [B], sisnr [B, C, 2]) with sisnr[..., 0] the estimate's SI-SNR and
    sisnr[..., 1] the mixture's (the baseline), in dB.
    """
    for t, what in ((padded_source, "padded_source"), (reorder_estimate_source, "reorder_estimate_source"),
                    (padded_mixture, "padded_mixture")):
        nat.require_cuda(t, what)
        if t.dtype != torch.float32:
            raise RuntimeError("%s must be float32" % what)
    if padded_source.dim() != 3 or padded_source.size() != reorder_estimate_source.size():
        raise RuntimeError("source and estimate must both be [B, C, T]")
    B, C, T = padded_source.shape
    if tuple(padded_mixture.shape) != (B, T):
        raise RuntimeError("mixture must be [B, T] = [%d, %d]" % (B, T))
    dev = padded_source.device
    lengths = None
    if mixture_lengths is not None:
        lengths = mixture_lengths.to(device=dev, dtype=torch.int64).contiguous()
    sisnr = torch.empty(B, C, 2, dtype=torch.float32, device=dev)
    sisnri = torch.empty(B, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        nat.check(nat.lib().ctn_sisnri(padded_source.contiguous().data_ptr(),
                                       reorder_estimate_source.contiguous().data_ptr(),
                                       padded_mixture.contiguous().data_ptr(), nat.ptr(lengths), B, C, T,
                                       sisnr.data_ptr(), sisnri.data_ptr(), nat.stream()))
    return sisnri, sisnr


def _dev(x):
    if isinstance(x, np.ndarray):
        if not torch.cuda.is_available():
            raise RuntimeError("a CUDA device is required: this implementation has no CPU path")
        return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).cuda(), True
    nat.require_cuda(x, "signal")
    return x.float(), False


def cal_SISNRi(src_ref, src_est, mix):
    """evaluate.py:108-125.  src_ref, src_est: [C, T]; mix: [T] (numpy arrays, as the reference
    passes, or CUDA tensors).  Returns average_SISNRi: a Python float for numpy inputs, a 0-dim CUDA
    tensor for tensor inputs.  Averages over all C sources (the reference hard-codes two)."""
    r, was_np = _dev(src_ref)
    e, _ = _dev(src_est)
    m, _ = _dev(mix)
    sisnri, _ = batch_SISNRi(r[None], e[None], m[None])
    return float(sisnri[0]) if was_np else sisnri[0]


def cal_SISNR(ref_sig, out_sig, eps=1e-8):
    """evaluate.py:128-144.  ref_sig, out_sig: [T].  Returns the SI-SNR in dB."""
    if eps != 1e-8:
        raise ValueError("eps is fixed at 1e-8 in the kernel (the reference's default)")
    r, was_np = _dev(ref_sig)
    o, _ = _dev(out_sig)
    assert r.numel() == o.numel()
    _, sisnr = batch_SISNRi(r[None, None], o[None, None], o[None])
    return float(sisnr[0, 0, 0]) if was_np else sisnr[0, 0, 0]


def cal_SDRi(src_ref, src_est, mix):
    raise NotImplementedError("SDRi is mir_eval.separation.bss_eval_sources (evaluate.py:88-105), a host-side "
                              "third-party routine; call mir_eval on remove_pad() outputs if it is installed")


@torch.no_grad()
def evaluate_batch(model, padded_mixture, mixture_lengths, padded_source):
    """One iteration of the reference's evaluation loop (evaluate.py:53-80) without leaving the device:
    forward, PIT loss / reordering, SI-SNRi per utterance.  Returns a dict of CUDA tensors
    {loss [], max_snr [B,1], reorder_estimate_source [B,C,T], sisnri [B], sisnr [B,C,2]}."""
    from pit_criterion import cal_loss
    estimate_source = model(padded_mixture)
    loss, max_snr, estimate_source, reorder_estimate_source = cal_loss(padded_source, estimate_source,
                                                                       mixture_lengths)
    sisnri, sisnr = batch_SISNRi(padded_source, reorder_estimate_source, padded_mixture, mixture_lengths)
    return {"loss": loss, "max_snr": max_snr, "reorder_estimate_source": reorder_estimate_source,
            "sisnri": sisnri, "sisnr": sisnr}


def _parser():
    import argparse
    parser = argparse.ArgumentParser('Evaluate separation performance using Conv-TasNet')
    parser.add_argument('--model_path', type=str, required=True, help='Path to model file created by training')
    parser.add_argument('--data_dir', type=str, required=True,
                        help='directory including mix.json, s1.json and s2.json')
    parser.add_argument('--cal_sdr', type=int, default=0,
                        help='SDR needs mir_eval (third-party, host): must stay 0 here')
    parser.add_argument('--use_cuda', type=int, default=0,
                        help='Accepted for compatibility; the model always runs on the GPU (evaluate.py:41-43)')
    parser.add_argument('--sample_rate', default=8000, type=int, help='Sample rate')
    parser.add_argument('--batch_size', default=1, type=int, help='Batch size')
    return parser


def evaluate(args, verbose=True, num_workers=2):
    """The reference's evaluation script (evaluate.py:33-84, same command line and printed lines): whole
    utterances of data_dir, forward, PIT reordering, SI-SNRi per utterance and its average.  Batches arrive
    packed and are padded on the device; the metric is formed on the device for the padded batch
    (ctn_sisnri) and only B floats per batch come back, instead of remove_pad of three tensors and numpy
    arithmetic per utterance.  Returns the average SI-SNRi."""
    from conv_tasnet import ConvTasNet
    from ctn_b200 import batch as _batch
    from data import AudioDataLoader, AudioDataset
    if args.cal_sdr:
        cal_SDRi(None, None, None)
    model = ConvTasNet.load_model(args.model_path)
    if verbose:
        print(model)
    model.eval()
    model.cuda()
    dataset = AudioDataset(args.data_dir, args.batch_size, sample_rate=args.sample_rate, segment=-1)
    data_loader = AudioDataLoader(dataset, batch_size=1, num_workers=num_workers, packed=True, pin_memory=True)
    total_SISNRi, total_cnt = 0.0, 0
    for mix_packed, ilens, src_packed in data_loader:
        padded_mixture, mixture_lengths, padded_source = _batch.assemble(mix_packed, ilens, src_packed, "cuda")
        out = evaluate_batch(model, padded_mixture, mixture_lengths, padded_source)
        for avg_SISNRi in out["sisnri"].tolist():
            if verbose:
                print("Utt", total_cnt + 1)
                print("\tSI-SNRi={0:.2f}".format(avg_SISNRi))
            total_SISNRi += avg_SISNRi
            total_cnt += 1
    print("Average SISNR improvement: {0:.2f}".format(total_SISNRi / total_cnt))
    return total_SISNRi / total_cnt


if __name__ == '__main__':
    args = _parser().parse_args()
    print(args)
    evaluate(args)
