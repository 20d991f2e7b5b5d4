"""Generate golden fixtures from the UNMODIFIED reference (build container only).

Run:  PYTHONDONTWRITEBYTECODE=1 python oracle/make_golden.py
Needs /root/reference (read-only mount); the GPU box does not have it, which
is why the outputs are committed under tests/golden/.  The only harness-side
accommodation is the `.cuda()` identity shim: utils.py:40 hard-codes `.cuda()`
on the overlap-add index tensor, which raises on a CPU-only host
(SURVEY.md §0 fact 4).  No reference file is edited or copied.
"""
import os
import sys

import numpy as np
import torch

REF = os.environ.get("CTN_REFERENCE", "/root/reference/src")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def _import_reference():
    sys.dont_write_bytecode = True
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self  # harness shim, see docstring
    sys.path.insert(0, REF)
    import conv_tasnet as ref_model
    import pit_criterion as ref_loss
    import utils as ref_utils
    sys.path.pop(0)
    return ref_model, ref_loss, ref_utils


CASES = {
    # the reference's own demo tuple (conv_tasnet.py:360-363), float inputs
    "demo_gln": dict(cfg=dict(N=3, L=4, B=2, H=3, P=3, X=3, R=2, C=2, norm_type="gLN",
                              causal=False, mask_nonlinear="relu"), M=2, T=12, lengths=[12, 12]),
    # causal cLN, T not a multiple of hop (tail zero-fill), ragged lengths
    "causal_cln": dict(cfg=dict(N=8, L=4, B=8, H=16, P=3, X=3, R=2, C=2, norm_type="cLN",
                                causal=True, mask_nonlinear="relu"), M=2, T=51, lengths=[51, 37]),
    # C=3 softmax mask, gLN, ragged lengths
    "softmax_c3": dict(cfg=dict(N=16, L=8, B=8, H=16, P=3, X=2, R=2, C=3, norm_type="gLN",
                                causal=False, mask_nonlinear="softmax"), M=3, T=100,
                       lengths=[100, 77, 64]),
    # smallest shape that takes the tensor-core GEMM path (channels % 8 == 0, K > 128)
    "tc_small": dict(cfg=dict(N=64, L=16, B=64, H=128, P=3, X=3, R=1, C=2, norm_type="gLN",
                              causal=False, mask_nonlinear="relu"), M=2, T=1207, lengths=[1207, 900]),
    # causal cLN on the tensor-core path
    "tc_causal_cln": dict(cfg=dict(N=64, L=16, B=64, H=128, P=3, X=2, R=1, C=2, norm_type="cLN",
                                   causal=True, mask_nonlinear="relu"), M=1, T=1300, lengths=[1300]),
    # odd channel counts (SIMT GEMM / scalar backward kernels), ragged lengths.  Gradient fixture for
    # that path: demo_gln's own gradients sit on a PReLU kink (one depthwise output is 5e-8, so fp32
    # rounding decides the branch) and cannot be held to a tight tolerance.
    # BatchNorm1d blocks (conv_tasnet.py:300-303) in INFERENCE mode on the tensor-core path: running statistics
    # come from three training-mode forwards of the reference, the affine is perturbed; no gradients stored
    "tc_bn_eval": dict(cfg=dict(N=64, L=16, B=64, H=128, P=3, X=3, R=1, C=2, norm_type="BN",
                                causal=False, mask_nonlinear="relu"), M=2, T=1207, lengths=[1207, 900], bn_eval=True),
    # BatchNorm1d in TRAINING mode: batch statistics, running-statistics update, gradients (the product raises
    # for this mode; the fixture pins the oracle for it)
    "bn_train": dict(cfg=dict(N=8, L=4, B=8, H=16, P=3, X=2, R=2, C=2, norm_type="BN",
                              causal=False, mask_nonlinear="relu"), M=3, T=61, lengths=[61, 61, 40], bn_train=True),
    "bn_eval_causal": dict(cfg=dict(N=8, L=4, B=8, H=16, P=3, X=2, R=2, C=2, norm_type="BN",
                                    causal=True, mask_nonlinear="relu"), M=2, T=51, lengths=[51, 37], bn_eval=True),
    "odd_gln": dict(cfg=dict(N=5, L=4, B=6, H=7, P=3, X=3, R=1, C=2, norm_type="gLN",
                             causal=False, mask_nonlinear="relu"), M=2, T=60, lengths=[60, 41]),
}


def main():
    ref_model, ref_loss, ref_utils = _import_reference()
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(1)  # bit-reproducible reductions

    only = [a for a in sys.argv[1:] if not a.startswith("-")]
    for name, case in CASES.items():
        if only and name not in only:
            continue
        cfg, M, T = case["cfg"], case["M"], case["T"]
        torch.manual_seed(123)
        model = ref_model.ConvTasNet(**cfg)
        g = torch.Generator().manual_seed(2025)
        if case.get("bn_eval"):
            with torch.no_grad():
                model.train()
                for _ in range(3):
                    model(0.1 * torch.randn(M, T, generator=g))
                for k, v in model.named_parameters():
                    if v.dim() == 1 and v.numel() > 1:      # BatchNorm1d weight / bias
                        v.add_(0.2 * torch.randn(v.shape, generator=g))
            model.eval()
        if case.get("bn_train"):
            with torch.no_grad():
                model.train()
                model(0.1 * torch.randn(M, T, generator=g))      # buffers away from their initial 0 / 1 / 0
                for k, v in model.named_parameters():
                    if v.dim() == 1 and v.numel() > 1:
                        v.add_(0.2 * torch.randn(v.shape, generator=g))
        init_sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
        sources = 0.1 * torch.randn(M, cfg["C"], T, generator=g)
        mixture = sources.sum(1) + 0.01 * torch.randn(M, T, generator=g)
        lengths = torch.tensor(case["lengths"], dtype=torch.int64)

        model.zero_grad()
        est = model(mixture)
        est_unmasked = est.detach().clone()
        loss, max_snr, est_masked, reorder = ref_loss.cal_loss(sources.clone(), est, lengths)
        if not case.get("bn_eval"):
            loss.backward()
        _, perms, idx = ref_loss.cal_si_snr_with_pit(sources.clone(), est_unmasked.clone(), lengths)

        blob = {"mixture": mixture, "sources": sources, "lengths": lengths,
                "est": est_unmasked, "loss": loss.detach(), "max_snr": max_snr.detach(),
                "est_masked": est_masked.detach(), "reorder": reorder.detach(),
                "perms": perms, "idx": idx}
        for k, v in init_sd.items():
            blob["w:" + k] = v
        if case.get("bn_train"):
            for k, v in model.state_dict().items():                # buffers after the training-mode forward
                if "running_" in k or "num_batches" in k:
                    blob["u:" + k] = v.detach().clone()
        for k, v in model.named_parameters():
            if v.grad is not None:
                blob["g:" + k] = v.grad.detach()
        np.savez_compressed(os.path.join(OUT, name + ".npz"),
                            cfg=np.array(repr(cfg)),
                            **{k: v.numpy() for k, v in blob.items()})
        print(name, "loss", float(loss), "idx", idx.tolist(), "K", (T - cfg["L"]) // (cfg["L"] // 2) + 1)

    if only:
        return
    # PIT-only cases: the reference demo shape (pit_criterion.py:118-126) with float data,
    # plus a C=3 case whose best permutation is a 3-cycle (reorder quirk)
    # and an exact tie (first permutation must win).
    g = torch.Generator().manual_seed(7)
    pit = {}
    B, C, T = 2, 3, 12
    src = torch.randn(B, C, T, generator=g)
    est = torch.randn(B, C, T, generator=g)
    lens = torch.tensor([T, T - 3])
    e = est.clone()
    loss, max_snr, est_m, reo = ref_loss.cal_loss(src, e, lens)
    _, perms, idx = ref_loss.cal_si_snr_with_pit(src, est.clone(), lens)
    pit.update(a_src=src, a_est=est, a_len=lens, a_loss=loss, a_max=max_snr, a_masked=est_m,
               a_reo=reo, a_idx=idx, a_perms=perms)
    B, C, T = 4, 3, 200
    src = torch.randn(B, C, T, generator=g)
    cyc = [[1, 2, 0], [2, 0, 1], [0, 2, 1], [0, 1, 2]]
    est = torch.stack([src[b, cyc[b]] for b in range(B)]) + 0.05 * torch.randn(B, C, T, generator=g)
    lens = torch.tensor([200, 150, 99, 200])
    e = est.clone()
    loss, max_snr, est_m, reo = ref_loss.cal_loss(src, e, lens)
    _, perms, idx = ref_loss.cal_si_snr_with_pit(src, est.clone(), lens)
    pit.update(b_src=src, b_est=est, b_len=lens, b_loss=loss, b_max=max_snr, b_masked=est_m,
               b_reo=reo, b_idx=idx)
    B, C, T = 1, 2, 64
    s0 = torch.randn(1, 1, T, generator=g)
    src = torch.cat([s0, s0], 1)           # identical targets -> both permutations tie exactly
    est = torch.randn(B, C, T, generator=g)
    est[:, 1] = est[:, 0]
    lens = torch.tensor([T])
    _, perms2, idx = ref_loss.cal_si_snr_with_pit(src, est.clone(), lens)
    pit.update(c_src=src, c_est=est, c_len=lens, c_idx=idx)
    np.savez_compressed(os.path.join(OUT, "pit_cases.npz"), **{k: v.numpy() for k, v in pit.items()})
    print("pit idx", pit["a_idx"].tolist(), pit["b_idx"].tolist(), pit["c_idx"].tolist())

    # overlap_and_add: the reference demo shape (utils.py:71-75) in float, plus L=20/16/odd cases
    ola = {}
    for tag, (M, C, K, L, step) in {"demo": (2, 2, 3, 4, 2), "l20": (1, 2, 17, 20, 10),
                                    "l16": (2, 3, 9, 16, 8), "odd": (1, 1, 6, 5, 2)}.items():
        sig = torch.randn(M, C, K, L, generator=g)
        ola[tag + "_in"] = sig
        ola[tag + "_out"] = ref_utils.overlap_and_add(sig, step)
        ola[tag + "_step"] = torch.tensor(step)
    np.savez_compressed(os.path.join(OUT, "ola_cases.npz"), **{k: v.numpy() for k, v in ola.items()})


    make_sisnri_cases()
    make_data_cases()


def _reference_functions(path, names):
    """Compiles the named top-level functions of a reference file WITHOUT importing the module
    (evaluate.py imports librosa / mir_eval, which are not installed): the definitions are taken
    from the file's AST at generation time; nothing is copied into this repository."""
    import ast
    with open(path) as f:
        tree = ast.parse(f.read())
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert sorted(n.name for n in body) == sorted(names)
    ns = {"np": np}
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    return [ns[n] for n in names]


def make_sisnri_cases():
    """SI-SNRi fixtures from the reference's own evaluate.cal_SISNRi / cal_SISNR (evaluate.py:108-144)."""
    cal_SISNRi, cal_SISNR = _reference_functions(os.path.join(REF, "evaluate.py"), ["cal_SISNRi", "cal_SISNR"])
    rng = np.random.RandomState(11)
    out = {}
    # (T, noise level of the estimate): low, medium and very high SNR, short and long, a DC offset
    for i, (T, noise, dc) in enumerate([(12, 0.5, 0.0), (4000, 0.1, 0.0), (32000, 1e-3, 0.3), (16001, 1.0, -0.2),
                                        (8000, 1e-5, 0.0)]):
        ref = (0.1 * rng.randn(2, T) + dc).astype(np.float32)
        est = (ref * rng.uniform(0.5, 2.0) + noise * 0.1 * rng.randn(2, T)).astype(np.float32)
        mix = (ref.sum(0) + 0.01 * rng.randn(T)).astype(np.float32)
        out["ref%d" % i], out["est%d" % i], out["mix%d" % i] = ref, est, mix
        out["sisnri%d" % i] = np.float64(cal_SISNRi(ref, est, mix))
        out["sisnr%d" % i] = np.array([[cal_SISNR(ref[c], est[c]), cal_SISNR(ref[c], mix)] for c in range(2)],
                                      dtype=np.float64)
        print("sisnri case", i, out["sisnri%d" % i], out["sisnr%d" % i].tolist())
    np.savez_compressed(os.path.join(OUT, "sisnri_cases.npz"), **out)


def make_data_cases():
    """Batching / segmenting / padding fixtures from the reference's own data.py (AudioDataset,
    EvalDataset and both collate functions, data.py:32-232), imported unmodified.  Its two absent
    third-party imports are stubbed in the harness: `librosa.load` serves the synthetic waveforms of
    oracle/synth_audio.py (keyed by the fake path), `preprocess.preprocess_one_dir` is never reached
    (mix_json is given)."""
    import json
    import tempfile
    import types
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import synth_audio as sa
    sys.dont_write_bytecode = True
    librosa = types.ModuleType("librosa")
    librosa.load = lambda path, sr=None: (sa.synth(path), sr)
    preprocess = types.ModuleType("preprocess")
    preprocess.preprocess_one_dir = lambda *a, **k: (_ for _ in ()).throw(AssertionError("not expected"))
    sys.modules["librosa"], sys.modules["preprocess"] = librosa, preprocess
    sys.path.insert(0, REF)
    import data as ref_data
    sys.path.pop(0)
    del sys.modules["librosa"], sys.modules["preprocess"]

    plan, arrays = {}, {}
    for name, case in sa.DATA_CASES.items():
        infos = sa.make_infos(case["lengths"], case["shuffle"])
        with tempfile.TemporaryDirectory() as d:
            for kind, lst in zip(("mix", "s1", "s2"), infos):
                with open(os.path.join(d, kind + ".json"), "w") as f:
                    json.dump(lst, f)
            ds = ref_data.AudioDataset(d, **case["kw"])
        plan[name] = [[[i[0] for i in part] for part in mb[:3]] + [mb[3], mb[4]] for mb in ds.minibatch]
        for b, mb in enumerate(ds.minibatch):
            mix_pad, ilens, src_pad = ref_data._collate_fn([mb])
            arrays["%s.%d.mix" % (name, b)] = mix_pad.numpy()
            arrays["%s.%d.ilens" % (name, b)] = ilens.numpy()
            arrays["%s.%d.src" % (name, b)] = src_pad.numpy()
        print(name, [[os.path.basename(p) for p in mb[0]] for mb in plan[name]])
    for name, case in sa.EVAL_CASES.items():
        infos = sa.make_infos(case["lengths"], case["shuffle"])[0]
        with tempfile.TemporaryDirectory() as d:
            with open(os.path.join(d, "mix.json"), "w") as f:
                json.dump(infos, f)
            ds = ref_data.EvalDataset(None, os.path.join(d, "mix.json"), **case["kw"])
        plan[name] = [[[i[0] for i in mb[0]], mb[1]] for mb in ds.minibatch]
        for b, mb in enumerate(ds.minibatch):
            mix_pad, ilens, names = ref_data._collate_fn_eval([mb])
            assert names == plan[name][b][0]
            arrays["%s.%d.mix" % (name, b)] = mix_pad.numpy()
            arrays["%s.%d.ilens" % (name, b)] = ilens.numpy()
        print(name, [[os.path.basename(p) for p in mb[0]] for mb in plan[name]])
    # pad_list with a non-zero pad value and 2-D items (data.py:295-301)
    xs = [torch.arange(6.).view(3, 2), torch.arange(10.).view(5, 2), torch.arange(2.).view(1, 2)]
    arrays["pad_list.out"] = ref_data.pad_list(xs, -1.5).numpy()
    with open(os.path.join(OUT, "data_cases.json"), "w") as f:
        json.dump(plan, f, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(OUT, "data_cases.npz"), **arrays)


if __name__ == "__main__":
    if sys.argv[1:] == ["sisnri"]:
        make_sisnri_cases()
    elif sys.argv[1:] == ["data"]:
        make_data_cases()
    else:
        main()
