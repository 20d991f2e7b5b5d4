"""CPU-side checks (no GPU): C-ABI library loads and exports every symbol
include/ctn_b200.h declares, the drop-in's Python surface matches the reference
(state_dict keys / shapes, same-seed initial weights, checkpoint package), and
the product path refuses to run without CUDA instead of falling back."""
import io
import os
import re

import pytest
import torch

from conftest import MODEL_CASES, PKG, ROOT, load_golden


def test_library_exports_every_declared_symbol():
    from ctn_b200 import _native as nat
    header = open(os.path.join(ROOT, "include", "ctn_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(ctn_[a-z0-9_]+)\s*\(", header)))
    lib = nat.lib()
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == nat.exported_symbols()
    assert lib.ctn_version() >= 100
    assert lib.ctn_operand_format() == 3      # the product build: bf16 (hi, lo) operands, three MMAs per product
    # pure-host geometry queries (no device work)
    assert lib.ctn_gemm_gln_partials(3199, 512) == 25 * 4 * 4
    assert lib.ctn_dwconv_gln_partials(3199, 512, 3, 1) == 50      # marching kernel: 4*1*100 warps / 8
    assert lib.ctn_dwconv_gln_partials(3199, 512, 3, 128) == 64
    assert lib.ctn_dwconv_gln_partials(3199, 510, 3, 1) == 200     # generic kernel: 16-frame tiles
    assert lib.ctn_gemm_cln_groups(512) == 4
    assert lib.ctn_pit_workspace_doubles(3, 2, 32000) == 3 * 8 * 14


def test_invalid_arguments_return_error_codes_without_a_gpu():
    import ctypes as C
    from ctn_b200 import _native as nat
    lib = nat.lib()
    assert lib.ctn_gemm1x1(None, None) == -1
    assert b"NULL" in lib.ctn_last_error()
    g = nat.GemmArgs()
    assert lib.ctn_gemm1x1(C.byref(g), None) == -1          # zero shape
    with pytest.raises(RuntimeError):
        nat.check(-1)
    assert lib.ctn_encoder_fwd(1, 1, 1, 3, 8, 4, None, None, None, None, None) == -1   # T < L
    assert lib.ctn_decoder_fwd(1, 1, 1, 1, 5, 8, 4, 2, 100, 7, 1, None) == -1          # bad mask id
    with pytest.raises(ValueError):                                                     # conv_tasnet.py:208
        nat.check(-1)
    # batch assembly: mixture / source pointers come in pairs, something must be asked for
    assert lib.ctn_pad_batch(None, None, None, 1, 2, 8, 0.0, None, None, None) == -1
    assert b"offsets" in lib.ctn_last_error()
    assert lib.ctn_pad_batch(8, None, 8, 1, 2, 8, 0.0, None, None, None) == -1
    assert b"together" in lib.ctn_last_error()
    assert lib.ctn_pad_batch(None, None, 8, 1, 2, 8, 0.0, None, None, None) == -1
    assert lib.ctn_remove_pad(None, None, 1, 2, 8, None, None) == -1
    assert lib.ctn_remove_pad(8, 8, 70000, 2, 8, 8, None) == -1                         # grid limit
    assert lib.ctn_pad_batch(8, None, 8, 0, 0, 8, 0.0, 8, None, None) == 0              # empty batch: no launch


@pytest.mark.parametrize("name", MODEL_CASES)
def test_same_seed_init_and_state_dict_layout_match_reference(name):
    import conv_tasnet
    cfg, z = load_golden(name)
    torch.manual_seed(123)                       # oracle/make_golden.py seeds the reference the same way
    model = conv_tasnet.ConvTasNet(**cfg)
    ref = {k[2:]: v for k, v in z.items() if k.startswith("w:")}
    sd = model.state_dict()
    assert list(sd) == list(ref)
    for k in sd:
        assert sd[k].shape == ref[k].shape, k
        assert torch.equal(sd[k], ref[k]), k
    assert [k for k, _ in model.named_parameters()] == list(ref)
    for attr in ("N", "L", "B", "H", "P", "X", "R", "C", "norm_type", "causal", "mask_nonlinear"):
        assert getattr(model, attr) == cfg[attr]


def test_paper_config_parameter_count():
    import conv_tasnet
    m = conv_tasnet.ConvTasNet(256, 20, 256, 512, 3, 8, 4, 2)
    assert sum(p.numel() for p in m.parameters()) == 8710720
    assert len(m.state_dict()) == 294
    m5 = conv_tasnet.ConvTasNet(512, 16, 128, 512, 3, 8, 3, 3, mask_nonlinear="softmax")
    assert sum(p.numel() for p in m5.parameters()) == 3511344


def test_checkpoint_package_round_trip(tmp_path):
    import conv_tasnet
    torch.manual_seed(1)
    m = conv_tasnet.ConvTasNet(8, 4, 8, 16, 3, 2, 2, 2, norm_type="cLN", causal=True)
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    pkg = conv_tasnet.ConvTasNet.serialize(m, opt, epoch=3, tr_loss=torch.zeros(5), cv_loss=torch.ones(5))
    assert set(pkg) == {"N", "L", "B", "H", "P", "X", "R", "C", "norm_type", "causal", "mask_nonlinear",
                        "state_dict", "optim_dict", "epoch", "tr_loss", "cv_loss"}
    path = tmp_path / "final.pth.tar"
    torch.save(pkg, path)
    m2 = conv_tasnet.ConvTasNet.load_model(str(path))
    assert m2.causal is True and m2.norm_type == "cLN"
    for (k1, v1), (k2, v2) in zip(m.state_dict().items(), m2.state_dict().items()):
        assert k1 == k2 and torch.equal(v1, v2)
    # causal models carry the Chomp1d index shift in their key names
    assert "separator.network.2.0.0.net.3.net.4.weight" in m.state_dict()
    buf = io.BytesIO()
    torch.save(m, buf)                              # the module itself pickles (engine state is dropped)


def test_no_cpu_fallback():
    import conv_tasnet
    import pit_criterion
    import utils
    m = conv_tasnet.ConvTasNet(8, 4, 8, 8, 3, 1, 1, 2).eval()
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 100))
    with pytest.raises(RuntimeError, match="CUDA"):
        pit_criterion.cal_loss(torch.zeros(1, 2, 10), torch.zeros(1, 2, 10), torch.tensor([10]))
    with pytest.raises(RuntimeError, match="CUDA"):
        utils.overlap_and_add(torch.zeros(1, 3, 4), 2)
    with pytest.raises(ValueError):
        conv_tasnet.ConvTasNet(8, 4, 8, 8, 3, 1, 1, 2, mask_nonlinear="sigmoid")


def test_remove_pad_and_get_mask():
    import pit_criterion
    import utils
    x = torch.arange(24.).view(2, 3, 4)
    out = utils.remove_pad(x, torch.tensor([4, 2]))
    assert out[0].shape == (3, 4) and out[1].shape == (3, 2)
    assert (out[1] == x[1, :, :2].numpy()).all()
    y = utils.remove_pad(x[:, 0], torch.tensor([3, 1]))
    assert y[0].shape == (3,) and y[1].shape == (1,)
    mask = pit_criterion.get_mask(torch.zeros(2, 3, 5), torch.tensor([5, 2]))
    assert mask.shape == (2, 1, 5) and mask[1, 0].tolist() == [1, 1, 0, 0, 0]


def test_product_path_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "conv-tasnet_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower() or f == "never", (dirpath, f)


def test_precision_variants_of_the_library():
    """CTN_PRECISION picks the library per process; each library states its operand format and exports the
    whole ABI; a mismatch or an unknown precision is refused."""
    import ctypes as C
    import subprocess
    import sys
    from ctn_b200 import _native as nat
    assert nat.PRECISION == "bf16x3"
    fp16 = os.path.join(PKG, "libctn_b200_fp16.so")
    assert os.path.exists(fp16), "python conv-tasnet_b200/build_native.py --fp16 (done by __graft_entry__.build())"
    h = C.CDLL(fp16)
    assert h.ctn_operand_format() == 101
    for name in nat.exported_symbols():
        assert hasattr(h, name), name
    code = "from ctn_b200 import _native as n; n.lib(); print(n.PRECISION, n.lib().ctn_operand_format())"
    env = dict(os.environ, PYTHONPATH=PKG, CTN_PRECISION="fp16")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.split() == ["fp16", "101"], r.stderr[-500:]
    r = subprocess.run([sys.executable, "-c", code], env=dict(env, CTN_PRECISION="fp8"), capture_output=True, text=True)
    assert r.returncode != 0 and "CTN_PRECISION" in r.stderr
