"""ctn_pit_sisnr_fwd_static: the allocation-free variant of ctn_pit_sisnr_fwd (scratch in a static __device__
array inside the library; include/ctn_b200.h).  It has no counterpart in the reference
-- so its oracle is the existing entry point: same inputs through both, through the
C ABI, on two calls in a row (the second call finds the first call's partials in the shared array).
This file sorts after every other test file on purpose."""
import ctypes as C

import pytest
import torch

from conftest import PKG  # noqa: F401  (puts the package on sys.path)

pytestmark = pytest.mark.gpu
DEV = "cuda"
DB_TOL = 1e-4     # dB; the two entry points run the same kernels, so the results are in fact bit-equal


def _perms(Cn):
    from itertools import permutations
    return torch.tensor(list(permutations(range(Cn))), dtype=torch.int64, device=DEV)


def _run(fn_name, src, est, lengths, with_workspace, strict=False):
    from ctn_b200 import _native as nat
    lib = nat.lib()
    B, Cn, T = src.shape
    perms = _perms(Cn)
    est = est.clone()
    out = dict(max_snr=torch.full((B,), float("nan"), device=DEV),
               max_idx=torch.full((B,), -1, dtype=torch.int64, device=DEV),
               loss=torch.full((1,), float("nan"), device=DEV),
               pair=torch.full((B, Cn, Cn), float("nan"), device=DEV),
               coef=torch.full((B, Cn, 4), float("nan"), device=DEV))
    args = [src.data_ptr(), est.data_ptr(), lengths.data_ptr(), B, Cn, T, perms.data_ptr(), perms.shape[0]]
    if with_workspace:
        ws = torch.empty(lib.ctn_pit_workspace_doubles(B, Cn, T), dtype=torch.float64, device=DEV)
        args.append(ws.data_ptr())
    args += [out["max_snr"].data_ptr(), out["max_idx"].data_ptr(), out["loss"].data_ptr(), out["pair"].data_ptr(),
             out["coef"].data_ptr(), nat.stream()]
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    stats0 = torch.cuda.memory_stats().get("num_alloc_retries", 0), torch.cuda.memory_allocated()
    rc = getattr(lib, fn_name)(*args)
    torch.cuda.synchronize()
    if strict:      # nothing was allocated on the device by the call, by torch or behind its back (checked on the
        # second call: the first use of a library symbol may make the driver load its module lazily)
        assert torch.cuda.mem_get_info()[0] == free0
        assert (torch.cuda.memory_stats().get("num_alloc_retries", 0), torch.cuda.memory_allocated()) == stats0
    out["rc"], out["est"] = rc, est
    return out


CASES = [
    # B, C, T, lengths                        (workspace doubles: B * ceil(T / 4096) * (5 C + C^2))
    (3, 2, 32000, [32000, 24000, 16001]),     # BASELINE configs[1]'s batch
    (8, 3, 32000, None),                      # BASELINE configs[4]'s batch: 8 x 8 x 24 = 1536 doubles
    (2, 1, 500, [500, 3]),
    (64, 2, 32000, None),                     # a bench micro-batch: 7168 doubles
]


@pytest.mark.parametrize("B,Cn,T,lens", CASES)
def test_static_scratch_entry_point_matches_the_workspace_one_twice(B, Cn, T, lens):
    from ctn_b200 import _native as nat
    g = torch.Generator().manual_seed(100 + B)
    for call in range(2):                    # two calls in a row, different data: the array keeps the last contents
        src = torch.randn(B, Cn, T, generator=g).to(DEV)
        est = (src[:, torch.randperm(Cn, generator=g)] + 0.3 * torch.randn(B, Cn, T, generator=g).to(DEV))
        lengths = torch.tensor(lens if lens else [T] * B, dtype=torch.int64, device=DEV)
        want = _run("ctn_pit_sisnr_fwd", src, est, lengths, True)
        got = _run("ctn_pit_sisnr_fwd_static", src, est, lengths, False, strict=(call == 1))
        assert want["rc"] == 0
        nat.check(got["rc"])
        assert torch.equal(got["max_idx"], want["max_idx"]), call            # permutation choice: bit-exact
        assert torch.allclose(got["max_snr"], want["max_snr"], atol=DB_TOL, rtol=0), call
        assert abs(float(got["loss"]) - float(want["loss"])) <= DB_TOL, call
        assert torch.allclose(got["pair"], want["pair"], atol=DB_TOL, rtol=0), call
        assert torch.allclose(got["coef"], want["coef"], rtol=1e-6, atol=1e-12), call
        assert torch.equal(got["est"], want["est"]), call                    # the in-place masking side effect
        assert not torch.isnan(got["max_snr"]).any()


def test_static_scratch_entry_point_rejects_sizes_beyond_its_array():
    from ctn_b200 import _native as nat
    lib = nat.lib()
    B, Cn, T = 700, 2, 32000                 # 700 x 8 x 14 = 78 400 doubles > 65 536
    assert lib.ctn_pit_workspace_doubles(B, Cn, T) > 65536
    one = torch.zeros(1, device=DEV)         # never dereferenced: the size check comes first
    p, i64 = one.data_ptr(), torch.zeros(B, dtype=torch.int64, device=DEV)
    rc = lib.ctn_pit_sisnr_fwd_static(p, p, i64.data_ptr(), B, Cn, T, _perms(Cn).data_ptr(), 2, p, i64.data_ptr(), p,
                                      None, None, nat.stream())
    assert rc == -2                          # CTN_ERR_UNSUPPORTED
    msg = lib.ctn_last_error().decode()
    assert "65536" in msg and "78400" in msg, msg
    with pytest.raises(RuntimeError):
        nat.check(rc)
    torch.cuda.synchronize()
    # the limit itself is accepted: 65 536 doubles = 585 x 8 x 14 (B = 585, C = 2, T = 32000) -> 65 520
    assert lib.ctn_pit_workspace_doubles(585, 2, 32000) <= 65536 < lib.ctn_pit_workspace_doubles(586, 2, 32000)


def test_static_scratch_symbol_is_declared_and_exported():
    import os
    from ctn_b200 import _native as nat
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(os.path.join(root, "include", "ctn_b200.h")) as f:
        header = f.read()
    assert "ctn_pit_sisnr_fwd_static" in header and "CTN_PIT_STATIC_SCRATCH_DOUBLES 65536" in header
    assert "ctn_pit_sisnr_fwd_static" in nat.exported_symbols()
    assert isinstance(nat.lib().ctn_pit_sisnr_fwd_static, C._CFuncPtr)
