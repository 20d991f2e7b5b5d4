"""Data parallelism over real GPUs (NCCL): two ranks, each with its utterance shard, must end up
with the gradient of the reference's single-process global-mean loss (train.py:121 DataParallel
semantics).  Needs >= 2 GPUs; skipped otherwise (the same host logic is covered on
CPU over gloo in tests/test_dist_cpu.py)."""
import os
import socket

import pytest
import torch

from conftest import PKG, ROOT  # noqa: F401

pytestmark = pytest.mark.gpu

CFG = dict(N=64, L=16, B=64, H=128, P=3, X=3, R=1, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, M, out, norm="gLN"):
    # BatchNorm: the kink-free variant (PReLU slopes 1, softmax mask).  Batch normalisation centres every channel on
    # the PReLU kink, so a 1e-7 difference in summation order between 1 and 2 shards can flip the branch of an
    # element and move the gradient by 1e-4 (tests/test_gpu_bn.py); without branch points the comparison is tight.
    smooth = norm == "BN"
    cfg = dict(CFG, norm_type=norm, mask_nonlinear="softmax" if smooth else "relu")
    import sys
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    import conv_tasnet
    import pit_criterion
    from ctn_b200 import _dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        torch.manual_seed(5)                       # same weights on every rank ...
        ref_model = conv_tasnet.ConvTasNet(**cfg).to(dev).train()
        torch.manual_seed(6 + rank)                # ... but rank-dependent ones here: attach() must broadcast
        model = conv_tasnet.ConvTasNet(**cfg).to(dev).train()
        with torch.no_grad():
            if smooth:
                for p in ref_model.parameters():
                    if p.numel() == 1:
                        p.fill_(1.0)
            if rank == 0:
                model.load_state_dict(ref_model.state_dict())
        n_global = None if M % world == 0 else M
        _dist.attach(model, global_batch=n_global)
        for a, b in zip(model.parameters(), ref_model.parameters()):
            assert torch.equal(a, b)
        g = torch.Generator().manual_seed(11)
        src = (0.1 * torch.randn(M, 2, 2400, generator=g)).to(dev)
        mix = src.sum(1)
        lengths = torch.tensor([2400, 2000, 2400, 1700, 2222][:M], device=dev)
        # the reference semantics: one process, whole batch
        est = ref_model(mix)
        loss_ref, *_ = pit_criterion.cal_loss(src, est, lengths)
        loss_ref.backward()
        full = torch.cat([p.grad.flatten() for p in ref_model.parameters()])
        # this rank's shard
        lo, hi = _dist.shard_bounds(M, world, rank)
        est = model(mix[lo:hi])
        loss, *_ = pit_criterion.cal_loss(src[lo:hi], est, lengths[lo:hi])
        loss.backward()
        mine = torch.cat([p.grad.flatten() for p in model.parameters()])
        gl = _dist.global_loss(loss, hi - lo, n_global)
        err = float((mine.double() - full.double()).norm() / full.double().norm())
        # BatchNorm1d: the batch statistics were summed over the ranks, so the updated running buffers equal
        # the single-process ones as well
        berr = 0.0
        for (ka, a), (kb, b) in zip(model.named_buffers(), ref_model.named_buffers()):
            assert ka == kb
            berr = max(berr, float((a.double() - b.double()).abs().max()))
        out[rank] = (err, abs(float(gl) - float(loss_ref)), berr)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("M", [4, 5])
def test_two_rank_gradient_equals_single_process(M):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), M, out), nprocs=2, join=True)
    for r in range(2):
        err, dl, _ = out[r]
        assert err < 2e-5, (r, err)      # fp32 summation order differs between 1 and 2 shards
        assert dl < 1e-4


@pytest.mark.parametrize("M", [4, 5])
def test_two_rank_batchnorm_training_equals_single_process(M):
    """norm_type="BN" (conv_tasnet.py:300-303): batch statistics couple the utterances, so the [2, H] column sums
    of every norm are all-reduced across the ranks (forward and backward); two ranks with (un)equal shards
    must reproduce the single-process gradient AND running statistics."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), M, out, "BN"), nprocs=2, join=True)
    for r in range(2):
        err, dl, berr = out[r]
        assert err < 2e-5, (r, err)
        assert dl < 1e-4
        assert berr < 1e-6, (r, berr)


def test_reference_dataparallel_wrapper():
    """train.py:120-122 / separate.py:43-46 wrap the model in torch.nn.DataParallel: one process, the
    batch scattered over the GPUs, replicas called from one Python thread per GPU.  The drop-in must
    give the single-GPU result (inference) and the single-GPU gradients (training) through it."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import conv_tasnet
    import pit_criterion
    torch.manual_seed(5)
    model = conv_tasnet.ConvTasNet(**CFG).cuda(0)
    g = torch.Generator().manual_seed(12)
    src = (0.1 * torch.randn(4, 2, 2400, generator=g)).cuda(0)
    mix = src.sum(1)
    lengths = torch.tensor([2400, 2000, 2400, 1700], device="cuda:0")
    dp = torch.nn.DataParallel(model, device_ids=[0, 1])
    model.eval()
    with torch.no_grad():
        single = model(mix)
        multi = dp(mix)
    assert multi.device == single.device and torch.equal(single, multi)   # utterances are independent
    model.train()
    model.zero_grad()
    loss, *_ = pit_criterion.cal_loss(src, model(mix), lengths)
    loss.backward()
    want = torch.cat([p.grad.flatten() for p in model.parameters()]).clone()
    model.zero_grad()
    loss2, *_ = pit_criterion.cal_loss(src, dp(mix), lengths)
    loss2.backward()
    got = torch.cat([p.grad.flatten() for p in model.parameters()])
    assert abs(float(loss) - float(loss2)) < 1e-4
    err = float((got.double() - want.double()).norm() / want.double().norm())
    assert err < 2e-5, err
