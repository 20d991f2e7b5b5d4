"""Host-side logic of the utterance-sharded data parallelism (ctn_b200/_dist.py), world size 2
over gloo on CPU.  The per-rank gradients come from the oracle (the checker); what is under
test is the sharding and the ONE weighted all-reduce that must reproduce the reference's
single-process global-mean-loss gradient (pit_criterion.py:22, solver.py:178-186)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import PKG  # noqa: F401  (puts the package on sys.path)
from oracle import convtasnet_oracle as O

CFG = dict(N=8, L=4, B=8, H=8, P=3, X=2, R=1, C=2, norm_type="gLN", causal=False, mask_nonlinear="relu")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _flat_grad(params, mix, src, lengths):
    p = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    est = O.forward(mix, p, CFG)
    loss, *_ = O.cal_loss(src, est, lengths)
    loss.backward()
    return torch.cat([v.grad.flatten() for v in p.values()]), loss.detach()


def _worker(rank, world, port, M, out):
    from ctn_b200 import _dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        params = O.random_params(*(CFG[k] for k in "NLBHPXRC"), False, seed=3, dtype=torch.float64)
        g = torch.Generator().manual_seed(11)
        src = 0.1 * torch.randn(M, 2, 60, generator=g, dtype=torch.float64)
        mix = src.sum(1)
        lengths = torch.tensor([60, 41, 60, 33, 50][:M])
        full, full_loss = _flat_grad(params, mix, src, lengths)
        lo, hi = _dist.shard_bounds(M, world, rank)
        assert torch.equal(_dist.shard(mix), mix[lo:hi])
        local, local_loss = _flat_grad(params, mix[lo:hi], src[lo:hi], lengths[lo:hi])
        n_global = None if M % world == 0 else M
        red = _dist.reduce_flat_grad(local.clone(), hi - lo, n_global)
        gl = _dist.global_loss(local_loss, hi - lo, n_global)
        err = float((red - full).norm() / full.norm())
        out[rank] = (err, abs(float(gl) - float(full_loss)), lo, hi)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("M", [4, 5])
def test_sharded_gradient_equals_global_mean_gradient(M):
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), M, out), nprocs=world, join=True)
    assert len(out) == world
    covered = []
    for r in range(world):
        err, dl, lo, hi = out[r]
        assert err < 1e-12, (r, err)
        assert dl < 1e-12
        covered += list(range(lo, hi))
    assert covered == list(range(M))


def test_shard_bounds_cover_and_balance():
    from ctn_b200 import _dist
    for n in (0, 1, 3, 8, 512, 513):
        for w in (1, 2, 3, 8):
            b = [_dist.shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        _dist.shard_bounds(4, 2, 2)


def test_shard_minibatch_partitions_dataset_items():
    """Every utterance of a minibatch goes to exactly one rank, in order, with the non-list fields intact."""
    from ctn_b200 import _dist
    mix = [["mix/u%d.wav" % i, 100 - i] for i in range(5)]
    s1 = [["s1/u%d.wav" % i, 100 - i] for i in range(5)]
    s2 = [["s2/u%d.wav" % i, 100 - i] for i in range(5)]
    item = [mix, s1, s2, 8000, 32000]
    for world in (1, 2, 3, 8):
        parts = [_dist.shard_minibatch(item, world, r) for r in range(world)]
        assert sum((p[0] for p in parts), []) == mix and sum((p[2] for p in parts), []) == s2
        assert all(p[3:] == [8000, 32000] for p in parts)
        assert all(len(p[0]) == len(p[1]) == len(p[2]) for p in parts)
        sizes = [len(p[0]) for p in parts]
        assert max(sizes) - min(sizes) <= 1
    ev = _dist.shard_minibatch([mix, 8000], 2, 1)
    assert ev == [mix[3:], 8000]


def _empty_worker(rank, world, port, out):
    from ctn_b200 import _dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        model = torch.nn.Linear(5, 3)          # stands for the replica: only its parameter list matters here
        n = sum(p.numel() for p in model.parameters())
        lo, hi = _dist.shard_bounds(1, world, rank)      # one utterance, two ranks: rank 1 is empty
        if hi > lo:
            flat = torch.arange(n, dtype=torch.float32) + 1.0
            _dist.reduce_flat_grad(flat, hi - lo, 1)
        else:
            flat = _dist.empty_shard_step(model)
            assert all(p.grad is not None for p in model.parameters())
            assert torch.equal(torch.cat([p.grad.flatten() for p in model.parameters()]), flat)
        out[rank] = flat.tolist()
    finally:
        dist.destroy_process_group()


def test_rank_with_an_empty_shard_joins_the_all_reduce():
    """Fewer utterances than ranks: the empty rank contributes zeros to the one collective of the step and ends
    with the global gradient, instead of leaving the other ranks waiting."""
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_empty_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert out[0] == out[1] == [float(i + 1) for i in range(len(out[0]))]
