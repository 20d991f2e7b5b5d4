"""Data parallelism over the GPUs of one box: one process per GPU, full weight replica,
utterances of the batch sharded across ranks.

The reference's only multi-GPU mechanism is torch.nn.DataParallel (train.py:121): one
process, per-step parameter broadcast, output gather and gradient reduce-to-root.  Here
each rank runs the whole hot path on its own shard; the ONLY collective on the path is one
all-reduce (sum) of the flat fp32 gradient buffer per step (NCCL over NVLink/NVSwitch on
the GPU box, gloo in the CPU tests), issued inside the backward of the autograd Function
(ctn_b200/_autograd.py).  Inference has no collective at all.

Equivalence with the reference's single-process loss  L = -mean_b max_snr_b  over the global
batch of M utterances (pit_criterion.py:22): rank g holds n_g utterances and a local mean
loss L_g, so  L = sum_g (n_g / M) L_g  and  dL/dtheta = sum_g (n_g / M) dL_g/dtheta.  Each
rank scales its flat gradient by n_g / M before the sum; for equal shards that is 1 / G.
"""
import torch
import torch.distributed as dist


def shard_bounds(n_items, world, rank):
    """[lo, hi) of the contiguous shard of `n_items` utterances owned by `rank`: the atom is
    one utterance; the first n_items % world ranks get one extra."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad world/rank %r/%r" % (world, rank))
    base, extra = divmod(int(n_items), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard(tensor, group=None):
    """This rank's utterance shard of a batch-major tensor (dim 0 = utterance)."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = shard_bounds(tensor.shape[0], world, rank)
    return tensor[lo:hi]


def shard_minibatch(minibatch, world, rank):
    """This rank's part of one dataset item (data.py: AudioDataset / EvalDataset items are whole minibatches,
    [mix_infos, s1_infos, s2_infos, sample_rate, segment_len] or [mix_infos, sample_rate]): the utterance lists
    are cut with shard_bounds, everything else is passed through, so each rank loads and pads only its own
    utterances.  Shards differ in size when the minibatch does not divide (and, with segmenting, in their number
    of segments): pass the global segment count to attach(global_batch=...) / reduce_flat_grad(n_global=...)."""
    n = len(minibatch[0])
    lo, hi = shard_bounds(n, world, rank)
    return [part[lo:hi] if isinstance(part, (list, tuple)) and len(part) == n else part for part in minibatch]


def reduce_flat_grad(flat, n_local, n_global=None, group=None):
    """In place: flat <- sum over ranks of (n_local / n_global) * flat.  ONE all-reduce.
    n_global None = equal shards (scale 1 / world)."""
    world = dist.get_world_size(group)
    scale = (1.0 / world) if n_global is None else (float(n_local) / float(n_global))
    flat.mul_(scale)
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def empty_shard_step(model, group=None):
    """The step of a rank whose shard of the minibatch is EMPTY (fewer utterances than ranks: shard_bounds gives
    lo == hi).  Such a rank has no forward and no backward, so it would never reach the all-reduce the other
    ranks wait in; it calls this instead: it contributes a zero flat gradient to the same collective and
    receives the global gradient as .grad of its parameters, so its optimizer step keeps the replica in sync."""
    ps = [p for p in model.parameters() if p.requires_grad]
    flat = torch.zeros(sum(p.numel() for p in ps), dtype=torch.float32, device=ps[0].device)
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    off = 0
    for p in ps:
        p.grad = flat[off:off + p.numel()].view_as(p)
        off += p.numel()
    return flat


def global_loss(local_loss, n_local, n_global=None, group=None):
    """The reference's batch-mean loss from the ranks' local means (reporting only; one
    scalar all-reduce, not on the gradient path)."""
    t = local_loss.detach().clone().reshape(1)
    return reduce_flat_grad(t, n_local, n_global, group)[0]


def broadcast_parameters(model, src=0, group=None):
    """Replicates rank `src`'s weights on every rank (what DataParallel's per-step replicate
    does, done once)."""
    with torch.no_grad():
        for p in model.parameters():
            dist.broadcast(p.data, src=src, group=group)
        for b in model.buffers():
            dist.broadcast(b.data, src=src, group=group)
        # `.data` writes do not advance the version counters the engines key their prepared weights on
        ts = list(model.parameters()) + list(model.buffers())
        if ts:
            torch.autograd.graph.increment_version(ts)


def attach(model, group=None, global_batch=None, broadcast=True):
    """Turns `model` (the ConvTasNet drop-in) into a data-parallel replica: weights are
    broadcast from rank 0 and every backward all-reduces the flat gradient buffer.
    global_batch: utterances per step over ALL ranks when shards are unequal; None = equal."""
    if not dist.is_initialized():
        raise RuntimeError("torch.distributed is not initialised (init_process_group('nccl') first)")
    if broadcast:
        broadcast_parameters(model, 0, group)
    eng = model.train_engine()
    eng.process_group = group if group is not None else dist.group.WORLD
    eng.global_batch = global_batch
    return model


def detach(model):
    eng = model.train_engine()
    eng.process_group = None
    eng.global_batch = None
    return model
