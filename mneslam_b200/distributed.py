"""Multi-GPU plumbing of the ray-sharded mapping iteration.

Every rank holds the full map and renders its own shard of the ray batch.  The one real exchange per iteration
is the gradient reduction; it is done ZeRO-1 style over the flat buffers of ``FusedMapper``:

    reduce-scatter(flat gradient)  ->  Adam on this rank's 1/G slice of the parameters  ->  all-gather(parameters)

which moves the same bytes over NVLink as one all-reduce but divides the optimiser's HBM traffic (32 B per
parameter, the largest single cost of an iteration on one GPU) by the number of ranks.  The loss accumulators
(16 doubles) are all-reduced separately, before the losses are finalised, because the masked means are ratios of
global sums.

The helpers below are backend-agnostic (NCCL on GPUs; gloo in the CPU tests, where the Adam kernel is replaced
by a torch stand-in).
"""
import torch
import torch.distributed as dist


def shard_rays(n_rays, rank, world_size):
    """Contiguous, balanced split of a global ray batch: returns (begin, end) of this rank's rays."""
    base, rem = divmod(n_rays, world_size)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


class FlatShards:
    """Slice bookkeeping of a flat parameter buffer split evenly over ranks.

    ``seg_ends`` / ``hparams``: exclusive end offsets (elements, multiples of ``align``) and per-segment
    hyper-parameters of the optimiser's parameter groups over the unpadded buffer."""

    def __init__(self, total, world_size, rank, seg_ends, hparams, align=4):
        q = world_size * align
        self.total = total
        self.padded = (total + q - 1) // q * q
        self.shard_len = self.padded // world_size
        self.lo = rank * self.shard_len
        self.hi = self.lo + self.shard_len
        self.world_size, self.rank = world_size, rank
        ends, hp = [], []
        begin = 0
        for i, (end, h) in enumerate(zip(seg_ends, hparams)):
            last = i == len(seg_ends) - 1
            e = self.padded if last else end          # the zero padding rides along with the last segment
            a, b = max(begin, self.lo), min(e, self.hi)
            if b > a:
                ends.append(b - self.lo)
                hp.append(h)
            begin = end
        self.local_ends, self.local_hparams = ends, hp

    def shard_of(self, flat):
        return flat[self.lo:self.hi]


def reduce_scatter_sum(out_shard, flat, shards, group=None):
    """out_shard <- this rank's slice of the element-wise sum of ``flat`` over ranks."""
    backend = dist.get_backend(group)
    if backend == 'nccl':
        dist.reduce_scatter_tensor(out_shard, flat, op=dist.ReduceOp.SUM, group=group)
    else:                                   # gloo has no reduce-scatter: all-reduce, then keep the slice
        tmp = flat.clone()
        dist.all_reduce(tmp, op=dist.ReduceOp.SUM, group=group)
        out_shard.copy_(shards.shard_of(tmp))


def all_gather_params(flat, shards, group=None):
    """Every rank's updated slice back into everybody's full buffer (in place)."""
    backend = dist.get_backend(group)
    mine = shards.shard_of(flat)
    if backend == 'nccl':
        dist.all_gather_into_tensor(flat, mine, group=group)
    else:
        parts = [torch.empty_like(mine) for _ in range(shards.world_size)]
        dist.all_gather(parts, mine.clone(), group=group)
        flat.copy_(torch.cat(parts))


def sharded_optimizer_step(flat_param, flat_grad, grad_shard, shards, adam_fn, group=None):
    """One data-parallel optimiser step.  ``adam_fn(param_slice, grad_slice, local_ends, local_hparams)`` updates
    the slice in place.  ``flat_grad`` is cleared for the next iteration."""
    reduce_scatter_sum(grad_shard, flat_grad, shards, group)
    flat_grad.zero_()
    adam_fn(shards.shard_of(flat_param), grad_shard, shards.local_ends, shards.local_hparams)
    all_gather_params(flat_param, shards, group)
