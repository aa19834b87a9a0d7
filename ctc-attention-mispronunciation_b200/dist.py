"""Batch sharding of the CTC path over the GPUs of one box (one process per GPU, torch.distributed).

Utterances are independent in every stage, so the path shards with no data-path collective; the only exchange is the
scalar loss: each rank's kernel writes its local sum into a one-element buffer that is all-reduced
over NCCL/NVLink.  The gradient needs no collective (each rank owns its slice).  The reference has no multi-GPU code.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_bounds(n_items: int, world_size: int, rank: int):
    """Contiguous slice [lo, hi) of rank; sizes differ by at most one."""
    base, rem = divmod(n_items, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def balanced_buckets(lengths, world_size: int, bucket_width: int = 64):
    """Length-bucketed assignment of ragged utterances to ranks (BASELINE config 5): sort by bucket, deal buckets
    round-robin so every rank gets ~equal total frames.  Returns a list of index lists, one per rank."""
    order = sorted(range(len(lengths)), key=lambda i: (int(lengths[i]) // bucket_width, i))
    shards = [[] for _ in range(world_size)]
    loads = [0] * world_size
    for i in reversed(order):                      # longest first, always to the lightest rank (LPT)
        r = min(range(world_size), key=lambda k: (loads[k], k))
        shards[r].append(i)
        loads[r] += int(lengths[i])
    for s in shards:
        s.sort(key=lambda i: (int(lengths[i]) // bucket_width, i))
    return shards


def all_reduce_loss(local_sum: torch.Tensor, group=None) -> torch.Tensor:
    """In-place SUM all-reduce of the one-element loss buffer (NCCL on GPUs, gloo in the CPU tests)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(local_sum, op=dist.ReduceOp.SUM, group=group)
    return local_sum


class ShardedCTCLoss:
    """Data-parallel CTC loss: every rank runs the kernel on its batch slice; the summed loss is all-reduced.

    ``loss_fn_raw`` defaults to mddctc.loss.ctc_loss_raw; it is injectable so that the sharding / reduction logic can be
    exercised on CPU (gloo) without a GPU.
    """

    def __init__(self, blank=0, zero_infinity=False, group=None, loss_fn_raw=None):
        self.blank, self.zero_infinity, self.group = blank, zero_infinity, group
        if loss_fn_raw is None:
            from .loss import ctc_loss_raw
            loss_fn_raw = ctc_loss_raw
        self._raw = loss_fn_raw

    def __call__(self, log_probs, targets, input_lengths, target_lengths, grad_scalar=1.0):
        """Arguments are this rank's LOCAL slice.  Returns (global loss sum [1], local nll, local grad)."""
        buf = torch.zeros(1, dtype=torch.float32, device=log_probs.device)
        nll, grad, _ = self._raw(log_probs, targets, input_lengths, target_lengths, blank=self.blank,
                                 zero_infinity=self.zero_infinity, grad_scalar=grad_scalar, loss_sum=buf, validate=False)
        all_reduce_loss(buf, self.group)
        return buf, nll, grad
