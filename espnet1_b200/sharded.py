"""Batch-sharded CTC training loss: the only exchange step of the path.

Each rank owns B_local utterances (logits, labels and the gradient w.r.t. the
logits stay rank-local); ONE all-reduce of two fp32 scalars
``[sum_b nll_b, B_local]`` yields the batch-size-weighted global loss - the
semantics of /root/reference/espnet2/train/trainer.py:604-613 together with
espnet2/torch_utils/recursive_op.py:8-47 (stats*weight summed, weight summed,
then divided) applied to ``CTC.forward``'s per-rank ``sum()/B_local``
(espnet2/asr/ctc.py:56-60).  ``torch.distributed`` (NCCL over NVLink on the
GPU box, gloo in the CPU tests) is plumbing only.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int):
    """Contiguous utterance range [lo, hi) of ``rank`` (first ranks get the remainder)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def pack_local(nll_local: torch.Tensor) -> torch.Tensor:
    """[sum of local per-utterance losses, local weight (= utterance count)]."""
    return torch.stack([nll_local.float().sum(),
                        torch.tensor(float(nll_local.numel()), device=nll_local.device)])


def reduce_sharded_loss(nll_local: torch.Tensor, group=None):
    """Global batch-averaged CTC loss from rank-local per-utterance NLLs.

    Returns (loss, total_weight); identical on every rank.  Without an
    initialised process group (single process) no collective is issued.
    """
    buf = pack_local(nll_local)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf[0] / buf[1], buf[1]


def ddp_loss_for_backward(nll_local: torch.Tensor, group=None) -> torch.Tensor:
    """Loss whose backward reproduces the reference's DDP scaling: local mean
    times ``weight_local/weight_total * world_size`` (trainer.py:614-617), to be
    used when parameter gradients are averaged over ranks by DDP."""
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    _, total = reduce_sharded_loss(nll_local.detach(), group)
    return nll_local.sum() / total * world
