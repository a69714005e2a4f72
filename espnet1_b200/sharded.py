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


_CONST: dict = {}


def _pack_consts(n_local: int, device: torch.device):
    """Persistent [n, 0] and [0, n] device tensors: packing must not touch the host
    (a Python scalar written into a CUDA tensor is a synchronous pageable H2D copy,
    which stalls the launch queue of a 0.4 ms step)."""
    key = (int(n_local), str(device))
    c = _CONST.get(key)
    if c is None:
        c = (torch.tensor([float(n_local), 0.0], device=device),
             torch.tensor([0.0, float(n_local)], device=device))
        _CONST[key] = c
    return c


def pack_local(nll_local: torch.Tensor) -> torch.Tensor:
    """[sum of local per-utterance losses, local weight (= utterance count)]."""
    n = nll_local.numel()
    scale, offs = _pack_consts(n, nll_local.device)
    return torch.addcmul(offs, scale, nll_local.float().sum() / max(n, 1))


def pack_sum(total: torch.Tensor, mul: float, n_local: int) -> torch.Tensor:
    """[total * mul, n_local] from a device scalar, one launch, no host round trip."""
    key = ("sum", float(mul), int(n_local), str(total.device))
    c = _CONST.get(key)
    if c is None:
        c = (torch.tensor([float(mul), 0.0], device=total.device),
             torch.tensor([0.0, float(n_local)], device=total.device))
        _CONST[key] = c
    return torch.addcmul(c[1], c[0], total.float().reshape(()))


def pack_local_mean(loss_mean: torch.Tensor, n_local: int) -> torch.Tensor:
    """Same pair from the rank-local batch mean that ``CTC.forward`` returns
    (espnet2/asr/ctc.py:56-60): [mean * n, n] in one launch, no host round trip."""
    scale, offs = _pack_consts(n_local, loss_mean.device)
    return torch.addcmul(offs, scale, loss_mean.detach().float())


def reduce_sharded_mean(loss_mean: torch.Tensor, n_local: int, group=None):
    """Global batch-weighted loss from the rank-local means (see reduce_sharded_loss)."""
    buf = pack_local_mean(loss_mean, n_local)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf[0] / buf[1], buf[1]


def reduce_sharded_loss(nll_local: torch.Tensor, group=None):
    """Global batch-averaged CTC loss from rank-local per-utterance NLLs.

    Returns (loss, total_weight); identical on every rank.  Without an
    initialised process group (single process) no collective is issued.
    """
    buf = pack_local(nll_local)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf[0] / buf[1], buf[1]


class PeerLossExchange:
    """The sharded-loss all-reduce as the tail of the loss kernel (csrc/peer.cuh).

    One per process and process group.  Construction is collective: every rank
    creates a small exchange buffer on its GPU, the cudaIpc handles are
    all-gathered through ``torch.distributed`` (plumbing) and mapped.  After
    ``F.ctc_loss(..., peer=ex)`` (or ``CTC.loss_exchange = ex``) the two floats
    ``ex.out`` = [global batch-weighted mean loss, total utterance count] are valid
    on the stream of that call, identical on all ranks - the numbers of
    ``reduce_sharded_loss`` without a collective launch or any host work.
    """

    def __init__(self, group=None, device=None, deferred: bool = True, timeout_s: float = 0.0):
        """``deferred`` (default): the loss kernel only pushes this rank's pair; the wait for the
        peers - a per-step barrier on the slowest rank that nothing on the device needs - happens
        when ``out`` is read (and, for the previous step, at the start of the next exchange).
        ``timeout_s``: how long a wait may last before it gives NaN + a sticky error flag
        (0 = the library default, 30 s)."""
        import ctypes

        from ._lib import check, lib
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        handle = ctypes.c_void_p()
        raw = (ctypes.c_ubyte * 64)()
        with torch.cuda.device(self.device):
            check(lib().b200ctc_peer_create(self.world, self.rank, ctypes.byref(handle), raw),
                  "peer_create")
            self.handle = handle
            mine = bytes(raw)
            if self.world > 1:
                gathered = [None] * self.world
                dist.all_gather_object(gathered, mine, group=group)
                blob = b"".join(gathered)
            else:
                blob = mine
            buf = (ctypes.c_ubyte * len(blob)).from_buffer_copy(blob)
            check(lib().b200ctc_peer_connect(self.handle, buf), "peer_connect")
            check(lib().b200ctc_peer_configure(self.handle, int(bool(deferred)),
                                               int(round(timeout_s * 1000))), "peer_configure")
        self.deferred = bool(deferred)
        # written by the loss / exchange kernels; read it through ``out``
        self.out_buffer = torch.zeros(2, dtype=torch.float32, device=self.device)
        if self.world > 1:
            dist.barrier(group=group)  # every rank has mapped every buffer before the first push

    @property
    def out(self) -> torch.Tensor:
        """[global batch-weighted mean loss, total utterance count] of the LATEST exchange, valid on
        the current stream.  In deferred mode reading it issues the (tiny) wait kernel there."""
        if self.deferred:
            from ._lib import check, lib, ptr, stream_ptr
            with torch.cuda.device(self.device):
                check(lib().b200ctc_peer_wait(self.handle, ptr(self.out_buffer),
                                              stream_ptr(self.device)), "peer_wait")
        return self.out_buffer

    def exchange(self, pair: torch.Tensor) -> torch.Tensor:
        """Stand-alone collective on a device pair [sum, weight]: ranks without utterances, and
        calls that cannot use the fused form (no gradient wanted, deferred gradient mode)."""
        from ._lib import check, lib, ptr, stream_ptr
        pair = pair.to(device=self.device, dtype=torch.float32).contiguous()
        with torch.cuda.device(self.device):
            check(lib().b200ctc_peer_exchange(self.handle, ptr(pair), ptr(self.out_buffer),
                                              stream_ptr(self.device)), "peer_exchange")
        return self.out_buffer

    def timed_out(self) -> bool:
        """True if a wait ever ran into the time-out (synchronises the device)."""
        from ._lib import lib
        return lib().b200ctc_peer_error(self.handle) != 0

    def raise_if_timed_out(self) -> None:
        """Call at a natural synchronisation point (end of epoch, logging): a peer that never
        showed up turned ``out`` into NaN; this makes it an exception."""
        if self.timed_out():
            raise RuntimeError("sharded CTC loss exchange: a rank did not push its loss within the "
                               "time-out; `out` is NaN (see PeerLossExchange(timeout_s=...))")

    def close(self):
        from ._lib import lib
        if getattr(self, "handle", None) is not None:
            lib().b200ctc_peer_destroy(self.handle)
            self.handle = None

    def __del__(self):  # pragma: no cover - best effort
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def ddp_loss_for_backward(nll_local: torch.Tensor, group=None) -> torch.Tensor:
    """Loss whose backward reproduces the reference's DDP scaling: local mean
    times ``weight_local/weight_total * world_size`` (trainer.py:614-617), to be
    used when parameter gradients are averaged over ranks by DDP."""
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    _, total = reduce_sharded_loss(nll_local.detach(), group)
    return nll_local.sum() / total * world
