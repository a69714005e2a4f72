"""torch-facing wrappers over the C ABI (autograd Functions + plain ops).

PyTorch is used for device memory, streams and autograd plumbing only; every
arithmetic step of the hot path runs in libb200ctc.so.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from ._lib import check, dtype_code, lib, on_tensor_device, ptr, require_cuda, stream_ptr


def _rows(x: torch.Tensor):
    V = x.shape[-1]
    return x.numel() // max(V, 1), V


class _SoftmaxFn(torch.autograd.Function):
    @staticmethod
    @on_tensor_device
    def forward(ctx, x, mode, bias=None):
        require_cuda(x, "softmax input")
        xc = x.contiguous()
        y = torch.empty_like(xc)
        rows, V = _rows(xc)
        if bias is None:
            check(lib().b200ctc_softmax_fwd(ptr(xc), ptr(y), rows, V, dtype_code(xc), mode,
                                            stream_ptr(xc.device)), "softmax_fwd")
        else:
            if bias.dtype != xc.dtype or bias.numel() != V or bias.device != xc.device:
                raise TypeError("softmax bias must be V values of the input's dtype on its device")
            bc = bias.contiguous()
            check(lib().b200ctc_softmax_fwd_bias(ptr(xc), ptr(bc), ptr(y), rows, V, dtype_code(xc), mode,
                                                 stream_ptr(xc.device)), "softmax_fwd_bias")
        ctx.mode = mode
        ctx.has_bias = bias is not None
        ctx.save_for_backward(y)
        return y

    @staticmethod
    @on_tensor_device
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dyc = dy.contiguous()
        dx = torch.empty_like(y)
        rows, V = _rows(y)
        check(lib().b200ctc_softmax_bwd(ptr(y), ptr(dyc), ptr(dx), rows, V, dtype_code(y),
                                        ctx.mode, stream_ptr(y.device)), "softmax_bwd")
        dbias = dx.reshape(-1, V).sum(0) if (ctx.has_bias and ctx.needs_input_grad[2]) else None
        return dx, None, dbias


def log_softmax(x: torch.Tensor, bias: torch.Tensor = None) -> torch.Tensor:
    """log-softmax over the last dim (espnet2/asr/ctc.py:117); with ``bias`` (V,) of ``x + bias`` -
    the projection's bias absorbed into the kernel's one read of the row."""
    return _SoftmaxFn.apply(x, 0, bias)


def softmax(x: torch.Tensor, bias: torch.Tensor = None) -> torch.Tensor:
    """softmax over the last dim (espnet2/asr/ctc.py:107); ``bias`` as in :func:`log_softmax`."""
    return _SoftmaxFn.apply(x, 1, bias)


@on_tensor_device
def argmax(x: torch.Tensor, bias: torch.Tensor = None) -> torch.Tensor:
    """argmax over the last dim, int64, first-max tie rule (espnet2/asr/ctc.py:127); with ``bias``
    (V,) of ``x + bias`` (see :func:`log_softmax`)."""
    require_cuda(x, "argmax input")
    xc = x.detach().contiguous()
    rows, V = _rows(xc)
    out = torch.empty(xc.shape[:-1], dtype=torch.int64, device=xc.device)
    if bias is None:
        check(lib().b200ctc_argmax(ptr(xc), ptr(out), rows, V, dtype_code(xc), stream_ptr(xc.device)),
              "argmax")
    else:
        if bias.dtype != xc.dtype or bias.numel() != V or bias.device != xc.device:
            raise TypeError("argmax bias must be V values of the input's dtype on its device")
        bc = bias.detach().contiguous()
        check(lib().b200ctc_argmax_bias(ptr(xc), ptr(bc), ptr(out), rows, V, dtype_code(xc),
                                        stream_ptr(xc.device)), "argmax_bias")
    return out


@on_tensor_device
def greedy_decode(logits: torch.Tensor, in_lens=None, blank: int = 0):
    """Best-path decode: (ids (B,T) padded with -1, lens (B,))."""
    require_cuda(logits, "greedy_decode input")
    xc = logits.detach().contiguous()
    B, T, V = xc.shape
    ids = torch.empty(B, T, dtype=torch.int64, device=xc.device)
    lens = torch.empty(B, dtype=torch.int64, device=xc.device)
    il = None if in_lens is None else in_lens.to(device=xc.device, dtype=torch.int64).contiguous()
    check(lib().b200ctc_greedy_decode(ptr(xc), ptr(il), B, T, V, dtype_code(xc), blank, ptr(ids),
                                      ptr(lens), stream_ptr(xc.device)), "greedy_decode")
    return ids, lens


def _upstream(g):
    """(fp32 device vector, stride) describing the upstream gradient of nll (B,)."""
    if g.dim() == 0 or g.numel() == 1 or g.stride(0) == 0:  # scalar / expanded scalar
        return g.reshape(-1)[:1].to(torch.float32).contiguous(), 0
    return g.to(torch.float32).contiguous(), 1


class _CTCLossFn(torch.autograd.Function):
    """nll (B,) fp32 = CTCLoss(reduction='none')(log_softmax(logits)); with
    ``reduce_scale`` the scalar ``reduce_scale * nll.sum()`` instead (written by
    the same launches).

    eager mode (default when the logits need a gradient): forward computes the
    loss AND d nll/d logits scaled by ``grad_hint`` in one call - the logits are
    read in full once, the gradient is written once (no (T,B,V) log-prob tensor
    ever exists); backward only rescales if the upstream gradient differs from
    the hint (device-side check, no host sync).
    deferred mode: forward keeps the lattices in a workspace, backward runs the
    gradient row pass with the real upstream gradient."""

    @staticmethod
    @on_tensor_device
    def forward(ctx, logits, targets, tgt_offsets, in_lens, tgt_lens, max_tgt_len, blank,
                zero_infinity, grad_hint, eager, reduce_scale, peer=None, hint_vec=None):
        require_cuda(logits, "CTC logits")
        x = logits.contiguous()
        B, T, V = x.shape
        dev = x.device
        need_grad = bool(ctx.needs_input_grad[0])
        nbytes = lib().b200ctc_loss_workspace_bytes(B, T, V, max_tgt_len)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        nll = torch.empty(B, dtype=torch.float32, device=dev)
        reduced = reduce_scale is not None
        loss = torch.zeros((), dtype=torch.float32, device=dev) if (reduced and B == 0) else (
            torch.empty((), dtype=torch.float32, device=dev) if reduced else None)
        scale = float(reduce_scale) if reduced else 0.0
        ctx.eager = bool(need_grad and eager and B > 0)
        ctx.gmul = float(reduce_scale) if reduced else 1.0
        if peer is not None and not (ctx.eager):  # ctc_loss() routes such calls to peer.exchange
            raise RuntimeError("the fused sharded-loss exchange needs the eager gradient mode, logits "
                               "that require a gradient and at least one utterance on every rank")
        if ctx.eager and hint_vec is not None:
            # one expected upstream gradient PER UTTERANCE (several heads in one batched call)
            if reduced or peer is not None:
                raise RuntimeError("per-utterance gradient hints: no batch reduction / exchange")
            grad = torch.empty_like(x)
            check(lib().b200ctc_loss_fwd_grad_hints(ptr(x), dtype_code(x), ptr(targets), ptr(tgt_offsets),
                                                    ptr(in_lens), ptr(tgt_lens), B, T, V, max_tgt_len,
                                                    blank, int(zero_infinity), ptr(hint_vec), ptr(nll),
                                                    ptr(grad), ptr(ws), nbytes, stream_ptr(dev)),
                  "loss_fwd_grad_hints")
            ctx.grad_buf = grad
            ctx.hint_vec = hint_vec
            ctx.cfg = (B, T, V, None)
            return nll
        if ctx.eager:
            ctx.hint_vec = None
            grad = torch.empty_like(x)
            if peer is not None:  # sharded: the loss kernel also all-reduces {sum, B} over the ranks
                check(lib().b200ctc_loss_fwd_grad_sharded(
                    ptr(x), dtype_code(x), ptr(targets), ptr(tgt_offsets), ptr(in_lens),
                    ptr(tgt_lens), B, T, V, max_tgt_len, blank, int(zero_infinity), grad_hint,
                    ptr(nll), ptr(loss), scale, ptr(grad), ptr(ws), nbytes, peer.handle,
                    ptr(peer.out_buffer), stream_ptr(dev)), "loss_fwd_grad_sharded")
            else:
                check(lib().b200ctc_loss_fwd_grad(ptr(x), dtype_code(x), ptr(targets),
                                                  ptr(tgt_offsets), ptr(in_lens), ptr(tgt_lens), B, T,
                                                  V, max_tgt_len, blank, int(zero_infinity), grad_hint,
                                                  ptr(nll), ptr(loss), scale, ptr(grad), ptr(ws),
                                                  nbytes, stream_ptr(dev)), "loss_fwd_grad")
            ctx.grad_buf = grad
            ctx.cfg = (B, T, V, grad_hint)
            return loss if reduced else nll
        check(lib().b200ctc_loss_fwd(ptr(x), dtype_code(x), ptr(targets), ptr(tgt_offsets),
                                     ptr(in_lens), ptr(tgt_lens), B, T, V, max_tgt_len, blank,
                                     int(zero_infinity), int(need_grad), ptr(nll), ptr(loss), scale,
                                     ptr(ws), nbytes, stream_ptr(dev)), "loss_fwd")
        if need_grad:
            ctx.save_for_backward(x, in_lens, ws)
            ctx.cfg = (B, T, V, max_tgt_len, blank, int(zero_infinity), nbytes)
        return loss if reduced else nll

    @staticmethod
    @on_tensor_device
    def backward(ctx, g):
        none10 = (None,) * 12
        gs, stride = _upstream(g)
        if ctx.eager:
            grad = ctx.grad_buf
            if grad is None:
                raise RuntimeError(
                    "the eager CTC gradient was already consumed by an earlier backward(); "
                    "use ctc_loss(..., grad_mode='deferred') to backpropagate twice")
            ctx.grad_buf = None
            B, T, V, hint = ctx.cfg
            if ctx.hint_vec is not None:
                check(lib().b200ctc_grad_rescale_hints(ptr(grad), dtype_code(grad), ptr(gs), stride,
                                                       ctx.gmul, ptr(ctx.hint_vec), B, T, V,
                                                       stream_ptr(grad.device)), "grad_rescale_hints")
                return (grad,) + none10
            check(lib().b200ctc_grad_rescale(ptr(grad), dtype_code(grad), ptr(gs), stride, ctx.gmul,
                                             hint, B, T, V, stream_ptr(grad.device)),
                  "grad_rescale")
            return (grad,) + none10
        x, in_lens, ws = ctx.saved_tensors
        B, T, V, lmax, blank, zinf, nbytes = ctx.cfg
        grad = torch.empty_like(x)
        check(lib().b200ctc_loss_bwd(ptr(x), dtype_code(x), ptr(in_lens), B, T, V, lmax, blank,
                                     zinf, ptr(gs), stride, ctx.gmul, ptr(grad), ptr(ws), nbytes,
                                     stream_ptr(x.device)), "loss_bwd")
        return (grad,) + none10


def _f32_recip(n: int) -> float:
    """fp32 1/n exactly as torch's `x / n` produces it."""
    return float(np.float32(1.0) / np.float32(max(n, 1)))


def ctc_loss(logits, targets, in_lens, tgt_lens, blank: int = 0, zero_infinity: bool = True,
             max_tgt_len: int | None = None, tgt_offsets=None, grad_hint: float | None = None,
             grad_mode: str = "eager", reduce_scale: float | None = None,
             peer=None) -> torch.Tensor:
    """Per-utterance CTC negative log-likelihood from LOGITS.

    logits (B,T,V) batch-major activations (fp32/fp16/bf16, CUDA); targets 1-D
    concatenated int64 labels; in_lens / tgt_lens (B,).  Equivalent to
    ``F.ctc_loss(logits.transpose(0,1).log_softmax(2), ..., reduction="none")``.
    ``max_tgt_len`` may be given to avoid a device->host sync.
    ``reduce_scale``: return the scalar ``reduce_scale * nll.sum()`` instead of
    the (B,) vector (``1/B`` gives espnet2/asr/ctc.py:56-60) with no extra launch.
    ``grad_hint``: the upstream gradient the caller expects for every nll[b]
    (default 1/B, i.e. ``nll.sum()/B``); any other upstream gradient is still
    exact, it only costs one rescaling pass.
    ``peer``: a ``sharded.PeerLossExchange``; the kernel that finishes the loss then also
    all-reduces [sum of losses, utterance count] over the data-parallel ranks through
    NVLink peer memory and leaves [global mean loss, total count] in ``peer.out``.
    """
    require_cuda(logits, "CTC logits")
    if grad_mode not in ("eager", "deferred"):
        raise ValueError(f"grad_mode must be 'eager' or 'deferred': {grad_mode}")
    dev = logits.device
    tl_dev = torch.as_tensor(tgt_lens).to(device=dev, dtype=torch.int64).contiguous()
    il_dev = torch.as_tensor(in_lens).to(device=dev, dtype=torch.int64).contiguous()
    tg_dev = torch.as_tensor(targets).to(device=dev, dtype=torch.int64).contiguous()
    if max_tgt_len is None:
        src = tgt_lens if not (isinstance(tgt_lens, torch.Tensor) and tgt_lens.is_cuda) else tl_dev
        max_tgt_len = int(torch.as_tensor(src).max()) if tl_dev.numel() else 0
    hint_vec = None
    if isinstance(grad_hint, torch.Tensor):  # one expected upstream gradient per utterance
        hint_vec = grad_hint.to(device=dev, dtype=torch.float32).contiguous()
        if hint_vec.numel() != logits.shape[0]:
            raise ValueError("grad_hint tensor must have one entry per utterance")
        grad_hint = 1.0
    if grad_hint is None:
        grad_hint = float(reduce_scale) if reduce_scale is not None else _f32_recip(logits.shape[0])
    # the exchange is fused into the loss kernel only in the eager gradient mode; any other call
    # with a `peer` (validation under no_grad, deferred mode, an empty shard) computes the local
    # loss and makes the SAME collective with the stand-alone exchange kernel
    fused = (peer is not None and grad_mode == "eager" and torch.is_grad_enabled()
             and logits.requires_grad and logits.shape[0] > 0)
    out = _CTCLossFn.apply(logits, tg_dev, tgt_offsets, il_dev, tl_dev, int(max_tgt_len),
                           int(blank), bool(zero_infinity), float(grad_hint),
                           grad_mode == "eager",
                           None if reduce_scale is None else float(reduce_scale),
                           peer if fused else None, hint_vec)
    if peer is not None and not fused:
        from .sharded import pack_sum

        total = out.detach() if reduce_scale is not None else out.detach().sum()
        inv = 1.0 / float(reduce_scale) if reduce_scale else 1.0
        peer.exchange(pack_sum(total, inv, logits.shape[0]))
    return out


@on_tensor_device
def forced_align(lp: torch.Tensor, target: torch.Tensor, blank: int = 0) -> torch.Tensor:
    """Viterbi alignment of one utterance; lp (T,V) fp32 log-probs, target (L,)."""
    require_cuda(lp, "forced_align input")
    lpc = lp.detach().to(torch.float32).contiguous()
    tg = target.to(device=lpc.device, dtype=torch.int64).contiguous()
    T, V = lpc.shape
    L = tg.numel()
    n = lib().b200ctc_align_workspace_bytes(T, L)
    ws = torch.empty(n, dtype=torch.uint8, device=lpc.device)
    out = torch.empty(T, dtype=torch.int64, device=lpc.device)
    check(lib().b200ctc_forced_align(ptr(lpc), ptr(tg), T, V, L, blank, ptr(out), ptr(ws), n,
                                     stream_ptr(lpc.device)), "forced_align")
    return out


launch_count = _lib.launch_count
reset_launch_count = _lib.reset_launch_count


# ---- beam pruning (espnet/nets/batch_beam_search.py:206-283) ----------
def _score_args(mats, weights):
    """host-side pointer / stride / weight arrays of the (n, V) fp32 score matrices"""
    import ctypes

    K = len(mats)
    for m in mats:
        require_cuda(m, "beam score matrix")
        if m.dtype != torch.float32 or m.dim() != 2:
            raise TypeError("beam score matrices must be 2-D fp32")
    ptrs = (ctypes.c_void_p * K)(*[m.data_ptr() for m in mats])
    strides = (ctypes.c_int64 * (2 * K))(*[s for m in mats for s in m.stride()])
    ws = (ctypes.c_float * K)(*[float(w) for w in weights])
    return ptrs, strides, ws, K


@on_tensor_device
def beam_prebeam(mats, weights, k: int) -> torch.Tensor:
    """Top-k token ids (n, k) of every row of ``sum_q weights[q] * mats[q]`` (pre-beam)."""
    ptrs, strides, ws, K = _score_args(mats, weights)
    n, V = mats[0].shape
    out = torch.empty(n, k, dtype=torch.int64, device=mats[0].device)
    check(lib().b200ctc_beam_prebeam(ptrs, strides, ws, K, n, V, k, ptr(out),
                                     stream_ptr(out.device)), "beam_prebeam")
    return out


@on_tensor_device
def beam_select(mats, weights, run_score, cum_in, yseq_in, cur_len, active, maxlen, step_i,
                U: int, W: int, k: int, eos: int, yseq_out):
    """One fused pruning step (include/b200ctc.h ``b200ctc_beam_select``).

    Returns (out_val (U,k), out_flat (U,k) int64, out_run (U,k), cum_out (K,U*k),
    pack (3+K,U,k)); ``yseq_out`` (a preallocated int64 buffer) receives the new prefixes."""
    ptrs, strides, ws, K = _score_args(mats, weights)
    n, V = mats[0].shape
    dev = mats[0].device
    out_val = torch.empty(U, k, dtype=torch.float32, device=dev)
    out_flat = torch.empty(U, k, dtype=torch.int64, device=dev)
    out_run = torch.empty(U, k, dtype=torch.float32, device=dev)
    cum_out = torch.empty(K, U * k, dtype=torch.float32, device=dev)
    pack = torch.empty(3 + K, U, k, dtype=torch.float32, device=dev)
    check(lib().b200ctc_beam_select(ptrs, strides, ws, K, ptr(run_score), ptr(cum_in), ptr(yseq_in),
                                    yseq_in.stride(0), cur_len, ptr(active), ptr(maxlen), step_i,
                                    U, W, V, k, eos, ptr(out_val), ptr(out_flat), ptr(out_run),
                                    ptr(cum_out), ptr(yseq_out), yseq_out.stride(0), ptr(pack),
                                    stream_ptr(dev)), "beam_select")
    return out_val, out_flat, out_run, cum_out, pack
