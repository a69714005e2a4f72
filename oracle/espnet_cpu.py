"""Port of the reference's CPU path on top of torch CPU kernels.

TEST INFRASTRUCTURE / CPU BASELINE ONLY (see oracle/__init__).  The reference
does its CTC arithmetic by calling its pinned third-party dependency ``torch``
(``setup.py:24``); this file restates the thin wrappers the reference puts
around those calls so the same CPU arithmetic can be run (and timed) on a box
where ``/root/reference`` does not exist.  Citations relative to
/root/reference.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

LOGZERO = -10000000000.0  # espnet/nets/ctc_prefix_score.py:33


# --------------------------------------------------------------------------
# training path
# --------------------------------------------------------------------------
def loss_fn(th_pred, th_target, th_ilen, th_olen, reduce=True, zero_infinity=True):
    """espnet2/asr/ctc.py:52-62 (``ctc_type == "builtin"``).

    th_pred: (T, B, V) activations; th_target: (sum L,) ; lengths (B,).
    """
    lp = th_pred.log_softmax(2)
    loss = F.ctc_loss(
        lp, th_target, th_ilen, th_olen, reduction="none", zero_infinity=zero_infinity
    )
    size = lp.size(1)
    return loss.sum() / size if reduce else loss / size


def ctc_forward(hs_pad, hlens, ys_pad, ys_lens, weight, bias, reduce=True,
                zero_infinity=True, dropout_rate=0.0):
    """espnet2/asr/ctc.py:72-97: Linear -> transpose -> concat targets -> loss_fn."""
    ys_hat = F.linear(F.dropout(hs_pad, p=dropout_rate), weight, bias)
    ys_hat = ys_hat.transpose(0, 1)
    ys_true = torch.cat([ys_pad[i, :l] for i, l in enumerate(ys_lens)])
    return loss_fn(ys_hat, ys_true, hlens, ys_lens, reduce, zero_infinity).to(
        device=hs_pad.device, dtype=hs_pad.dtype
    )


def legacy_ctc_forward(hs_pad, hlens, ys_pad, weight, bias, reduce=True, ignore_id=-1):
    """espnet/nets/pytorch_backend/ctc.py:73-133 (builtin): labels parsed from
    ``-1`` padding, ``CTCLoss(reduction="sum" if reduce else "none",
    zero_infinity=True)``, ``/B``, then ``.sum()`` when reduce."""
    ys = [y[y != ignore_id] for y in ys_pad]
    ys_hat = F.linear(hs_pad, weight, bias).transpose(0, 1)
    olens = torch.tensor([len(s) for s in ys], dtype=torch.long)
    lp = ys_hat.log_softmax(2)
    loss = F.ctc_loss(lp, torch.cat(ys), hlens.long(), olens,
                      reduction="sum" if reduce else "none", zero_infinity=True)
    loss = loss / lp.size(1)
    return loss.sum() if reduce else loss


def loss_and_backward(x_leaf, targets_1d, in_lens, tgt_lens, zero_infinity=True):
    """The timed part of the CPU baseline: ``x_leaf`` is a (B,T,V) leaf that already requires a
    gradient and whose ``.grad`` is None (the caller prepares both OUTSIDE the clock): log_softmax
    + ctc_loss + ``sum()/B`` + backward, espnet2/asr/ctc.py:54-60.  No input copy is made - the
    reference's forward makes none; the gradient buffer is allocated by autograd, as it is for
    the reference's (non-leaf) logits.  Returns nll (B,); the gradient is in ``x_leaf.grad``."""
    lp = x_leaf.transpose(0, 1).log_softmax(2)
    nll = F.ctc_loss(lp, targets_1d, in_lens, tgt_lens, reduction="none",
                     zero_infinity=zero_infinity)
    (nll.sum() / x_leaf.size(0)).backward()
    return nll.detach()


def logits_loss_and_grad(logits_btv, targets_1d, in_lens, tgt_lens, zero_infinity=True):
    """Timed CPU baseline at the logits boundary: log_softmax + ctc_loss
    (reduction none) + ``sum()/B`` + backward, exactly the arithmetic of
    espnet2/asr/ctc.py:54-60 with autograd providing the gradient.
    Returns (nll (B,), grad (B,T,V))."""
    x = logits_btv.detach().clone().requires_grad_(True)
    lp = x.transpose(0, 1).log_softmax(2)
    nll = F.ctc_loss(lp, targets_1d, in_lens, tgt_lens, reduction="none",
                     zero_infinity=zero_infinity)
    (nll.sum() / x.size(0)).backward()
    return nll.detach(), x.grad
