"""numpy restatement of the CTC training path (TEST INFRASTRUCTURE, see oracle/__init__).

What the reference computes (citations relative to /root/reference):

* ``espnet2/asr/ctc.py:52-62``  ``loss_fn``: ``th_pred.log_softmax(2)`` then
  ``torch.nn.CTCLoss(reduction="none", zero_infinity=...)`` then ``/B``.
* ``espnet2/asr/ctc.py:99-127`` ``softmax`` / ``log_softmax`` / ``argmax``.
* ``espnet/nets/pytorch_backend/ctc.py:56-65`` legacy ``loss_fn``.

The loss arithmetic lives in the reference's third-party dependency
``torch`` (``setup.py:24`` pins ``torch>=1.3.0``; 2.11.0 installed here):
``torch.nn.functional.ctc_loss`` implements Graves et al. 2006, eq. (6)-(16)
in log space.  This file restates that published algorithm independently;
``tests/test_oracle_golden.py`` pins it against fixtures produced by the
real reference modules.
"""
from __future__ import annotations

import numpy as np

NEG_INF = -np.inf


def log_softmax(x: np.ndarray, dtype=np.float32) -> np.ndarray:
    """Row-wise log-softmax over the last axis (espnet2/asr/ctc.py:54,117)."""
    x = np.asarray(x, dtype=dtype)
    m = np.max(x, axis=-1, keepdims=True)
    m = np.where(np.isfinite(m), m, 0).astype(dtype)
    z = x - m
    lse = np.log(np.sum(np.exp(z), axis=-1, keepdims=True, dtype=dtype))
    return (z - lse).astype(dtype)


def softmax(x: np.ndarray, dtype=np.float32) -> np.ndarray:
    """Row-wise softmax over the last axis (espnet2/asr/ctc.py:107)."""
    x = np.asarray(x, dtype=dtype)
    m = np.max(x, axis=-1, keepdims=True)
    m = np.where(np.isfinite(m), m, 0).astype(dtype)
    e = np.exp(x - m)
    return (e / np.sum(e, axis=-1, keepdims=True, dtype=dtype)).astype(dtype)


def argmax(x: np.ndarray) -> np.ndarray:
    """torch.argmax semantics (espnet2/asr/ctc.py:127): first index of the max,
    a NaN compares greater than everything (first NaN wins)."""
    x = np.asarray(x)
    nan = np.isnan(x)
    has_nan = nan.any(axis=-1)
    out = np.argmax(np.where(nan, -np.inf, x), axis=-1).astype(np.int64)
    if has_nan.any():
        out = np.where(has_nan, np.argmax(nan, axis=-1), out).astype(np.int64)
    return out


def _lse3(a, b, c):
    """log(exp a + exp b + exp c) with the max-shift used by the published
    recursion; -inf inputs are allowed."""
    m = np.maximum(np.maximum(a, b), c)
    m0 = np.where(np.isneginf(m), 0, m).astype(a.dtype)
    with np.errstate(divide="ignore"):
        return (np.log(np.exp(a - m0) + np.exp(b - m0) + np.exp(c - m0)) + m0).astype(a.dtype)


def extended_labels(target: np.ndarray, blank: int) -> np.ndarray:
    """blank, l1, blank, l2, ..., lL, blank  (length 2L+1)."""
    L = len(target)
    ext = np.full(2 * L + 1, blank, dtype=np.int64)
    ext[1::2] = target
    return ext


def ctc_alpha_beta(lp: np.ndarray, target: np.ndarray, blank: int = 0):
    """Log-space alpha/beta lattices for ONE utterance.

    lp: (T, V) log-probs of the valid frames only; target: (L,) int labels.
    Returns (log_alpha (T,S), log_beta (T,S), nll) with S = 2L+1.
    Graves 2006 eq. (6)-(8) forward, (9)-(11) backward.
    """
    dtype = lp.dtype
    T = lp.shape[0]
    L = len(target)
    S = 2 * L + 1
    ext = extended_labels(target, blank)
    # skip transition s-2 -> s allowed iff label differs (never for blanks)
    skip = np.zeros(S, dtype=bool)
    if S > 2:
        skip[2:] = ext[2:] != ext[:-2]
    ninf = np.array(NEG_INF, dtype=dtype)
    la = np.full((T, S), NEG_INF, dtype=dtype)
    lb = np.full((T, S), NEG_INF, dtype=dtype)
    if T == 0:
        return la, lb, dtype.type(np.inf if L > 0 else 0.0)
    em = lp[:, ext]  # (T, S) emissions gathered along the lattice
    la[0, 0] = em[0, 0]
    if L > 0:
        la[0, 1] = em[0, 1]
    for t in range(1, T):
        p = la[t - 1]
        p1 = np.concatenate(([ninf], p[:-1]))
        p2 = np.concatenate(([ninf, ninf], p[:-2]))[:S]
        p2 = np.where(skip, p2, ninf)
        la[t] = _lse3(p, p1, p2) + em[t]
    l1 = la[T - 1, S - 1]
    l2 = la[T - 1, S - 2] if L > 0 else ninf
    m = max(l1, l2)
    m = dtype.type(0) if np.isneginf(m) else m
    with np.errstate(divide="ignore"):
        nll = -(np.log(np.exp(l1 - m) + np.exp(l2 - m)) + m)
    nll = dtype.type(nll)
    lb[T - 1, S - 1] = em[T - 1, S - 1]
    if L > 0:
        lb[T - 1, S - 2] = em[T - 1, S - 2]
    skip_b = np.zeros(S, dtype=bool)  # s -> s+2 allowed
    if S > 2:
        skip_b[:-2] = ext[2:] != ext[:-2]
    for t in range(T - 2, -1, -1):
        n = lb[t + 1]
        n1 = np.concatenate((n[1:], [ninf]))
        n2 = np.concatenate((n[2:], [ninf, ninf]))[:S]
        n2 = np.where(skip_b, n2, ninf)
        lb[t] = _lse3(n, n1, n2) + em[t]
    return la, lb, nll


def ctc_loss_and_grad(
    logits: np.ndarray,
    targets,
    in_lens,
    blank: int = 0,
    zero_infinity: bool = True,
    grad_out=None,
    dtype=np.float32,
):
    """Per-utterance NLL and d(sum_b grad_out[b]*nll[b])/d(logits).

    logits: (B, T, V) pre-softmax activations (batch-major, as ``ctc_lo``
    emits them, espnet2/asr/ctc.py:82).  targets: list of B int arrays.
    The gradient is chained through log-softmax exactly as autograd does for
    espnet2/asr/ctc.py:54-55: ``g_lp = (exp(lp) - exp(lcab + nll - lp)) * g``
    for valid frames (0 beyond ``in_lens[b]``), then
    ``g_logit = g_lp - softmax * sum_v g_lp``.
    """
    logits = np.asarray(logits, dtype=dtype)
    B, T, V = logits.shape
    lp_all = log_softmax(logits, dtype=dtype)
    nll = np.zeros(B, dtype=dtype)
    grad = np.zeros_like(logits)
    if grad_out is None:
        grad_out = np.ones(B, dtype=dtype)
    grad_out = np.broadcast_to(np.asarray(grad_out, dtype=dtype), (B,))
    for b in range(B):
        Tb = int(in_lens[b])
        tgt = np.asarray(targets[b], dtype=np.int64)
        lp = lp_all[b, :Tb]
        la, lb, n = ctc_alpha_beta(lp, tgt, blank)
        if np.isinf(n) and zero_infinity:
            nll[b] = 0
            continue
        nll[b] = n
        if Tb == 0:
            continue
        ext = extended_labels(tgt, blank)
        g_lp = np.zeros((T, V), dtype=dtype)
        ab = la + lb  # (Tb, S)
        lcab = np.full((Tb, V), NEG_INF, dtype=dtype)
        for s, c in enumerate(ext):
            lcab[:, c] = np.logaddexp(lcab[:, c], ab[:, s]).astype(dtype)
        with np.errstate(over="ignore", invalid="ignore"):
            g_lp[:Tb] = (np.exp(lp) - np.exp(lcab + n - lp)) * grad_out[b]
        sm = np.exp(lp_all[b])
        grad[b] = g_lp - sm * np.sum(g_lp, axis=-1, keepdims=True, dtype=dtype)
    return nll, grad


def espnet2_loss(nll: np.ndarray, reduce: bool = True):
    """``loss.sum()/B`` or ``loss/B`` (espnet2/asr/ctc.py:56-62)."""
    B = nll.shape[0]
    return nll.sum() / B if reduce else nll / B


def greedy_collapse(ids: np.ndarray, blank: int = 0):
    """Best-path decode: unique-consecutive then drop blanks
    (espnet/nets/e2e_asr_common.py:159-191 ``calculate_cer_ctc``)."""
    out = []
    prev = None
    for i in ids.tolist():
        if i != prev and i != blank:
            out.append(i)
        prev = i
    return out


def viterbi_align(lp: np.ndarray, target: np.ndarray, blank: int = 0):
    """Forced alignment, max-semiring lattice with back-pointers
    (espnet/nets/pytorch_backend/ctc.py:163-226 ``forced_align``; that code
    uses fp64 lattice values and a finite log-zero of -1e11)."""
    T = lp.shape[0]
    ext = extended_labels(np.asarray(target), blank)
    S = len(ext)
    logzero = -100000000000.0
    delta = np.zeros((T, S)) + logzero
    back = np.zeros((T, S), dtype=np.int16) - 1
    delta[0, 0] = lp[0, ext[0]]
    if S > 1:
        delta[0, 1] = lp[0, ext[1]]
    for t in range(1, T):
        for s in range(S):
            if ext[s] == blank or s < 2 or ext[s] == ext[s - 2]:
                cand = [delta[t - 1, s], delta[t - 1, s - 1]]  # s-1 == -1 wraps like the reference
                prev = [s, s - 1]
            else:
                cand = [delta[t - 1, s], delta[t - 1, s - 1], delta[t - 1, s - 2]]
                prev = [s, s - 1, s - 2]
            k = int(np.argmax(cand))
            delta[t, s] = cand[k] + lp[t, ext[s]]
            back[t, s] = prev[k]
    seq = np.zeros(T, dtype=np.int64)
    cand = [delta[-1, S - 1], delta[-1, S - 2]]
    seq[-1] = [S - 1, S - 2][int(np.argmax(cand))]
    for t in range(T - 2, -1, -1):
        seq[t] = back[t + 1, seq[t + 1]]
    return [int(ext[s]) for s in seq]
