#!/usr/bin/env python3
"""Intermediate-CTC pattern at C2: H heads (encoder taps) against the same labels, sequential
``CTC.forward`` calls vs ``CTC.forward_multi`` (heads on alternating streams)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from espnet1_b200.espnet2.asr.ctc import CTC  # noqa: E402

dev = torch.device("cuda:0")
B, T, V, D, H = 64, 500, 5000, 256, int(os.environ.get("HEADS", "3"))
g = torch.Generator().manual_seed(3)
LOGITS = os.environ.get("LOGITS", "1") == "1"  # heads given as logits (ctc_lo = identity): the hot path alone
if LOGITS:
    D = V
hs = [(torch.randn(B, T, D, generator=g) * (2 if LOGITS else 1)).to(dev).requires_grad_(True) for _ in range(H)]
hl = torch.full((B,), T, dtype=torch.long, device=dev)
yl = torch.randint(75, 151, (B,), generator=g)
ys = torch.full((B, 150), -1, dtype=torch.long)
for b in range(B):
    ys[b, : yl[b]] = torch.randint(1, V, (int(yl[b]),), generator=g)
ys, yl = ys.to(dev), yl.to(dev)
ctc = CTC(V, 8 if LOGITS else D).to(dev)
if LOGITS:
    ctc.ctc_lo = torch.nn.Identity()


def seq():
    for h in hs:
        h.grad = None
    ctc.zero_grad(set_to_none=True)
    losses = [ctc(h, hl, ys, yl) for h in hs]
    sum(losses).backward()
    return losses


def multi():
    for h in hs:
        h.grad = None
    ctc.zero_grad(set_to_none=True)
    losses = ctc.forward_multi(hs, hl, ys, yl)
    sum(losses).backward()
    return losses


def streams():
    for h in hs:
        h.grad = None
    ctc.zero_grad(set_to_none=True)
    losses = ctc.forward_multi_streams(hs, hl, ys, yl)
    sum(losses).backward()
    return losses


# what forward_multi runs AFTER its single ctc_lo GEMM: one batched loss call on (H*B, T, V) logits
# with per-utterance hints (in LOGITS mode forward_multi itself would first copy the 640 MB "hidden
# states" per head in torch.cat - in real use those are 32 MB and the GEMM writes the stacked logits)
from espnet1_b200 import functional as Fn  # noqa: E402

big = torch.cat([h.detach() for h in hs], 0).requires_grad_(True) if LOGITS else None
tg1 = torch.cat([ys[b, : yl[b]] for b in range(B)])
hintv = torch.full((H * B,), Fn._f32_recip(B), device=dev)


def stacked():
    big.grad = None
    nll = Fn.ctc_loss(big, tg1.repeat(H), hl.repeat(H), yl.repeat(H), max_tgt_len=150, grad_hint=hintv)
    per = nll.view(H, B)
    losses = [per[h].sum() * Fn._f32_recip(B) for h in range(H)]
    sum(losses).backward()
    return losses


def timeit(fn, iters=20):
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


a = [float(v) for v in seq()]
ga = [h.grad.clone() for h in hs]
out = {"heads": H, "logits_given": LOGITS, "sequential_ms": timeit(seq), "alternating_streams_ms": timeit(streams)}
if LOGITS:
    c = [float(v) for v in stacked()]
    gc = big.grad.view(H, B, T, V)
    out["stacked_one_batched_call_ms"] = timeit(stacked)
    out["stacked_losses_rel_diff"] = max(abs(x - y) / abs(x) for x, y in zip(a, c))
    out["stacked_input_grads_bit_identical"] = all(torch.equal(ga[h], gc[h]) for h in range(H))
else:  # hidden states in, ctc_lo GEMM included: forward_multi as a user calls it
    out["forward_multi_ms"] = timeit(multi)
print(out)
