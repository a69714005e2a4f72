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
wa = None if LOGITS else ctc.ctc_lo.weight.grad.clone()
b = [float(v) for v in multi()]
same = a == b and all(torch.equal(x, h.grad) for x, h in zip(ga, hs))
print({"heads": H, "sequential_ms": timeit(seq), "multi_ms": timeit(multi), "identical_losses_and_input_grads": same,
       "logits_given": LOGITS,
       "weight_grad_max_diff": None if LOGITS else float((wa - ctc.ctc_lo.weight.grad).abs().max())})
