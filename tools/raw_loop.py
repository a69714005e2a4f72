#!/usr/bin/env python3
"""C2 step through the C ABI only (no autograd, preallocated buffers), plain launches vs one
CUDA graph replay per step: separates host / launch-gap time from kernel time."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import C2, synth  # noqa: E402
from espnet1_b200 import _lib  # noqa: E402
from espnet1_b200._lib import check, ptr  # noqa: E402

L = _lib.lib()
dev = torch.device("cuda:0")
cfg = C2
B, T, V = cfg["B"], cfg["T"], cfg["V"]
logits, tg, il, tl = synth(cfg, 1234)
x = logits.to(dev)
tg, il, tl = tg.to(dev), il.to(dev), tl.to(dev)
n = L.b200ctc_loss_workspace_bytes(B, T, V, cfg["lmax"])
ws = torch.empty(n, dtype=torch.uint8, device=dev)
nll = torch.empty(B, device=dev)
loss = torch.empty((), device=dev)
grad = torch.empty_like(x)
one = torch.ones((), device=dev)


def step(stream):
    check(L.b200ctc_loss_fwd_grad(ptr(x), 0, ptr(tg), None, ptr(il), ptr(tl), B, T, V, cfg["lmax"], 0, 1,
                                  1.0 / B, ptr(nll), ptr(loss), 1.0 / B, ptr(grad), ptr(ws), n, stream), "fwd_grad")
    check(L.b200ctc_grad_rescale(ptr(grad), 0, ptr(one), 0, 1.0 / B, 1.0 / B, B, T, V, stream), "rescale")


def timeit(fn, iters=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


s0 = torch.cuda.current_stream().cuda_stream
print("plain launches  ms/step:", timeit(lambda: step(s0)))
side = torch.cuda.Stream()
g = torch.cuda.CUDAGraph()
with torch.cuda.stream(side):
    step(side.cuda_stream)
    torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=side):
        step(side.cuda_stream)
print("graph replay    ms/step:", timeit(g.replay), "loss", float(loss))
