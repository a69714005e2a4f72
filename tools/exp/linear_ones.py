#!/usr/bin/env python3
"""The CTC projection in TRAINING (forward + backward of ctc_lo at C2: 32 000 frames, 256 -> 5000,
fp32): nn.Linear (cuBLASLt adds the bias in a separate pass, autograd sums the gradient's columns
for d bias) against a GEMM over [hs, 1] x [W, b]^T (the bias as a 257th input column)."""
import torch

B, T, D, V = 64, 500, 256, 5000
dev = "cuda"
lin = torch.nn.Linear(D, V).to(dev)
hs = torch.randn(B, T, D, device=dev, requires_grad=True)
gout = torch.randn(B, T, V, device=dev)


def plain():
    y = lin(hs)
    y.backward(gout)


def ones_col():
    x1 = torch.cat((hs, hs.new_ones(B, T, 1)), -1)
    w1 = torch.cat((lin.weight, lin.bias.unsqueeze(1)), 1)
    y = torch.nn.functional.linear(x1, w1)
    y.backward(gout)


def fwd_plain():
    with torch.no_grad():
        lin(hs)


def fwd_ones():
    with torch.no_grad():
        torch.nn.functional.linear(torch.cat((hs, hs.new_ones(B, T, 1)), -1),
                                   torch.cat((lin.weight, lin.bias.unsqueeze(1)), 1))


for name, fn in (("fwd nn.Linear", fwd_plain), ("fwd ones column", fwd_ones), ("fwd+bwd nn.Linear", plain),
                 ("fwd+bwd ones column", ones_col)):
    for _ in range(3):
        lin.zero_grad(); hs.grad = None
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:22s} {e0.elapsed_time(e1) / 10:.3f} ms")
with torch.no_grad():
    a = lin(hs)
    b = torch.nn.functional.linear(torch.cat((hs, hs.new_ones(B, T, 1)), -1),
                                   torch.cat((lin.weight, lin.bias.unsqueeze(1)), 1))
    print("max |diff|", float((a - b).abs().max()), "max |y|", float(a.abs().max()))
