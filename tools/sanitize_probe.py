#!/usr/bin/env python3
"""Small shapes through the newest kernels, for `compute-sanitizer --tool memcheck|racecheck`:
split lattice (2 CTAs per lattice, ragged), WIDE row pass, dense-gather switch, fused beam
pruning / pre-beam, register argmax."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from espnet1_b200 import functional as F  # noqa: E402

dev = "cuda:0"
g = torch.Generator().manual_seed(1)
B, T, V = 3, 700, 96
tl = torch.tensor([600, 520, 130])
il = torch.tensor([700, 690, 400])
tg = torch.randint(1, V, (int(tl.sum()),), generator=g)
x = (torch.randn(B, T, V, generator=g) * 2).to(dev).requires_grad_(True)
nll = F.ctc_loss(x, tg.to(dev), il.to(dev), tl.to(dev))
(nll.sum() / B).backward()
with torch.no_grad():
    nll2 = F.ctc_loss(x.detach(), tg.to(dev), il.to(dev), tl.to(dev))
print("loss", nll.tolist(), "fwd-only equal", torch.allclose(nll, nll2, rtol=1e-6))
print("argmax", F.argmax(torch.randn(5, 5000, device=dev)).tolist())
U, W, Vb = 2, 3, 200
mats = [torch.randn(U * W, Vb, device=dev), torch.tensor([1.0], device=dev).expand(U * W, Vb)]
ids = F.beam_prebeam(mats, [0.5, 0.2], 7)
yin = torch.randint(0, Vb, (U * W, 6), device=dev)
yout = torch.zeros(U * W, 8, dtype=torch.int64, device=dev)
out = F.beam_select(mats, [0.5, 0.2], torch.zeros(U * W, device=dev), torch.zeros(2, U * W, device=dev), yin, 4,
                    torch.ones(U, dtype=torch.uint8, device=dev), torch.full((U,), 9, dtype=torch.int32, device=dev),
                    2, U, W, W, Vb - 1, yout)
torch.cuda.synchronize()
print("beam", ids.shape, out[1].tolist())
