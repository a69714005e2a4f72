#!/usr/bin/env python3
"""Per-step device times of a K=20 region that starts from an idle GPU (sync before), autograd step."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import C2, synth  # noqa: E402
from espnet1_b200.espnet2.asr.ctc import CTC  # noqa: E402

dev = torch.device("cuda:0")
logits, tg, il, tl = synth(C2, 1234)
x = logits.to(dev).requires_grad_(True)
tg, il, tl = tg.to(dev), il.to(dev), tl.to(dev)
m = CTC(C2["V"], 8).to(dev)


def step():
    x.grad = None
    loss = m.loss_fn(x.transpose(0, 1), tg, il, tl, max_tgt_len=C2["lmax"])
    loss.backward()


for _ in range(8):
    step()
for rep in range(3):
    torch.cuda.synchronize()
    if rep == 2:
        time.sleep(0.05)  # a longer idle period before the region
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(21)]
    ev[0].record()
    for i in range(20):
        step()
        ev[i + 1].record()
    torch.cuda.synchronize()
    d = [ev[i].elapsed_time(ev[i + 1]) for i in range(20)]
    print("rep", rep, "first 6 steps ms:", [round(v, 3) for v in d[:6]], "mean of last 10:", round(sum(d[10:]) / 10, 4),
          "total/20:", round(ev[0].elapsed_time(ev[20]) / 20, 4))
