#!/usr/bin/env python3
"""Host-side cost of one autograd step (CTC.loss_fn + backward) at C2: wall time per call with the
GPU queue empty (sync before every step) and a cProfile of 200 steps."""
import cProfile
import os
import pstats
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


for _ in range(10):
    step()
torch.cuda.synchronize()
fw, bw = [], []
for _ in range(30):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    x.grad = None
    loss = m.loss_fn(x.transpose(0, 1), tg, il, tl, max_tgt_len=C2["lmax"])
    t1 = time.perf_counter()
    loss.backward()
    t2 = time.perf_counter()
    fw.append(t1 - t0)
    bw.append(t2 - t1)
fw.sort(); bw.sort()
print("host us per step (median): forward", round(fw[15] * 1e6, 1), "backward", round(bw[15] * 1e6, 1))
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    step()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
