#!/usr/bin/env python3
"""Minimal C2 fwd+bwd loop for ncu captures / quick timing:  python tools/prof_step.py [steps] [B T V lmin lmax]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import C2, synth  # noqa: E402
from espnet1_b200 import functional as F  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
cfg = dict(C2)
if len(sys.argv) > 6:
    cfg = dict(B=int(sys.argv[2]), T=int(sys.argv[3]), V=int(sys.argv[4]), lmin=int(sys.argv[5]), lmax=int(sys.argv[6]))
dev = torch.device("cuda:0")
logits, tg, il, tl = synth(cfg, 1234)
x = logits.to(dev).requires_grad_(True)
tg, il, tl = tg.to(dev), il.to(dev), tl.to(dev)
import ctypes  # noqa: E402

from espnet1_b200 import _lib  # noqa: E402

L = _lib.lib()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
for i in range(steps):
    if i == 2 and os.environ.get("PHASES"):
        L.b200ctc_profile_enable(1)
    ev[i].record()
    x.grad = None
    nll = F.ctc_loss(x, tg, il, tl, max_tgt_len=cfg["lmax"], reduce_scale=1.0 / cfg["B"])
    nll.backward()
ev[steps].record()
torch.cuda.synchronize()
print("ms per step:", [round(ev[i].elapsed_time(ev[i + 1]), 4) for i in range(steps)], "loss", float(nll.detach()))
if os.environ.get("PHASES"):
    pm, pc = (ctypes.c_float * 8)(), (ctypes.c_int * 8)()
    L.b200ctc_profile_read(pm, pc, 8)
    names = ["prep", "gather", "lattice", "grad_rows|dense", "prefix", "softmax", "rescale", "fixup"]
    print({n: round(pm[i] / pc[i], 4) for i, n in enumerate(names) if pc[i]})
