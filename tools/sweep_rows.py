#!/usr/bin/env python3
"""Tuning sweep: per-kernel CUDA-event times of the C2 step for several
(CTAs/SM, stages) settings of the two row passes.  python tools/sweep_rows.py"""
import ctypes
import itertools
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import C2, synth  # noqa: E402
from espnet1_b200 import _lib, functional as F  # noqa: E402

cfg = dict(C2)
if len(sys.argv) > 5:
    cfg = dict(B=int(sys.argv[1]), T=int(sys.argv[2]), V=int(sys.argv[3]), lmin=int(sys.argv[4]), lmax=int(sys.argv[5]))
dev = torch.device("cuda:0")
logits, tg, il, tl = synth(cfg, 1234)
x = logits.to(dev).requires_grad_(True)
tg, il, tl = tg.to(dev), il.to(dev), tl.to(dev)
L = _lib.lib()


def run(n=8):
    for _ in range(3):
        x.grad = None
        (F.ctc_loss(x, tg, il, tl, max_tgt_len=cfg["lmax"]).sum() / cfg["B"]).backward()
    torch.cuda.synchronize()
    L.b200ctc_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        x.grad = None
        (F.ctc_loss(x, tg, il, tl, max_tgt_len=cfg["lmax"]).sum() / cfg["B"]).backward()
    e1.record()
    torch.cuda.synchronize()
    L.b200ctc_profile_enable(0)
    pm, pc = (ctypes.c_float * 8)(), (ctypes.c_int * 8)()
    L.b200ctc_profile_read(pm, pc, 8)
    ph = [pm[i] / max(pc[i], 1) * 1e3 for i in range(4)]
    return e0.elapsed_time(e1) / n * 1e3, ph


combos = [(5, 2), (4, 2), (3, 3), (3, 2), (2, 4), (2, 3), (6, 1), (8, 1), (5, 1)]
print("cfg", cfg)
print("ctas stages | step_us | prep A lattice B (us)")
for ctas, st in combos:
    os.environ["B200CTC_A_CTAS"] = os.environ["B200CTC_B_CTAS"] = str(ctas)
    os.environ["B200CTC_A_STAGES"] = os.environ["B200CTC_B_STAGES"] = str(st)
    tot, ph = run()
    print(f"{ctas:4d} {st:6d} | {tot:8.1f} | " + " ".join(f"{p:7.1f}" for p in ph), flush=True)
