#!/usr/bin/env python3
"""Pre-beam prefix step (snum=15, beam 10, T=500, V=5000) for growing batches: where does the
time-parallel kernel (3x the arithmetic, all SMs) stop beating the serial one (one thread per
pair)?  Tuning build; B200CTC_PREFIX_SCAN=0 selects the serial kernel."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from espnet1_b200.espnet.nets.ctc_prefix_score import CTCPrefixScoreTH  # noqa: E402

dev = "cuda:0"
W, V, T, snum = 10, 5000, 500, 15
for B in (16, 32, 48, 64, 96, 128, 192):
    g = torch.Generator().manual_seed(B)
    lp = torch.log_softmax(torch.randn(B, T, V, generator=g), -1).to(dev)
    sc = CTCPrefixScoreTH(lp, torch.full((B,), T), 0, V - 1)
    n = B * W
    y = torch.full((n, 3), V - 1, dtype=torch.long, device=dev)
    ids = torch.stack([torch.randperm(V - 2, generator=g)[:snum] + 1 for _ in range(n)]).to(dev)
    _, st5 = sc(y[:, :1], None, ids)
    best = torch.randint(0, W * V, (B, W), device=dev)
    best = (best // V) * V + ids.view(B, W, snum)[:, :, 0].gather(1, (best // V))  # valid candidates
    st = sc.index_select_state(st5, best)
    for _ in range(3):
        sc(y, st, ids)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        sc(y, st, ids)
    e1.record()
    torch.cuda.synchronize()
    print(f"B={B:4d} pairs={n * snum:6d}  step {e0.elapsed_time(e1) / 20 * 1e3:8.1f} us", flush=True)
    del lp, sc, st5, st
    torch.cuda.empty_cache()
