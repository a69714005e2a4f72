#!/usr/bin/env python3
"""A few prefix steps of one C3 variant (for ncu launch lists): prof_prefix_one.py snum margin steps"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from espnet1_b200.espnet.nets.ctc_prefix_score import CTCPrefixScoreTH

snum = int(sys.argv[1]); margin = int(sys.argv[2]); n = int(sys.argv[3])
dev = torch.device("cuda:0")
B, W, V, T = 32, 10, 5000, 500
g = torch.Generator().manual_seed(1234)
lpz = torch.log_softmax(torch.randn(B, T, V, generator=g) * 2, dim=2)
xl = torch.randint(250, T + 1, (B,), generator=g); xl[0] = T
sc = CTCPrefixScoreTH(lpz.to(dev), xl, 0, V - 1, margin=margin)
tt = torch.arange(T, dtype=torch.float32)
state = None
for i in range(n):
    gi = torch.Generator().manual_seed(77 + i)
    local = torch.randn(B * W, V, generator=gi)
    part = torch.topk(local, snum, dim=-1)[1].to(dev) if snum else None
    att = None
    if margin:
        c = T * (i + 1) / 60.0 + (torch.rand(B * W, generator=gi) * 10 - 5)
        att = torch.softmax(-((tt[None] - c[:, None]) ** 2) / 128.0, dim=1).to(dev)
    best = torch.topk(local.view(B, -1), W, 1)[1].to(dev)
    y = torch.full((B * W, i + 1), V - 1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    _, st5 = sc(y, state, part, att)
    state = sc.index_select_state(st5, best)
torch.cuda.synchronize()
print("done")
