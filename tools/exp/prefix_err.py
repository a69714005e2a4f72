#!/usr/bin/env python3
"""Error of the CUDA prefix scores against the same recursion in fp64, next to the error of the
reference's fp32 arithmetic (oracle), for long inputs.  Tuning build: B200CTC_PREFIX_SCAN=0
selects the serial kernel."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_prefix import _drive, _th  # noqa: E402
from oracle.prefix_torch import PrefixOracle  # noqa: E402

for (B, W, V, T, snum, margin) in ((3, 4, 200, 333, 16, 0), (2, 3, 120, 640, 7, 0), (4, 5, 300, 500, 15, 0)):
    g = torch.Generator().manual_seed(1000 + V + snum + margin)
    lpz = torch.log_softmax(torch.randn(B, T, V, generator=g) * 2, dim=2)
    xlens = torch.randint(T // 2, T + 1, (B,), generator=g)
    xlens[0] = T
    eos = V - 1
    ora = PrefixOracle(lpz.clone(), xlens, 0, eos, margin)
    ora64 = PrefixOracle(lpz.double().clone(), xlens, 0, eos, margin)
    cu = _th()(lpz.clone().to("cuda:0"), xlens, 0, eos, margin)
    ref, _ = _drive(lambda *a: ora.score(*a), ora.select, B, W, V, T, snum, margin, 6, 7, "cpu")
    tru, _ = _drive(lambda *a: ora64.score(*a), ora64.select, B, W, V, T, snum, margin, 6, 7, "cpu")
    got, _ = _drive(cu, cu.index_select_state, B, W, V, T, snum, margin, 6, 7, "cuda:0")
    for i, ((s0, p0, _, _), (s1, p1, _, _), (s2, p2, _, _)) in enumerate(zip(ref, got, tru)):
        live = p2 > -1e9
        e_ref = (p0.double() - p2)[live].abs().max().item()
        e_cu = (p1.double() - p2)[live].abs().max().item()
        print(f"T={T} snum={snum} step {i}: log_psi err ref32 {e_ref:.2e}  cuda {e_cu:.2e}   |psi| ~ {p2[live].abs().median().item():.0f}")
