#!/usr/bin/env python3
"""In-situ (warm-cache) kernel times of one batch-32 decode via torch.profiler (CUPTI), and the
marginal cost of a position: (t(POS=60) - t(POS=30)) / 30."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from espnet1_b200.espnet.nets.batch_beam_search import BatchBeamSearch  # noqa: E402
from espnet1_b200.espnet.nets.scorers.ctc import CTCPrefixScorer  # noqa: E402
from espnet1_b200.espnet.nets.scorers.length_bonus import LengthBonus  # noqa: E402
from espnet1_b200.espnet2.asr.ctc import CTC  # noqa: E402
from bench_beam import TableScorer  # noqa: E402

dev = "cuda:0"
U, W, V, T, D = 32, 10, 5000, 500, 256
g = torch.Generator().manual_seed(5)
xs = torch.randn(U, T, D, generator=g).to(dev)
table = torch.log_softmax(torch.randn(V, V, generator=g) * 2, dim=1).to(dev)
ctc = CTC(V, D).to(dev)
eos = V - 1
bs = BatchBeamSearch({"ctc": CTCPrefixScorer(ctc, eos), "decoder": TableScorer(table), "length_bonus": LengthBonus(V)},
                     {"ctc": 0.3, "decoder": 0.7, "length_bonus": 0.2}, W, V, eos, eos, pre_beam_score_key="full")


def run(pos, reps=5):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        bs.forward_batch(xs, [T] * U, maxlenratio=-pos)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best


with torch.no_grad():
    run(30, 2)
    t30, t60, t120 = run(30), run(60), run(120)
    print(f"decode 30 pos {1e3 * t30:.3f} ms, 60 pos {1e3 * t60:.3f} ms, 120 pos {1e3 * t120:.3f} ms; marginal "
          f"{1e3 * (t60 - t30) / 30:.4f} / {1e3 * (t120 - t60) / 60:.4f} ms per position")
    from torch.profiler import ProfilerActivity, profile

    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        bs.forward_batch(xs, [T] * U, maxlenratio=-30)
        torch.cuda.synchronize()
    rows = [(e.key, e.device_time_total, e.count) for e in prof.key_averages() if e.device_time_total > 0]
    rows.sort(key=lambda r: -r[1])
    tot = sum(r[1] for r in rows)
    print(f"GPU busy {tot / 1e3:.3f} ms over one 30-position decode")
    for k, t, c in rows[:24]:
        print(f"{t / 1e3:9.3f} ms  x{c:4d}  {t / max(c, 1):8.1f} us  {k[:90]}")
