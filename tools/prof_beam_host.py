#!/usr/bin/env python3
"""cProfile of the host side of BatchBeamSearch.forward_batch at C3 widths (32 utterances)."""
import cProfile
import os
import pstats
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from espnet1_b200.espnet.nets.batch_beam_search import BatchBeamSearch  # noqa: E402
from espnet1_b200.espnet.nets.scorers.ctc import CTCPrefixScorer  # noqa: E402
from espnet1_b200.espnet.nets.scorers.length_bonus import LengthBonus  # noqa: E402
from espnet1_b200.espnet2.asr.ctc import CTC  # noqa: E402
from bench_beam import TableScorer  # noqa: E402

dev = "cuda:0"
U, W, V, T, D, POS = 32, 10, 5000, 500, 256, 30
g = torch.Generator().manual_seed(5)
xs = torch.randn(U, T, D, generator=g).to(dev)
table = torch.log_softmax(torch.randn(V, V, generator=g) * 2, dim=1).to(dev)
ctc = CTC(V, D).to(dev)
eos = V - 1
bs = BatchBeamSearch({"ctc": CTCPrefixScorer(ctc, eos), "decoder": TableScorer(table), "length_bonus": LengthBonus(V)},
                     {"ctc": 0.3, "decoder": 0.7, "length_bonus": 0.2}, W, V, eos, eos, pre_beam_score_key="full")
with torch.no_grad():
    bs.forward_batch(xs.clone(), [T] * U, maxlenratio=-POS)
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(3):
        bs.forward_batch(xs.clone(), [T] * U, maxlenratio=-POS)
    torch.cuda.synchronize()
    pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(22)
