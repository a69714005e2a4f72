#!/usr/bin/env python3
"""Debug aid: the wave-pipelined eager step must be bit-identical to the single-wave step."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import C2, synth  # noqa: E402
from espnet1_b200 import functional as F  # noqa: E402

cfg = dict(C2)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
dev = torch.device("cuda:0")
if os.environ.get("RAGGED", "1") == "1":  # the parity test's generator: ragged input lengths
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
    from test_gpu_ctc_random import make_case  # noqa: E402
    logits, tg, il, tl = make_case(cfg["B"], cfg["T"], cfg["V"], cfg["lmin"], cfg["lmax"], seed=1234)
else:
    logits, tg, il, tl = synth(cfg, 1234)
x = logits.to(dev)


def run(waves):
    os.environ["B200CTC_WAVES"] = str(waves)
    xr = x.clone().requires_grad_(True)
    nll = F.ctc_loss(xr, tg, il, tl, max_tgt_len=cfg["lmax"])
    (nll.sum() / x.shape[0]).backward()
    torch.cuda.synchronize()
    return nll.detach().clone(), xr.grad.clone()


n1, g1 = run(1)
for waves in (1, 2, 4, 8):
    for rep in range(reps):
        n, g = run(waves)
        if not torch.equal(n, n1):
            bad = (n != n1).nonzero().flatten().tolist()
            print(f"waves {waves} rep {rep}: nll differs at utts {bad[:16]}", [(float(n[b]), float(n1[b])) for b in bad[:4]])
        if not torch.equal(g, g1):
            rows = (g != g1).any(-1).nonzero()
            bs = sorted(set(rows[:, 0].tolist()))
            r0 = rows[0].tolist()
            print(f"waves {waves} rep {rep}: grad differs in {rows.shape[0]} rows, utts {bs[:16]}, first (b,t)={r0} "
                  f"T_b={int(il[r0[0]])} maxdiff {float((g - g1).abs().max()):.3e}")
    print("waves", waves, "done")
