#!/usr/bin/env python3
"""Debug aid: per-utterance nll of the CUDA path vs torch's CUDA ctc_loss, repeated (race hunting)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import C2, synth  # noqa: E402
from espnet1_b200 import functional as F  # noqa: E402

cfg = dict(C2)
if len(sys.argv) > 5:
    cfg = dict(B=int(sys.argv[1]), T=int(sys.argv[2]), V=int(sys.argv[3]), lmin=int(sys.argv[4]), lmax=int(sys.argv[5]))
dev = torch.device("cuda:0")
logits, tg, il, tl = synth(cfg, 1234)
x = logits.to(dev)
lp = x.transpose(0, 1).log_softmax(2)
ref = torch.nn.functional.ctc_loss(lp, tg.to(dev), il.to(dev), tl.to(dev), reduction="none", zero_infinity=True)
for mode in ("eager", "deferred", "nograd"):
    for rep in range(12):
        if mode == "nograd":
            with torch.no_grad():
                nll = F.ctc_loss(x, tg, il, tl, max_tgt_len=cfg["lmax"])
        else:
            xr = x.clone().requires_grad_(True)
            nll = F.ctc_loss(xr, tg, il, tl, max_tgt_len=cfg["lmax"], grad_mode=mode)
        d = (nll.detach() - ref).abs() / ref.abs()
        bad = (d > 1e-5).nonzero().flatten().tolist()
        if bad:
            print(f"{mode} rep {rep}: max rel {float(d.max()):.3e} bad utts {bad[:16]} L {[int(tl[b]) for b in bad[:16]]}",
                  [round(float(nll[b].detach() - ref[b]) / 0.6931, 2) for b in bad[:8]])
    print(mode, "done")
