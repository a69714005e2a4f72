#!/usr/bin/env python3
"""A/B of tuning switches on the AUTOGRAD step (CTC.loss_fn + backward, what bench.py times):
  python tools/exp/autograd_ab.py "B200CTC_PDL=7" "B200CTC_PDL=31" ..."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
if sys.argv[1] == "--child":
    import torch

    from bench import C2, synth
    from espnet1_b200.espnet2.asr.ctc import CTC

    if os.environ.get("B200CTC_LIB"):
        from espnet1_b200 import _lib

        _lib.LIB_PATH = os.path.join(ROOT, "espnet1_b200", os.environ["B200CTC_LIB"])
    dev = torch.device("cuda:0")
    logits, tg, il, tl = synth(C2, 1234)
    x = logits.to(dev).requires_grad_(True)
    tg, il, tl = tg.to(dev), il.to(dev), tl.to(dev)
    m = CTC(C2["V"], 8).to(dev)

    def step():
        x.grad = None
        loss = m.loss_fn(x.transpose(0, 1), tg, il, tl, max_tgt_len=C2["lmax"])
        loss.backward()

    for _ in range(8):
        step()
    res = []
    for _ in range(3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            step()
        e1.record()
        torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) / 50)
    print("RESULT ms", [round(v, 4) for v in res], flush=True)
else:
    for cfg in sys.argv[1:]:
        env = dict(os.environ)
        for kv in cfg.split(","):
            if kv:
                k, v = kv.split("=")
                env[k] = v
        r = subprocess.run([sys.executable, __file__, "--child"], env=env, capture_output=True, text=True)
        out = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")]
        print(f"[{cfg}]", out[0] if out else "FAILED\n" + r.stderr[-1500:], flush=True)
