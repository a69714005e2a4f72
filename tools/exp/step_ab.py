#!/usr/bin/env python3
"""A/B of one tuning switch on the C-ABI step (tuning build only):
  python tools/exp/step_ab.py B200CTC_FUSED 0 1 [B T V lmin lmax] [ragged]
Runs the eager step under each value in a subprocess, prints ms/step, per-phase times and
whether loss / gradient are bit-identical between the settings."""
import hashlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def child():
    import ctypes

    import torch

    from bench import C2, synth
    from espnet1_b200 import _lib
    from espnet1_b200._lib import check, ptr

    a = sys.argv[3:]
    cfg = dict(C2)
    if len(a) >= 5:
        cfg = dict(B=int(a[0]), T=int(a[1]), V=int(a[2]), lmin=int(a[3]), lmax=int(a[4]))
    ragged = "ragged" in a
    if os.environ.get("B200CTC_LIB"):
        _lib.LIB_PATH = os.path.join(ROOT, "espnet1_b200", os.environ["B200CTC_LIB"])
    L = _lib.lib()
    dev = torch.device("cuda:0")
    B, T, V = cfg["B"], cfg["T"], cfg["V"]
    logits, tg, il, tl = synth(cfg, 1234)
    if ragged:
        g = torch.Generator().manual_seed(5)
        il = torch.randint(T * 3 // 5, T + 1, (B,), generator=g)
        il[0] = T
    x = logits.to(dev)
    tg, il, tl = tg.to(dev), il.to(dev), tl.to(dev)
    n = L.b200ctc_loss_workspace_bytes(B, T, V, cfg["lmax"])
    ws = torch.empty(n, dtype=torch.uint8, device=dev)
    nll = torch.empty(B, device=dev)
    loss = torch.empty((), device=dev)
    grad = torch.empty_like(x)
    one = torch.ones((), device=dev)
    s0 = torch.cuda.current_stream().cuda_stream

    def step():
        check(L.b200ctc_loss_fwd_grad(ptr(x), 0, ptr(tg), None, ptr(il), ptr(tl), B, T, V, cfg["lmax"], 0, 1,
                                      1.0 / B, ptr(nll), ptr(loss), 1.0 / B, ptr(grad), ptr(ws), n, s0), "fwd_grad")
        check(L.b200ctc_grad_rescale(ptr(grad), 0, ptr(one), 0, 1.0 / B, 1.0 / B, B, T, V, s0), "rescale")

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(40):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 40
    L.b200ctc_profile_enable(1)
    for _ in range(10):
        step()
    torch.cuda.synchronize()
    L.b200ctc_profile_enable(0)
    pm, pc = (ctypes.c_float * 8)(), (ctypes.c_int * 8)()
    L.b200ctc_profile_read(pm, pc, 8)
    names = ["prep", "gather", "lattice", "grad_rows", "prefix", "softmax", "rescale", "fixup"]
    ph = {nm: round(pm[i] / pc[i] * 1e3, 1) for i, nm in enumerate(names) if pc[i]}
    h = hashlib.sha1(nll.cpu().numpy().tobytes()).hexdigest()[:12]
    hg = hashlib.sha1(grad.cpu().numpy().tobytes()).hexdigest()[:12]
    print(f"RESULT ms={ms:.4f} phases_us={ph} loss={float(loss):.6f} nll_sha={h} grad_sha={hg} "
          f"nan={int(torch.isnan(grad).sum())}", flush=True)


if __name__ == "__main__":
    if sys.argv[1] == "--child":
        child()
    else:
        # python step_ab.py "K=V,K2=V2" "K=V3" ... [-- B T V lmin lmax [ragged]]
        args = sys.argv[1:]
        rest = []
        if "--" in args:
            i = args.index("--")
            args, rest = args[:i], args[i + 1:]
        for cfgs in args:
            env = dict(os.environ)
            for kv in cfgs.split(","):
                if kv:
                    k, v = kv.split("=")
                    env[k] = v
            r = subprocess.run([sys.executable, __file__, "--child", "x", *rest], env=env, capture_output=True, text=True)
            out = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")]
            print(f"[{cfgs}]", out[0] if out else ("FAILED rc=%d\n" % r.returncode) + r.stdout[-2000:] + r.stderr[-3000:], flush=True)
