#!/usr/bin/env python3
"""Debug aid: run b200ctc_loss_fwd_grad through ctypes with a caller-owned workspace and
diff the workspace arrays (em, alpha, beta, offsets) between single-wave and multi-wave runs."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch  # noqa: E402

from espnet1_b200._lib import lib  # noqa: E402
from test_gpu_ctc_random import make_case  # noqa: E402

B, T, V, Lmax = 64, 500, 5000, 150
dev = torch.device("cuda:0")
logits, tg, il, tl = make_case(B, T, V, 75, Lmax, seed=1234)
x = logits.to(dev)
tg, il, tl = tg.to(dev), il.to(dev), tl.to(dev)


def al(v, a):
    return (v + a - 1) // a * a


def layout():
    S = 2 * Lmax + 1
    Lp = 4 + al(max(Lmax, 1), 4)
    Sp = al(S, 8)
    R = 2
    NW = (S + 32 * R - 1) // (32 * R)
    NB = (T + 15) // 16
    o = 0
    out = {}

    def take(name, n):
        nonlocal o
        out[name] = (o, n)
        o = al(o + n, 256)

    take("counter", 4); take("utt", 40 * B); take("lab", 4 * B * Lp); take("dup", 4 * B * Lp)
    take("lse2", 4 * B * T); take("cmax", 4 * B * T); take("offA", 4 * B * NB * NW); take("offB", 4 * B * NB * NW)
    take("em", 4 * B * T * Lp); take("alpha", 4 * B * T * Sp); take("beta", 4 * B * T * Sp)
    return out, o, dict(Lp=Lp, Sp=Sp, NW=NW, NB=NB)


lay, total, dims = layout()
n = lib().b200ctc_loss_workspace_bytes(B, T, V, Lmax)
assert n == total, (n, total)


def run(waves, fill):
    os.environ["B200CTC_WAVES"] = str(waves)
    ws = torch.full((n,), fill, dtype=torch.uint8, device=dev)
    nll = torch.empty(B, device=dev)
    grad = torch.empty_like(x)
    rc = lib().b200ctc_loss_fwd_grad(x.data_ptr(), 0, tg.data_ptr(), None, il.data_ptr(), tl.data_ptr(), B, T, V, Lmax,
                                     0, 1, 1.0 / B, nll.data_ptr(), None, 0.0, grad.data_ptr(), ws.data_ptr(), n,
                                     torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    return ws, nll, grad


def view(ws, name, shape, dt=torch.float32):
    o, nb = lay[name]
    return ws[o:o + nb].view(dt).view(*shape)


ws1, n1, g1 = run(1, 0xFF)
Lp, Sp, NW, NB = dims["Lp"], dims["Sp"], dims["NW"], dims["NB"]
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 40):
    ws, nn, g = run(4, 0xFF if rep % 2 else 0x00)
    if torch.equal(g, g1) and torch.equal(nn, n1):
        continue
    print(f"rep {rep}: MISMATCH")
    for name, shape, dt in [("em", (B, T, Lp), torch.float32), ("alpha", (B, T, Sp), torch.float32),
                            ("beta", (B, T, Sp), torch.float32), ("offA", (B, NB, NW), torch.int32),
                            ("offB", (B, NB, NW), torch.int32), ("lab", (B, Lp), torch.int32),
                            ("cmax", (B, T), torch.float32)]:
        a, b = view(ws1, name, shape, dt), view(ws, name, shape, dt)
        for bb in range(B):
            Tb, L = int(il[bb]), int(tl[bb])
            S = 2 * L + 1
            if name in ("alpha", "beta"):
                d = a[bb, :Tb, :S] != b[bb, :Tb, :S]
            elif name == "em":
                d = torch.cat([a[bb, :Tb, :1] != b[bb, :Tb, :1], a[bb, :Tb, 4:4 + L] != b[bb, :Tb, 4:4 + L]], 1)
            elif name in ("offA", "offB"):
                d = a[bb, :(Tb + 15) // 16, :(S + 63) // 64] != b[bb, :(Tb + 15) // 16, :(S + 63) // 64]
            elif name == "lab":
                d = a[bb, :L + 1] != b[bb, :L + 1]
            else:
                d = a[bb, :Tb] != b[bb, :Tb]
            if d.any():
                idx = d.nonzero()
                print(f"   {name}: utt {bb} (T={Tb}, L={L}) {idx.shape[0]} diffs, first {idx[0].tolist()} last {idx[-1].tolist()}")
                if name in ("alpha", "beta"):
                    t0 = int(idx[0, 0])
                    step = t0 if name == "alpha" else Tb - 1 - t0
                    for tt in ([t0, t0 + 1] if name == "alpha" else [t0, t0 - 1]):
                        if 0 <= tt < Tb:
                            cols = d[tt].nonzero().flatten().tolist()
                            sps = cols if name == "alpha" else [S - 1 - c for c in cols]
                            print(f"      t={tt} step={tt if name == 'alpha' else Tb - 1 - tt} (mod8={(tt if name == 'alpha' else Tb - 1 - tt) % 8}, mod16={(tt if name == 'alpha' else Tb - 1 - tt) % 16}) "
                                  f"recursion positions {sorted(sps)[:12]} good {[round(float(a[bb, tt, c]), 4) for c in cols[:4]]} bad {[round(float(b[bb, tt, c]), 4) for c in cols[:4]]}")
print("done")
