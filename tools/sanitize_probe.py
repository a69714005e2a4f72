#!/usr/bin/env python3
"""Small shapes through every kernel family, for `compute-sanitizer --tool memcheck|racecheck|synccheck`
(ONE tool per gpurun call, profiles/README.md "sanitizer"):
split lattice (2 CTAs per lattice, ragged), WIDE row pass, dense-gather switch, fused beam
pruning / pre-beam, register argmax."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from espnet1_b200 import functional as F  # noqa: E402

dev = "cuda:0"
g = torch.Generator().manual_seed(1)
B, T, V = 3, 700, 96
tl = torch.tensor([600, 520, 130])
il = torch.tensor([700, 690, 400])
tg = torch.randint(1, V, (int(tl.sum()),), generator=g)
x = (torch.randn(B, T, V, generator=g) * 2).to(dev).requires_grad_(True)
nll = F.ctc_loss(x, tg.to(dev), il.to(dev), tl.to(dev))
(nll.sum() / B).backward()
with torch.no_grad():
    nll2 = F.ctc_loss(x.detach(), tg.to(dev), il.to(dev), tl.to(dev))
print("loss", nll.tolist(), "fwd-only equal", torch.allclose(nll, nll2, rtol=1e-6))
print("argmax", F.argmax(torch.randn(5, 5000, device=dev)).tolist())
U, W, Vb = 2, 3, 200
mats = [torch.randn(U * W, Vb, device=dev), torch.tensor([1.0], device=dev).expand(U * W, Vb)]
ids = F.beam_prebeam(mats, [0.5, 0.2], 7)
yin = torch.randint(0, Vb, (U * W, 6), device=dev)
yout = torch.zeros(U * W, 8, dtype=torch.int64, device=dev)
out = F.beam_select(mats, [0.5, 0.2], torch.zeros(U * W, device=dev), torch.zeros(2, U * W, device=dev), yin, 4,
                    torch.ones(U, dtype=torch.uint8, device=dev), torch.full((U,), 9, dtype=torch.int32, device=dev),
                    2, U, W, W, Vb - 1, yout)
torch.cuda.synchronize()
print("beam", ids.shape, out[1].tolist())

# ---- round 2: short lattices (sparse gather + the 7-CTA WIDE row pass), per-utterance hints ----
B2, T2, V2 = 4, 60, 512
tl2 = torch.tensor([20, 7, 0, 13])
il2 = torch.tensor([60, 41, 60, 33])
tg2 = torch.randint(1, V2, (int(tl2.sum()),), generator=g)
x2 = (torch.randn(B2, T2, V2, generator=g) * 2).to(dev).requires_grad_(True)
n2 = F.ctc_loss(x2, tg2.to(dev), il2.to(dev), tl2.to(dev), reduce_scale=0.25)
n2.backward()
hints = torch.tensor([0.1, 0.1, 0.2, 0.2], device=dev)
x3 = x2.detach().clone().requires_grad_(True)
n3 = F.ctc_loss(x3, tg2.to(dev), il2.to(dev), tl2.to(dev), grad_hint=hints)
(n3 * hints).sum().backward()
xb = x2.detach().to(torch.bfloat16).requires_grad_(True)
F.ctc_loss(xb, tg2.to(dev), il2.to(dev), tl2.to(dev), reduce_scale=0.25).backward()
print("short lattice loss", float(n2.detach()), "hints grad finite", bool(torch.isfinite(x3.grad).all()))

# ---- two loss calls overlapping on two streams (forward_multi_streams), one batched multi-head call
from espnet1_b200.espnet2.asr.ctc import CTC  # noqa: E402

ctc = CTC(V2, 16).to(dev)
hs = [torch.randn(B2, T2, 16, generator=g).to(dev).requires_grad_(True) for _ in range(2)]
ys = torch.randint(1, V2, (B2, 20), generator=g).to(dev)
la = ctc.forward_multi_streams(hs, il2.to(dev), ys, tl2.to(dev))
lb = ctc.forward_multi(hs, il2.to(dev), ys, tl2.to(dev), upstream_scales=[0.3, 0.7])
(sum(la) + 0.3 * lb[0] + 0.7 * lb[1]).backward()
torch.cuda.synchronize()
print("two-stream heads", [float(v.detach()) for v in la], "batched heads", [float(v.detach()) for v in lb])

# ---- soft-max family + backward, greedy decode, forced alignment (workspace and static scratch)
xs = torch.randn(3, 9, 520, device=dev, requires_grad=True)
F.softmax(xs).square().sum().backward()
xl = torch.randn(3, 9, 520, device=dev, requires_grad=True)
F.log_softmax(xl).sum().backward()
ids_g, lens_g = F.greedy_decode(torch.randn(2, 30, 50, device=dev), torch.tensor([30, 21]))
lp = torch.log_softmax(torch.randn(40, 30, device=dev), -1)
tga = torch.randint(1, 30, (9,), device=dev)
ali = F.forced_align(lp, tga)
from espnet1_b200 import _lib  # noqa: E402

out_s = torch.empty(40, dtype=torch.int64, device=dev)
_lib.check(_lib.lib().b200ctc_forced_align_static(lp.data_ptr(), tga.data_ptr(), 40, 30, 9, 0, out_s.data_ptr(),
                                                  _lib.stream_ptr(lp.device)), "forced_align_static")
print("greedy lens", lens_g.tolist(), "align static == workspace", bool(torch.equal(ali, out_s)))

# ---- prefix scoring: pre-beam + window, full vocabulary, select, streaming append
from espnet1_b200.espnet.nets.ctc_prefix_score import CTCPrefixScoreTH  # noqa: E402
from espnet1_b200.espnet.nets.scorers.ctc import CTCPrefixScorer  # noqa: E402

Bp, Wp, Tp, Vp = 2, 3, 50, 40
lpz = torch.log_softmax(torch.randn(Bp, Tp, Vp, generator=g) * 2, -1).to(dev)
sc = CTCPrefixScoreTH(lpz, torch.tensor([50, 37]), 0, Vp - 1, margin=5)
yy = [[Vp - 1] for _ in range(Bp * Wp)]
part = torch.randint(1, Vp - 1, (Bp * Wp, 6), generator=g).to(dev)
att = torch.softmax(torch.randn(Bp * Wp, Tp, generator=g), -1).to(dev)
s1, st5 = sc(yy, None, part, att)
st4 = sc.index_select_state(st5, torch.topk(s1.view(Bp, -1), Wp, 1)[1])
s2, _ = sc([[Vp - 1, 3] for _ in range(Bp * Wp)], st4, None, att)


class _Stub:
    def log_softmax(self, x):
        return F.log_softmax(x)


ps = CTCPrefixScorer(_Stub(), Vp - 1)
xf = torch.randn(120, Vp, generator=g).to(dev)
ps.batch_init_state(xf[:30].clone())
y1 = torch.full((Wp, 1), Vp - 1, dtype=torch.long, device=dev)
i1 = torch.randint(1, Vp - 1, (Wp, 5), generator=g).to(dev)
sco, stp = ps.batch_score_partial(y1, i1, [None] * Wp, xf[:30])
sts = [ps.select_state(stp, i, int(i1[i, 0])) for i in range(Wp)]
for Tn in (45, 120):
    ps.extend_prob(xf[:Tn].clone())
    sts = ps.extend_state(sts)
    sco, stp = ps.batch_score_partial(torch.cat((y1, i1[:, :1]), 1), i1, sts, xf[:Tn])
print("prefix", tuple(s1.shape), tuple(s2.shape), "stream", tuple(sco.shape))

# ---- device-resident beam search: one library call per position, positions issued ahead
from espnet1_b200.espnet.nets.batch_beam_search import BatchBeamSearch  # noqa: E402
from espnet1_b200.espnet.nets.scorers.length_bonus import LengthBonus  # noqa: E402


class _Tab:
    ignores_x = True

    def __init__(self, t):
        self.t = t

    def init_state(self, x):
        return None

    def batch_score(self, ys, states, xs):
        return self.t[ys[:, -1]], None


Vb2 = 300
tab = torch.log_softmax(torch.randn(Vb2, Vb2, generator=g) * 2, 1).to(dev)
bsr = BatchBeamSearch({"ctc": CTCPrefixScorer(_Stub(), Vb2 - 1), "decoder": _Tab(tab), "length_bonus": LengthBonus(Vb2)},
                      {"ctc": 0.3, "decoder": 0.7, "length_bonus": 0.2}, 5, Vb2, Vb2 - 1, Vb2 - 1,
                      pre_beam_score_key="full")
xb = (torch.randn(3, 80, Vb2, generator=g) * 3).to(dev)
with torch.no_grad():
    nb = bsr.forward_batch(xb, [80, 61, 72], maxlenratio=-12)
print("beam engine", [len(v) for v in nb], nb[0][0].yseq.tolist()[:6], round(nb[0][0].score, 3))

# ---- sharded-loss exchange, one rank, deferred consumption
from espnet1_b200.sharded import PeerLossExchange  # noqa: E402

ex = PeerLossExchange()
for _ in range(3):
    x4 = x2.detach().clone().requires_grad_(True)
    F.ctc_loss(x4, tg2.to(dev), il2.to(dev), tl2.to(dev), reduce_scale=0.25, peer=ex).backward()
print("exchange out", ex.out.tolist(), "timed out", ex.timed_out())
ex.close()
torch.cuda.synchronize()
print("probe done")
