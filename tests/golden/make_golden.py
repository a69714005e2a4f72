#!/usr/bin/env python3
"""Generate golden fixtures by running the REAL reference (read-only tree).

Run in the build container only:  ``python tests/golden/make_golden.py``
(needs /root/reference; the GPU box never runs this).  The reference has no
golden vectors of its own for the CTC path, so these fixtures -
outputs of the unmodified reference modules on seeded inputs - are what pins
the oracle and the CUDA path.  Everything is small (a few hundred KB total).

Reference modules exercised (relative to /root/reference):
  espnet2/asr/ctc.py                     CTC.forward/softmax/log_softmax/argmax
  espnet/nets/pytorch_backend/ctc.py     legacy CTC.forward, forced_align
  espnet/nets/ctc_prefix_score.py        CTCPrefixScoreTH, CTCPrefixScore
  espnet/nets/scorers/ctc.py             CTCPrefixScorer
  espnet/nets/beam_search.py, batch_beam_search.py   (callers, for n-best lists)
"""
import json
import os
import sys

import numpy as np
import torch

REF = os.environ.get("ESPNET_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
HERE = os.path.dirname(os.path.abspath(__file__))

import espnet2.asr.ctc as ref_ctc2  # noqa: E402

ref_ctc2.check_argument_types = lambda: True  # typeguard 4.x vs `bool = None`
from espnet.nets.batch_beam_search import BatchBeamSearch  # noqa: E402
from espnet.nets.beam_search import BeamSearch  # noqa: E402
from espnet.nets.ctc_prefix_score import CTCPrefixScore, CTCPrefixScoreTH  # noqa: E402
from espnet.nets.pytorch_backend.ctc import CTC as RefCTC1  # noqa: E402
from espnet.nets.scorer_interface import BatchScorerInterface  # noqa: E402
from espnet.nets.scorers.ctc import CTCPrefixScorer  # noqa: E402
from espnet.nets.scorers.length_bonus import LengthBonus  # noqa: E402

LOGZERO = -10000000000.0


def save(name, d):
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in d.items()})
    print("wrote", path, os.path.getsize(path), "bytes")


# --------------------------------------------------------------------------
def make_ctc_module():
    """espnet2 + espnet1 CTC modules on one small ragged batch with: repeated
    labels, a short utterance, an empty target, an infeasible utterance."""
    torch.manual_seed(7)
    B, T, D, V = 5, 30, 8, 12
    hs = torch.randn(B, T, D)
    hlens = torch.tensor([30, 24, 30, 12, 30])
    ys = [[3, 3, 5, 1, 1, 1, 9], [2, 4, 6, 8, 10, 11], [], [7] * 9, [1, 2, 3, 4]]
    # utt 3: 9 repeats of one label need 17 frames > 12 -> infeasible (inf -> 0)
    Lmax = max(len(y) for y in ys)
    ys_pad = torch.full((B, Lmax), -1, dtype=torch.long)
    for i, y in enumerate(ys):
        ys_pad[i, : len(y)] = torch.tensor(y, dtype=torch.long)
    ys_lens = torch.tensor([len(y) for y in ys])
    out = dict(hs=hs.numpy(), hlens=hlens.numpy(), ys_pad=ys_pad.numpy(), ys_lens=ys_lens.numpy())
    m = ref_ctc2.CTC(V, D)
    out["weight"] = m.ctc_lo.weight.detach().numpy()
    out["bias"] = m.ctc_lo.bias.detach().numpy()
    for reduce in (True, False):
        m.reduce = reduce
        m.zero_grad()
        x = hs.clone().requires_grad_(True)
        loss = m(x, hlens, ys_pad, ys_lens)
        (loss if reduce else (loss * torch.arange(1, B + 1)).sum()).backward()
        tag = "reduce" if reduce else "noreduce"
        out[f"loss_{tag}"] = loss.detach().numpy()
        out[f"ghs_{tag}"] = x.grad.numpy()
        out[f"gw_{tag}"] = m.ctc_lo.weight.grad.numpy().copy()
        out[f"gb_{tag}"] = m.ctc_lo.bias.grad.numpy().copy()
    # gradient w.r.t. the logits themselves (what the CUDA kernel emits)
    logits = m.ctc_lo(hs).detach().requires_grad_(True)
    m.reduce = True
    ys_true = torch.cat([ys_pad[i, :l] for i, l in enumerate(ys_lens)])
    m.loss_fn(logits.transpose(0, 1), ys_true, hlens, ys_lens).backward()
    out["logits"] = logits.detach().numpy()
    out["glogits_reduce"] = logits.grad.numpy()
    # zero_infinity=False variant
    m_inf = ref_ctc2.CTC(V, D, zero_infinity=False)
    m_inf.load_state_dict(m.state_dict())
    m_inf.reduce = False
    out["loss_noreduce_keepinf"] = m_inf(hs, hlens, ys_pad, ys_lens).detach().numpy()
    with torch.no_grad():
        out["softmax"] = m.softmax(hs).numpy()
        out["log_softmax"] = m.log_softmax(hs).numpy()
        out["argmax"] = m.argmax(hs).numpy()
    # legacy module, same weights
    for reduce in (True, False):
        m1 = RefCTC1(V, D, 0.0, reduce=reduce)
        m1.load_state_dict(m.state_dict(), strict=False)
        x = hs.clone().requires_grad_(True)
        loss = m1(x, hlens, ys_pad)
        loss.sum().backward()
        tag = "reduce" if reduce else "noreduce"
        out[f"legacy_loss_{tag}"] = loss.detach().numpy()
        out[f"legacy_ghs_{tag}"] = x.grad.numpy()
    # forced alignment of utt 0 (legacy API, ctc.py:163)
    m1 = RefCTC1(V, D, 0.0)
    m1.load_state_dict(m.state_dict(), strict=False)
    with torch.no_grad():
        ali = m1.forced_align(hs[0:1], ys[0])
    out["forced_align_utt0"] = np.asarray(ali, dtype=np.int64)
    save("ctc_module.npz", out)


def make_ctc_c1():
    """Config C1 at the logits boundary: B=8, T=200, V=30, L in [25,50]."""
    g = torch.Generator().manual_seed(1234)
    B, T, V = 8, 200, 30
    logits = torch.randn(B, T, V, generator=g)
    tlens = torch.randint(25, 51, (B,), generator=g)
    ilens = torch.randint(120, 201, (B,), generator=g)
    ilens[0] = T
    tgts = [torch.randint(1, V, (int(l),), generator=g) for l in tlens]
    x = logits.clone().requires_grad_(True)
    lp = x.transpose(0, 1).log_softmax(2)
    nll = torch.nn.CTCLoss(reduction="none", zero_infinity=True)(lp, torch.cat(tgts), ilens, tlens)
    (nll.sum() / B).backward()
    save("ctc_c1.npz", dict(logits=logits.numpy(), targets=torch.cat(tgts).numpy(),
                            tgt_lens=tlens.numpy(), in_lens=ilens.numpy(),
                            nll=nll.detach().numpy(), grad=x.grad.numpy()))


# --------------------------------------------------------------------------
def synth_att(n_bh, T, step, g):
    c = T * (step + 1) / 8.0 + (torch.rand(n_bh, generator=g) * 4 - 2)
    t = torch.arange(T, dtype=torch.float32)
    return torch.softmax(-((t[None] - c[:, None]) ** 2) / (2 * 3.0 ** 2), dim=1)


def prefix_trace(tag, B, T, V, W, snum, margin, xlens, steps, seed, out):
    """Drive CTCPrefixScoreTH the way recognize_beam_batch does
    (rnn/decoders.py:811-828,841,935-937) and record every tensor."""
    g = torch.Generator().manual_seed(seed)
    eos = V - 1
    lpz = torch.log_softmax(torch.randn(B, T, V, generator=g) * 2, dim=2)
    out[f"{tag}__lpz"] = lpz.numpy().copy()
    out[f"{tag}__xlens"] = np.asarray(xlens)
    out[f"{tag}__cfg"] = np.asarray([B, T, V, W, snum or 0, margin, eos, steps])
    scorer = CTCPrefixScoreTH(lpz, torch.tensor(xlens), 0, eos, margin=margin)
    out[f"{tag}__lpz_padded"] = lpz.numpy().copy()  # constructor pads in place
    n_bh = B * W
    yseq = [[eos] for _ in range(n_bh)]  # sos == eos in ESPnet
    vscores = torch.zeros(B, W)
    state = None
    pad_b = (torch.arange(B) * W).view(-1, 1)
    for i in range(steps):
        local = torch.randn(n_bh, V, generator=g)
        local[:, 0] = LOGZERO
        part_ids = torch.topk(local, snum, dim=-1)[1] if snum else None
        att_w = synth_att(n_bh, T, i, g) if margin > 0 else None
        sc, state = scorer(yseq, state, part_ids, att_w)
        r, log_psi, f_min, f_max, idmap = state
        p = f"{tag}__s{i}__"
        out[p + "y"] = np.asarray(yseq)
        out[p + "local"] = local.numpy().copy()
        if part_ids is not None:
            out[p + "part_ids"] = part_ids.numpy()
            out[p + "idmap"] = idmap.numpy()
        if att_w is not None:
            out[p + "att_w"] = att_w.numpy()
        out[p + "scores"] = sc.numpy().copy()
        out[p + "r"] = r.numpy().copy()
        out[p + "log_psi"] = log_psi.numpy().copy()
        out[p + "fminmax"] = np.asarray([f_min, f_max])
        tot = (local + 0.5 * sc).view(B, W, V)
        if i == 0:
            tot[:, 1:, :] = LOGZERO
        vs = vscores.view(B, W, 1).repeat(1, 1, V)
        vs[:, :, eos] = LOGZERO
        vs = (vs + tot).view(B, -1)
        vscores, best = torch.topk(vs, W, 1)
        out[p + "best_ids"] = best.numpy()
        tok = torch.fmod(best, V).view(-1).tolist()
        src = (best // V + pad_b).view(-1).tolist()
        yseq = [yseq[s][:] + [t] for s, t in zip(src, tok)]
        state = scorer.index_select_state(state, best)
        out[p + "sel_r"] = state[0].numpy().copy()
        out[p + "sel_s"] = state[1].numpy().copy()


def make_prefix():
    out = {}
    prefix_trace("pre_m0", 3, 40, 12, 4, 6, 0, [40, 33, 25], 6, 11, out)
    prefix_trace("pre_m5", 3, 40, 12, 4, 6, 5, [40, 33, 25], 6, 12, out)
    prefix_trace("full_m0", 2, 32, 10, 3, None, 0, [32, 20], 5, 13, out)
    prefix_trace("full_m5", 2, 32, 10, 3, None, 5, [32, 20], 5, 14, out)
    prefix_trace("b1_w3", 1, 50, 9, 3, 4, 0, [50], 8, 15, out)
    save("prefix_th.npz", out)
    # numpy single-hypothesis scorer: a short greedy walk
    g = torch.Generator().manual_seed(21)
    T, V = 25, 8
    x = torch.log_softmax(torch.randn(T, V, generator=g) * 2, dim=1).numpy()
    sc = CTCPrefixScore(x, 0, V - 1, np)
    r = sc.initial_state()
    o = {"x": x, "r0": r.copy()}
    y = [V - 1]
    for i in range(5):
        cs = np.array(sorted(torch.randperm(V, generator=g)[:5].tolist()))
        psi, rs = sc(y, cs, r)
        o[f"s{i}__y"] = np.asarray(y)
        o[f"s{i}__cs"] = cs
        o[f"s{i}__psi"] = psi.copy()
        valid_from = max(len(y) - 1, 1) - 1
        o[f"s{i}__r"] = rs[:, valid_from:].copy()
        o[f"s{i}__valid_from"] = np.asarray(valid_from)
        k = int(np.argmax(np.where((cs == 0) | (cs == V - 1), -np.inf, psi)))
        y = y + [int(cs[k])]
        r = rs[k]
    save("prefix_np.npz", o)


# --------------------------------------------------------------------------
class _StubCTC(torch.nn.Module):
    """`ctc` object handed to CTCPrefixScorer: x is already (T, V) logits."""

    def log_softmax(self, x):
        return torch.log_softmax(x, dim=-1)


class _Table(BatchScorerInterface):
    def __init__(self, table):
        self.table = table

    def score(self, y, state, x):
        return self.table[int(y[-1])].to(dtype=x.dtype), None

    def batch_score(self, ys, states, xs):
        return self.table.to(dtype=xs.dtype)[ys[:, -1]], None


def make_beam():
    res = {}
    arrays = {}
    cases = [
        # name, T, V, beam, weights, pre_beam_key, seed
        ("ctc_only_v5", 20, 5, 3, {"ctc": 1.0, "length_bonus": 0.1}, None, 123),
        ("joint_v5", 20, 5, 3, {"ctc": 0.5, "decoder": 0.5, "length_bonus": 0.1}, "full", 124),
        ("joint_v40", 40, 40, 5, {"ctc": 0.3, "decoder": 0.7, "length_bonus": 0.2}, "full", 125),
        ("ctc_only_v40", 40, 40, 4, {"ctc": 1.0, "length_bonus": 0.3}, None, 126),
    ]
    for name, T, V, beam, weights, key, seed in cases:
        g = torch.Generator().manual_seed(seed)
        x = torch.randn(T, V, generator=g) * 3
        table = torch.log_softmax(torch.randn(V, V, generator=g) * 2, dim=1)
        arrays[f"{name}__x"] = x.numpy()
        arrays[f"{name}__table"] = table.numpy()
        eos = V - 1
        entry = {"T": T, "V": V, "beam_size": beam, "weights": weights, "pre_beam_key": key,
                 "sos": eos, "eos": eos}
        for cls, tag in ((BeamSearch, "beam"), (BatchBeamSearch, "batch")):
            scorers = {"ctc": CTCPrefixScorer(_StubCTC(), eos), "length_bonus": LengthBonus(V)}
            if "decoder" in weights:
                scorers["decoder"] = _Table(table)
            bs = cls(scorers=scorers, weights=weights, beam_size=beam, vocab_size=V,
                     sos=eos, eos=eos, pre_beam_score_key=key)
            with torch.no_grad():
                nbest = bs(x=x.clone(), maxlenratio=0.0, minlenratio=0.0)
            entry[tag] = [{"yseq": [int(v) for v in h.yseq], "score": float(h.score),
                           "scores": {k: float(v) for k, v in h.scores.items()}} for h in nbest]
        res[name] = entry
    save("beam_inputs.npz", arrays)
    with open(os.path.join(HERE, "beam_nbest.json"), "w") as f:
        json.dump(res, f, indent=1)
    print("wrote beam_nbest.json")


def make_beam_multi():
    """n-best lists of the REAL BatchBeamSearch, one utterance at a time, for
    a ragged batch of utterances; the device-resident engine must reproduce them both one
    by one and decoded together (forward_batch)."""
    res, arrays = {}, {}
    cases = [
        # name, lengths, V, beam, weights, pre_beam_key, seed
        ("multi_joint_v60", [36, 29, 22, 36], 60, 4, {"ctc": 0.4, "decoder": 0.6, "length_bonus": 0.3},
         "full", 211),
        ("multi_ctc_v30", [24, 17, 24], 30, 3, {"ctc": 1.0, "length_bonus": 0.5}, None, 212),
        ("multi_joint_v300", [50, 41], 300, 6, {"ctc": 0.3, "decoder": 0.7, "length_bonus": 0.4},
         "decoder", 213),
    ]
    for name, lens, V, beam, weights, key, seed in cases:
        g = torch.Generator().manual_seed(seed)
        Tm = max(lens)
        xs = torch.randn(len(lens), Tm, V, generator=g) * 3
        table = torch.log_softmax(torch.randn(V, V, generator=g) * 2, dim=1)
        arrays[f"{name}__xs"] = xs.numpy()
        arrays[f"{name}__table"] = table.numpy()
        eos = V - 1
        entry = {"lens": lens, "V": V, "beam_size": beam, "weights": weights, "pre_beam_key": key,
                 "sos": eos, "eos": eos, "nbest": []}
        for u, T in enumerate(lens):
            scorers = {"ctc": CTCPrefixScorer(_StubCTC(), eos), "length_bonus": LengthBonus(V)}
            if "decoder" in weights:
                scorers["decoder"] = _Table(table)
            bs = BatchBeamSearch(scorers=scorers, weights=weights, beam_size=beam, vocab_size=V,
                                 sos=eos, eos=eos, pre_beam_score_key=key)
            with torch.no_grad():
                nbest = bs(x=xs[u, :T].clone(), maxlenratio=0.0, minlenratio=0.0)
            entry["nbest"].append([{"yseq": [int(v) for v in h.yseq], "score": float(h.score),
                                    "scores": {k: float(v) for k, v in h.scores.items()}}
                                   for h in nbest])
        res[name] = entry
    save("beam_multi_inputs.npz", arrays)
    with open(os.path.join(HERE, "beam_multi_nbest.json"), "w") as f:
        json.dump(res, f, indent=1)
    print("wrote beam_multi_nbest.json")


def make_beam_opts():
    """BatchBeamSearch options: fixed / ratio maximum lengths and a multi-token primer
    (beam_search.py:346-365, batch_beam_search.py:131-144), single utterance each."""
    res, arrays = {}, {}
    cases = [
        # name, T, V, beam, weights, pre_beam_key, seed, maxlenratio, minlenratio, primer
        ("opt_maxlen_fixed", 30, 25, 4, {"ctc": 0.5, "decoder": 0.5, "length_bonus": 0.2}, "full", 301, -7.0, 0.0, None),
        ("opt_maxlen_ratio", 30, 25, 3, {"ctc": 0.7, "decoder": 0.3, "length_bonus": 0.1}, "decoder", 302, 0.4, 0.0, None),
        ("opt_primer", 26, 20, 3, {"ctc": 0.4, "decoder": 0.6, "length_bonus": 0.2}, "full", 303, 0.0, 0.0, [19, 3, 5]),
        ("opt_beam_wide", 22, 12, 8, {"ctc": 1.0, "length_bonus": 0.4}, None, 304, 0.0, 0.0, None),
    ]
    for name, T, V, beam, weights, key, seed, mlr, minr, primer in cases:
        g = torch.Generator().manual_seed(seed)
        x = torch.randn(T, V, generator=g) * 3
        table = torch.log_softmax(torch.randn(V, V, generator=g) * 2, dim=1)
        arrays[f"{name}__x"] = x.numpy()
        arrays[f"{name}__table"] = table.numpy()
        eos = V - 1
        scorers = {"ctc": CTCPrefixScorer(_StubCTC(), eos), "length_bonus": LengthBonus(V)}
        if "decoder" in weights:
            scorers["decoder"] = _Table(table)
        bs = BatchBeamSearch(scorers=scorers, weights=weights, beam_size=beam, vocab_size=V,
                             sos=eos, eos=eos, pre_beam_score_key=key, hyp_primer=primer)
        with torch.no_grad():
            nbest = bs(x=x.clone(), maxlenratio=mlr, minlenratio=minr)
        res[name] = {"T": T, "V": V, "beam_size": beam, "weights": weights, "pre_beam_key": key,
                     "sos": eos, "eos": eos, "maxlenratio": mlr, "minlenratio": minr, "primer": primer,
                     "batch": [{"yseq": [int(v) for v in h.yseq], "score": float(h.score),
                                "scores": {k: float(v) for k, v in h.scores.items()}} for h in nbest]}
    save("beam_opts_inputs.npz", arrays)
    with open(os.path.join(HERE, "beam_opts_nbest.json"), "w") as f:
        json.dump(res, f, indent=1)
    print("wrote beam_opts_nbest.json")


def c3_beam_inputs(seed=401, T=500, V=5000):
    """Inputs of the C3-width beam-search case, regenerated from the seed wherever they are needed
    (a 5000 x 5000 bigram table is 100 MB: not a fixture); the fixture pins their checksums."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(T, V, generator=g) * 3
    table = torch.log_softmax(torch.randn(V, V, generator=g) * 2, dim=1)
    return x, table


def make_beam_c3():
    """beam-search identity was only pinned for V <= 300.  The REAL reference
    BatchBeamSearch at C3 widths: V=5000, T=500, beam 10, pre-beam 15 on the full scores, CTC 0.3 +
    bigram decoder 0.7 + length bonus 0.2, 20 positions (maxlenratio = -20)."""
    T, V, beam, seed, mlr = 500, 5000, 10, 401, -20.0
    weights = {"ctc": 0.3, "decoder": 0.7, "length_bonus": 0.2}
    x, table = c3_beam_inputs(seed, T, V)
    eos = V - 1
    scorers = {"ctc": CTCPrefixScorer(_StubCTC(), eos), "length_bonus": LengthBonus(V), "decoder": _Table(table)}
    bs = BatchBeamSearch(scorers=scorers, weights=weights, beam_size=beam, vocab_size=V, sos=eos, eos=eos,
                         pre_beam_score_key="full")
    with torch.no_grad():
        nbest = bs(x=x.clone(), maxlenratio=mlr, minlenratio=0.0)
    res = {"T": T, "V": V, "beam_size": beam, "weights": weights, "pre_beam_key": "full", "sos": eos, "eos": eos,
           "seed": seed, "maxlenratio": mlr,
           "x_sum": float(x.double().sum()), "table_sum": float(table.double().sum()),
           "batch": [{"yseq": [int(v) for v in h.yseq], "score": float(h.score),
                      "scores": {k: float(v) for k, v in h.scores.items()}} for h in nbest]}
    with open(os.path.join(HERE, "beam_c3_nbest.json"), "w") as f:
        json.dump(res, f, indent=1)
    print("wrote beam_c3_nbest.json:", len(nbest), "hypotheses, best", res["batch"][0]["score"])


def make_streaming():
    """extend_prob / extend_state (ctc_prefix_score.py:222-269) through the
    CTCPrefixScorer adapter the way batch_beam_search_online.py:306-309 does."""
    g = torch.Generator().manual_seed(31)
    T1, T2, V, W, snum = 12, 20, 7, 2, 3
    eos = V - 1
    x_full = torch.randn(T2, V, generator=g) * 2
    sc = CTCPrefixScorer(_StubCTC(), eos)
    sc.batch_init_state(x_full[:T1].clone())
    o = {"x_full": x_full.numpy(), "cfg": np.asarray([T1, T2, V, W, snum, eos])}
    y = torch.full((W, 1), eos, dtype=torch.long)
    ids = torch.stack([torch.randperm(V, generator=g)[:snum] for _ in range(W)])
    scores, st = sc.batch_score_partial(y, ids, [None] * W, x_full[:T1])
    o["ids0"] = ids.numpy()
    o["scores0"] = scores.numpy().copy()
    states = [sc.select_state(st, i, int(ids[i, 0])) for i in range(W)]
    toks = [int(ids[i, 0]) for i in range(W)]
    sc.extend_prob(x_full.clone())
    states = sc.extend_state(states)
    o["ext_r"] = torch.stack([s[0] for s in states]).numpy()
    y = torch.cat((y, torch.tensor(toks).view(-1, 1)), 1)
    ids = torch.stack([torch.randperm(V, generator=g)[:snum] for _ in range(W)])
    scores, st = sc.batch_score_partial(y, ids, states, x_full)
    o["toks"] = np.asarray(toks)
    o["ids1"] = ids.numpy()
    o["scores1"] = scores.numpy().copy()
    o["r1"] = st[0].numpy().copy()
    save("prefix_stream.npz", o)


def make_align_greedy():
    """(1) forced alignment (espnet/nets/pytorch_backend/ctc.py:163-226) on several shapes - short
    and long targets, repeated labels, a target nearly as long as the input; (2) the greedy CTC
    collapse of ``ErrorCalculator.calculate_cer_ctc`` (espnet/nets/e2e_asr_common.py:159-191):
    the reference joins the surviving characters and hands them to ``editdistance.eval``.  That
    package is not installed here, so a stand-in module records the strings the REFERENCE's own
    collapse code produced (one unique character per id: the strings map back to ids exactly)."""
    import types

    from espnet.nets.e2e_asr_common import ErrorCalculator

    out = {}
    torch.manual_seed(31)
    V, D = 14, 6
    m1 = RefCTC1(V, D, 0.0)
    cases = [(40, [3, 3, 5, 1, 1]), (80, list(np.random.RandomState(1).randint(1, V, 20))),
             (25, list(np.random.RandomState(2).randint(1, V, 11))), (60, [7]),
             (33, [2, 2, 2, 2, 9, 9, 4])]
    for i, (T, y) in enumerate(cases):
        h = torch.randn(1, T, D) * 3
        with torch.no_grad():
            lpz = m1.log_softmax(h).squeeze(0)
            ali = m1.forced_align(h, [int(v) for v in y])
        out[f"align{i}__lp"] = lpz.numpy()
        out[f"align{i}__y"] = np.asarray(y, dtype=np.int64)
        out[f"align{i}__ali"] = np.asarray(ali, dtype=np.int64)
    out["align_n"] = len(cases)

    captured = []
    fake = types.ModuleType("editdistance")
    fake.eval = lambda hyp, ref: captured.append(hyp) or 0
    sys.modules["editdistance"] = fake
    try:
        Vg, B, T = 40, 6, 70
        chars = [chr(0x4E00 + i) for i in range(Vg)]  # one unique character per id
        chars[0] = "<blank>"
        chars[Vg - 1] = "<space>"
        ec = ErrorCalculator(chars, "<space>", "<blank>", report_cer=False, report_wer=False)
        g = torch.Generator().manual_seed(17)
        logits = torch.randn(B, T, Vg, generator=g) * 2
        logits[:, :, 0] += 2.5      # plenty of blanks
        logits[:, :, Vg - 1] = -50  # the space symbol is never the arg-max (the collapse drops it)
        hold = torch.rand(B, T, generator=g) < 0.45  # repeat the previous frame: runs to collapse
        for t in range(1, T):
            logits[:, t][hold[:, t]] = logits[:, t - 1][hold[:, t]]
        ys_hat = logits.argmax(-1)
        ys_pad = torch.randint(1, Vg - 1, (B, 5))
        ec.calculate_cer_ctc(ys_hat, ys_pad)
    finally:
        del sys.modules["editdistance"]
    assert len(captured) == B
    ids = [[chars.index(c) for c in hyp] for hyp in captured]
    lmax = max(len(v) for v in ids)
    pad = np.full((B, lmax), -1, dtype=np.int64)
    for b, v in enumerate(ids):
        pad[b, : len(v)] = v
    out["greedy__logits"] = logits.numpy()
    out["greedy__ids"] = pad
    out["greedy__lens"] = np.asarray([len(v) for v in ids], dtype=np.int64)
    save("align_greedy.npz", out)


if __name__ == "__main__":
    torch.set_num_threads(1)
    if len(sys.argv) > 1:  # regenerate selected fixtures only: make_golden.py make_beam_multi ...
        for fn in sys.argv[1:]:
            globals()[fn]()
        sys.exit(0)
    make_ctc_module()
    make_ctc_c1()
    make_prefix()
    make_beam()
    make_beam_multi()
    make_beam_opts()
    make_streaming()
    make_align_greedy()
    make_beam_c3()
