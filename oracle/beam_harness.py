"""Label-synchronous beam search driver (the CALLER of the prefix scorer).

TEST INFRASTRUCTURE ONLY (see oracle/__init__).  A compact restatement of the
control flow of ``BeamSearch`` (espnet/nets/beam_search.py:291-536) and
``BatchBeamSearch`` (espnet/nets/batch_beam_search.py:98-347), relative to
/root/reference, so that token-sequence parity of a scorer can be checked on a
box without the reference tree.  Scorers are duck-typed on the reference's
scorer interface (espnet/nets/scorer_interface.py:9-186): full scorers expose
``score``/``batch_score``; partial scorers expose ``score_partial``/
``batch_score_partial``/``select_state``.
"""
from __future__ import annotations

import math

import torch

D_END = math.log(1 * math.exp(-10))  # e2e_asr_common.py:18


def end_detect(ended, i, M=3, d_end=D_END):
    """Eq. (50) of Watanabe et al. 2017 (espnet/nets/e2e_asr_common.py:18-49)."""
    if not ended:
        return False
    best = max(h["score"] for h in ended)
    hits = 0
    for m in range(M):
        same = [h["score"] for h in ended if len(h["yseq"]) == i - m]
        if same and max(same) - best < d_end:
            hits += 1
    return hits == M


def _is_partial(s):
    return hasattr(s, "score_partial")


def _select(scorer, state, i, new_id=None):
    if hasattr(scorer, "select_state"):
        return scorer.select_state(state, i, new_id) if new_id is not None else scorer.select_state(state, i)
    return None if state is None else state[i]


def _split(scorers, weights):
    act = {k: v for k, v in scorers.items() if v is not None and weights.get(k, 0) != 0}
    full = {k: v for k, v in act.items() if not _is_partial(v)}
    part = {k: v for k, v in act.items() if _is_partial(v)}
    return act, full, part


def _maxlen(x, maxlenratio):
    if maxlenratio == 0:
        return x.shape[0]
    if maxlenratio < 0:
        return -int(maxlenratio)
    return max(1, int(maxlenratio * x.size(0)))


def _finish(ended):
    out = sorted(ended, key=lambda h: float(h["score"]), reverse=True)
    return [
        {"yseq": [int(v) for v in h["yseq"]], "score": float(h["score"]),
         "scores": {k: float(v) for k, v in h["scores"].items()}}
        for h in out
    ]


def batch_beam_search(x, scorers, weights, beam_size, vocab_size, sos, eos,
                      pre_beam_ratio=1.5, pre_beam_score_key=None, maxlenratio=0.0):
    """``BatchBeamSearch.forward`` for one utterance ``x (T, D)``."""
    act, full, part = _split(scorers, weights)
    pre_beam = int(pre_beam_ratio * beam_size)
    do_pre = pre_beam_score_key is not None and pre_beam < vocab_size and len(part) > 0
    maxlen = _maxlen(x, maxlenratio)
    hyps = [{
        "yseq": torch.tensor([sos], device=x.device), "score": 0.0,
        "scores": {k: 0.0 for k in act},
        "states": {k: (d.batch_init_state(x) if hasattr(d, "batch_init_state") else d.init_state(x))
                   for k, d in act.items()},
    }]
    ended = []
    for i in range(maxlen):
        n = len(hyps)
        yseq = torch.stack([h["yseq"] for h in hyps])
        run_score = torch.tensor([float(h["score"]) for h in hyps])
        w = torch.zeros(n, vocab_size, dtype=x.dtype, device=x.device)
        f_scores, f_states = {}, {}
        for k, d in full.items():
            f_scores[k], f_states[k] = d.batch_score(
                yseq, [h["states"][k] for h in hyps], x.expand(n, *x.shape))
            w += weights[k] * f_scores[k]
        ids = None
        if do_pre:
            src = w if pre_beam_score_key == "full" else f_scores[pre_beam_score_key]
            ids = torch.topk(src, pre_beam, dim=-1)[1]
        p_scores, p_states = {}, {}
        for k, d in part.items():
            p_scores[k], p_states[k] = d.batch_score_partial(
                yseq, ids, [h["states"][k] for h in hyps], x)
            w += weights[k] * p_scores[k]
        w += run_score.to(dtype=x.dtype, device=x.device).unsqueeze(1)
        top = w.view(-1).topk(beam_size)[1]
        new = []
        for hid, tok in zip((top // vocab_size).tolist(), (top % vocab_size).tolist()):
            prev = hyps[hid]
            sc = {k: prev["scores"][k] + f_scores[k][hid][tok] for k in full}
            sc.update({k: prev["scores"][k] + p_scores[k][hid][tok] for k in part})
            st = {k: _select(full[k], f_states[k], hid) for k in full}
            st.update({k: part[k].select_state(p_states[k], hid, tok) for k in part})
            new.append({
                "yseq": torch.cat((prev["yseq"], torch.tensor([tok], dtype=prev["yseq"].dtype,
                                                              device=prev["yseq"].device))),
                "score": w[hid, tok], "scores": sc, "states": st})
        if i == maxlen - 1:
            for h in new:
                h["yseq"] = torch.cat((h["yseq"], torch.tensor([eos], dtype=h["yseq"].dtype,
                                                               device=h["yseq"].device)))
        hyps = []
        for h in new:
            (ended if int(h["yseq"][-1]) == eos else hyps).append(h)
        if maxlenratio == 0.0 and end_detect(
                [{"yseq": h["yseq"].tolist(), "score": float(h["score"])} for h in ended], i):
            break
        if not hyps:
            break
    return _finish(ended)


def beam_search(x, scorers, weights, beam_size, vocab_size, sos, eos,
                pre_beam_ratio=1.5, pre_beam_score_key=None, maxlenratio=0.0):
    """Per-hypothesis ``BeamSearch.forward`` (non-batched scorer methods)."""
    act, full, part = _split(scorers, weights)
    pre_beam = int(pre_beam_ratio * beam_size)
    do_pre = pre_beam_score_key is not None and pre_beam < vocab_size and len(part) > 0
    maxlen = _maxlen(x, maxlenratio)
    hyps = [{"yseq": torch.tensor([sos], device=x.device), "score": 0.0,
             "scores": {k: 0.0 for k in act},
             "states": {k: d.init_state(x) for k, d in act.items()}}]
    ended = []
    for i in range(maxlen):
        best = []
        for hyp in hyps:
            ids = torch.arange(vocab_size, device=x.device)
            w = torch.zeros(vocab_size, dtype=x.dtype, device=x.device)
            f_scores, f_states, p_scores, p_states = {}, {}, {}, {}
            for k, d in full.items():
                f_scores[k], f_states[k] = d.score(hyp["yseq"], hyp["states"][k], x)
                w += weights[k] * f_scores[k]
            if do_pre:
                src = w if pre_beam_score_key == "full" else f_scores[pre_beam_score_key]
                ids = torch.topk(src, pre_beam)[1]
            for k, d in part.items():
                p_scores[k], p_states[k] = d.score_partial(hyp["yseq"], ids, hyp["states"][k], x)
                w[ids] += weights[k] * p_scores[k]
            w += hyp["score"]
            if w.size(0) == ids.size(0):
                top = w.topk(beam_size)[1]
                loc = top
            else:
                keep = w[ids]
                w[:] = -float("inf")
                w[ids] = keep
                top = w.topk(beam_size)[1]
                loc = w[ids].topk(beam_size)[1]
            for j, pj in zip(top.tolist(), loc.tolist()):
                sc = {k: hyp["scores"][k] + f_scores[k][j] for k in full}
                sc.update({k: hyp["scores"][k] + p_scores[k][pj] for k in part})
                st = dict(f_states)
                st.update({k: part[k].select_state(p_states[k], pj) for k in part})
                best.append({
                    "yseq": torch.cat((hyp["yseq"], torch.tensor([j], dtype=hyp["yseq"].dtype,
                                                                 device=hyp["yseq"].device))),
                    "score": w[j], "scores": sc, "states": st})
            best = sorted(best, key=lambda h: h["score"], reverse=True)[:beam_size]
        if i == maxlen - 1:
            for h in best:
                h["yseq"] = torch.cat((h["yseq"], torch.tensor([eos], dtype=h["yseq"].dtype,
                                                               device=h["yseq"].device)))
        hyps = []
        for h in best:
            if int(h["yseq"][-1]) == eos:
                for k, d in act.items():
                    s = d.final_score(h["states"][k]) if hasattr(d, "final_score") else 0.0
                    h["scores"][k] = h["scores"][k] + s
                    h["score"] = h["score"] + weights[k] * s
                ended.append(h)
            else:
                hyps.append(h)
        if maxlenratio == 0.0 and end_detect(
                [{"yseq": h["yseq"].tolist(), "score": float(h["score"])} for h in ended], i):
            break
        if not hyps:
            break
    return _finish(ended)
