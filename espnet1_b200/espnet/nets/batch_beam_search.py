"""Device-resident batch beam search around the CUDA CTC prefix scorer.

Same constructor, ``forward(x, maxlenratio, minlenratio)`` contract and n-best
results as the reference's ``BatchBeamSearch``
(/root/reference/espnet/nets/batch_beam_search.py:26-347 on top of
``BeamSearch.__init__/forward``, beam_search.py:34-118,346-431), rebuilt so that
the hypotheses never leave the device between steps and so that
several utterances decode in one pass (the ``recognize_beam_batch``
pattern of rnn/decoders.py:752-937 with the BatchBeamSearch semantics):

* the beam of every utterance is ``beam_size`` fixed SLOTS; a slot that ended
  (or was never filled) carries a running score of -inf, so the per-utterance
  ``topk`` over (slots x vocabulary) selects exactly what the reference selects
  from its shrinking list of running hypotheses (batch_beam_search.py:98-118);
* prefixes, running scores and per-scorer scores are device tensors
  updated by gathers; the reference rebuilds ``Hypothesis`` objects and
  ``torch.stack``s n_beam CTC states every step (:206-283, scorers/ctc.py:113-121);
* the CTC state stays batched: one ``index_select_state`` gather per step
  replaces n_beam ``select_state`` slices (``CTCPrefixScorer.batch_select_state``);
* ONE device->host read per step (tokens, back-pointers, scores of the new
  slots) feeds the host-side bookkeeping that the algorithm needs anyway:
  collecting ended hypotheses and Eq. (50) end detection; the prefixes are read
  only at positions where a hypothesis ends, the n-best lists go back in one upload.

Scorers that only implement the reference's list-of-states interface
(``batch_score`` / ``batch_score_partial`` / ``select_state``) are driven through
it unchanged (single utterance for partial scorers).
"""
from __future__ import annotations

import logging
import os
import time
from typing import Any, Dict, List

import numpy as np
import torch

from .beam_search import Hypothesis
from .e2e_asr_common import D_END, end_detect  # noqa: F401
from .scorer_interface import PartialScorerInterface, ScorerInterface

_NEG_INF = float("-inf")


def _is_partial(v) -> bool:
    return isinstance(v, PartialScorerInterface) or hasattr(v, "batch_score_partial")


class BatchBeamSearch(torch.nn.Module):
    """Batch beam search (label-synchronous, joint CTC/attention scoring)."""

    def __init__(self, scorers: Dict[str, ScorerInterface], weights: Dict[str, float], beam_size: int,
                 vocab_size: int, sos: int, eos: int, token_list: List[str] = None,
                 pre_beam_ratio: float = 1.5, pre_beam_score_key: str = None,
                 hyp_primer: List[int] = None):
        super().__init__()
        self.weights = weights
        self.scorers, self.full_scorers, self.part_scorers = dict(), dict(), dict()
        self.nn_dict = torch.nn.ModuleDict()  # recursive cast, as beam_search.py:71-73
        for k, v in scorers.items():
            if weights.get(k, 0) == 0 or v is None:
                continue
            self.scorers[k] = v
            if _is_partial(v):
                self.part_scorers[k] = v
            else:
                self.full_scorers[k] = v
            if isinstance(v, torch.nn.Module):
                self.nn_dict[k] = v
        self.sos, self.eos = sos, eos
        self.hyp_primer = hyp_primer
        self.token_list = token_list
        self.pre_beam_size = int(pre_beam_ratio * beam_size)
        self.beam_size = beam_size
        self.n_vocab = vocab_size
        if (pre_beam_score_key is not None and pre_beam_score_key != "full"
                and pre_beam_score_key not in self.full_scorers):
            raise KeyError(f"{pre_beam_score_key} is not found in {self.full_scorers}")
        self.pre_beam_score_key = pre_beam_score_key
        self.do_pre_beam = (self.pre_beam_score_key is not None and self.pre_beam_size < self.n_vocab
                            and len(self.part_scorers) > 0)
        self.last_steps = 0  # decoding positions of the most recent call (for benchmarks)
        # B200CTC_BEAM_PROF=1: synchronise after every phase and accumulate wall time per phase
        self.prof: Dict[str, float] = {} if os.environ.get("B200CTC_BEAM_PROF") else None

    def _tick(self, name: str, t0: float, dev) -> float:
        if self.prof is None:
            return t0
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        t1 = time.perf_counter()
        self.prof[name] = self.prof.get(name, 0.0) + (t1 - t0)
        return t1

    def set_hyp_primer(self, hyp_primer: List[int] = None) -> None:
        self.hyp_primer = hyp_primer

    # ------------------------------------------------------------------ API
    def forward(self, x: torch.Tensor, maxlenratio: float = 0.0,
                minlenratio: float = 0.0) -> List[Hypothesis]:
        """N-best list of ONE utterance ``x (T, D)`` (beam_search.py:346-431)."""
        nbest = self._decode(x.unsqueeze(0), [x.shape[0]], maxlenratio)[0]
        if len(nbest) == 0:
            logging.warning("there is no N-best results, perform recognition "
                            "again with smaller minlenratio.")
            return [] if minlenratio < 0.1 else self.forward(x, maxlenratio, max(0.0, minlenratio - 0.1))
        return nbest

    def forward_batch(self, xs: torch.Tensor, xlens, maxlenratio: float = 0.0) -> List[List[Hypothesis]]:
        """N-best lists of several utterances ``xs (U, Tmax, D)`` decoded together; each list
        equals ``forward(xs[u, :xlens[u]])``.  Partial scorers must offer the batched-state
        API of :class:`..scorers.ctc.CTCPrefixScorer`.  Full scorers never see padding frames
        unless they ask for them (see :meth:`_score_full`)."""
        lens = [int(v) for v in (xlens.tolist() if isinstance(xlens, torch.Tensor) else xlens)]
        return self._decode(xs, lens, maxlenratio)

    # ------------------------------------------------------------- internals
    @staticmethod
    def _score_full(d, yseq, states, xs, xs_rep, lens, W):
        """``d.batch_score`` of a full scorer for all slots.  With several utterances of
        DIFFERENT lengths the rows of ``xs`` are padded: a scorer that looks at the encoder output
        (an attention decoder) must not see the padding frames, or its scores differ from
        decoding the utterance alone.  So, unless the scorer declares that it ignores ``xs``
        (``ignores_x = True``, e.g. a length bonus or an LM) or masks the padding itself
        (``batch_score_padded(ys, states, xs, xlens)``), every utterance is scored on its own
        unpadded frames - exactly what ``forward(xs[u, :xlens[u]])`` does."""
        U = len(lens)
        if U == 1 or min(lens) == xs.shape[1] or getattr(d, "ignores_x", False):
            return d.batch_score(yseq, states, xs_rep)
        if hasattr(d, "batch_score_padded"):
            xl = torch.tensor(lens, device=xs.device).repeat_interleave(W)
            return d.batch_score_padded(yseq, states, xs_rep, xl)
        outs, new_states = [], []
        for u in range(U):
            sl = slice(u * W, (u + 1) * W)
            xu = xs[u, : lens[u]].unsqueeze(0).expand(W, *xs[u, : lens[u]].shape)
            sc_u, st_u = d.batch_score(yseq[sl], states[sl], xu)
            outs.append(sc_u)
            new_states.append(st_u)
        if all(st is None for st in new_states):
            merged = None
        else:
            merged = [st for st_u in new_states for st in (st_u if st_u is not None else [None] * W)]
        return torch.cat(outs, dim=0), merged

    def _maxlen(self, T: int, maxlenratio: float) -> int:
        if maxlenratio == 0:
            return T
        if maxlenratio < 0:
            return -1 * int(maxlenratio)
        return max(1, int(maxlenratio * T))

    def _decode(self, xs: torch.Tensor, lens: List[int], maxlenratio: float) -> List[List[Hypothesis]]:
        U, W, V, dev = xs.shape[0], self.beam_size, self.n_vocab, xs.device
        n = U * W
        maxlen = [self._maxlen(T, maxlenratio) for T in lens]
        primer = [self.sos] if self.hyp_primer is None else list(self.hyp_primer)

        # ---- initial hypotheses (batch_beam_search.py:120-144): slot 0 of every utterance
        full_states: Dict[str, List[Any]] = {}
        for k, d in self.full_scorers.items():
            init = [d.batch_init_state(xs[u, : lens[u]]) if hasattr(d, "batch_init_state")
                    else d.init_state(xs[u, : lens[u]]) for u in range(U)]
            full_states[k] = [init[u] for u in range(U) for _ in range(W)]
        part_states: Dict[str, Any] = {}
        batched: Dict[str, bool] = {}
        for k, d in self.part_scorers.items():
            batched[k] = hasattr(d, "batch_select_state")
            if batched[k]:
                part_states[k] = d.batch_init_state_multi(xs, lens)
            else:
                if U != 1:
                    raise ValueError(f"partial scorer {k} has no batched-state API: one utterance at a time")
                part_states[k] = [d.batch_init_state(xs[0, : lens[0]])] * n
        yseq = torch.tensor(primer, dtype=torch.int64, device=dev).repeat(n, 1)
        score = torch.full((U, W), _NEG_INF, dtype=xs.dtype, device=dev)
        score[:, 0] = 0.0
        names = list(self.full_scorers) + list(self.part_scorers)  # accumulation order (:217-241)
        sc = {k: torch.zeros(n, dtype=xs.dtype, device=dev) for k in names}
        xs_rep = xs.expand(n, *xs.shape[1:]) if U == 1 else xs.repeat_interleave(W, dim=0)
        slot0 = (torch.arange(U, device=dev) * W).unsqueeze(1)            # first slot of each utterance
        maxlen_t = torch.tensor(maxlen, device=dev).unsqueeze(1)
        active = torch.ones(U, 1, dtype=torch.bool, device=dev)          # utterance still decoding
        # fused pruning kernel (csrc/beam.cu): CUDA, fp32, segment and winner counts within its limits
        fused = (dev.type == "cuda" and xs.dtype == torch.float32 and W * V <= 57344 and V <= 57344 and W <= 64
                 and (not self.do_pre_beam or self.pre_beam_size <= 64)
                 and not os.environ.get("B200CTC_BEAM_TORCH"))
        if fused:
            from ... import functional as F

            P0 = len(primer)
            ybuf = [torch.empty(n, P0 + max(maxlen) + 1, dtype=torch.int64, device=dev) for _ in range(2)]
            ybuf[0][:, :P0] = yseq
            cum = torch.zeros(len(names), n, dtype=torch.float32, device=dev)
            active_u8 = torch.ones(U, dtype=torch.uint8, device=dev)
            maxlen_i32 = torch.tensor(maxlen, dtype=torch.int32, device=dev)
            yseq = ybuf[0][:, :P0]
        ended: List[List[dict]] = [[] for _ in range(U)]
        best_all = [-np.inf] * U
        best_len: List[Dict[int, float]] = [dict() for _ in range(U)]
        done_np = np.zeros(U, dtype=bool)
        maxlen_np = np.asarray(maxlen)

        i = -1
        for i in range(max(maxlen)):
            tp = self._tick("host_other", time.perf_counter(), dev) if self.prof is not None else 0.0
            # ---- scoring (batch_beam_search.py:217-241), same accumulation order
            w = None if fused else torch.zeros(n, V, dtype=xs.dtype, device=dev)
            step_sc: Dict[str, torch.Tensor] = {}
            new_full: Dict[str, Any] = {}
            for k, d in self.full_scorers.items():
                step_sc[k], new_full[k] = self._score_full(d, yseq, full_states[k], xs, xs_rep, lens, W)
                if not fused:
                    w += self.weights[k] * step_sc[k]
                elif step_sc[k].dtype != torch.float32:
                    step_sc[k] = step_sc[k].float()  # the pruning kernel reads fp32 matrices
            tp = self._tick("full_scorers", tp, dev)
            part_ids = None
            if self.do_pre_beam and fused:
                if self.pre_beam_score_key == "full":
                    part_ids = F.beam_prebeam([step_sc[k] for k in self.full_scorers],
                                              [self.weights[k] for k in self.full_scorers], self.pre_beam_size)
                else:
                    part_ids = F.beam_prebeam([step_sc[self.pre_beam_score_key]], [1.0], self.pre_beam_size)
            elif self.do_pre_beam:
                src = w if self.pre_beam_score_key == "full" else step_sc[self.pre_beam_score_key]
                part_ids = torch.topk(src, self.pre_beam_size, dim=-1)[1]
            tp = self._tick("pre_beam_topk", tp, dev)
            new_part: Dict[str, Any] = {}
            for k, d in self.part_scorers.items():
                if batched[k]:
                    step_sc[k], new_part[k] = d.batch_score_partial_batched(yseq, part_ids, part_states[k])
                else:
                    step_sc[k], new_part[k] = d.batch_score_partial(yseq, part_ids, part_states[k], xs[0, : lens[0]])
                if not fused:
                    w += self.weights[k] * step_sc[k]
            tp = self._tick("part_scorers", tp, dev)

            # ---- beam (batch_beam_search.py:98-118), per utterance
            if fused:  # weighted sum, top-k, score gathers, prefix append, slot masks: one kernel
                cur_len = yseq.shape[1]
                nxt = ybuf[(i + 1) & 1]
                _, top, score, cum, pack_d = F.beam_select(
                    [step_sc[k] for k in names], [self.weights[k] for k in names], score, cum,
                    yseq, cur_len, active_u8, maxlen_i32, i, U, W, W, self.eos, nxt)
                yseq = nxt[:, : cur_len + 1]
            else:
                w += score.view(n, 1)
                top_v, top = w.view(U, W * V).topk(W)
                prev, tok = top // V, top % V
                fprev = (prev + slot0).view(-1)
                ftok = tok.view(-1)
                for k in names:
                    sc[k] = sc[k][fprev] + step_sc[k][fprev, ftok]
                yseq = torch.cat((yseq[fprev], ftok.unsqueeze(1)), 1)
                last = maxlen_t - 1 == i
                alive = active & ~last & (tok != self.eos) & torch.isfinite(top_v)
                score = torch.where(alive, top_v, torch.full_like(top_v, _NEG_INF))
                # slot ids and tokens are < 2^24: exact in fp32
                pack_d = torch.stack([top_v.float(), prev.float(), tok.float()]
                                     + [sc[k].view(U, W).float() for k in names])
            for k, d in self.part_scorers.items():
                if batched[k]:
                    part_states[k] = d.batch_select_state(new_part[k], top)
            tp = self._tick("beam_topk_gathers", tp, dev)
            pack = pack_d.cpu().numpy()  # the one device->host read of the step
            h_val = pack[0]
            h_prev = pack[1].astype(np.int64)
            h_tok = pack[2].astype(np.int64)
            tp = self._tick("d2h_read", tp, dev)
            for k, d in self.full_scorers.items():
                st = new_full[k]
                if st is None:  # stateless scorer (ScorerInterface.select_state(None, i) is None)
                    continue
                full_states[k] = [d.select_state(st, u * W + int(h_prev[u, j])) if hasattr(d, "select_state")
                                  else st[u * W + int(h_prev[u, j])]
                                  for u in range(U) for j in range(W)]
            for k, d in self.part_scorers.items():
                if not batched[k]:
                    part_states[k] = [d.select_state(new_part[k], int(h_prev[0, j]), int(h_tok[0, j]))
                                      for j in range(W)]

            # ---- post-process (batch_beam_search.py:285-347) + end detection, on the host.
            # The prefixes stay on the device; they are read (one copy) only at a position
            # where some hypothesis ends.
            valid = h_val != -np.inf            # fewer than beam_size candidates: not a hypothesis
            is_last = maxlen_np - 1 == i        # eos in the final loop (:318-335)
            ends = valid & ((h_tok == self.eos) | is_last[:, None]) & ~done_np[:, None]
            running = (valid & ~ends).sum(1)
            changed = False
            y_host = yseq.cpu().numpy() if ends.any() else None
            for u, j in zip(*np.nonzero(ends)):
                seq = y_host[int(u) * W + int(j)].tolist()
                if is_last[u]:
                    seq.append(self.eos)
                states = {k: None for k in self.full_scorers}
                for k, d in self.part_scorers.items():
                    states[k] = (d.slot_state(part_states[k], int(u) * W + int(j)) if batched[k]
                                 else part_states[k][int(j)])
                v = float(h_val[u, j])
                ended[u].append({"yseq": seq, "score": v,
                                 "scores": {k: float(pack[3 + q][u, j]) for q, k in enumerate(names)},
                                 "states": states})
                best_all[u] = max(best_all[u], v)
                best_len[u][len(seq)] = max(best_len[u].get(len(seq), -np.inf), v)
            for u in range(U):
                if done_np[u]:
                    continue
                if (maxlenratio == 0.0 and self._end_detect(best_all[u], best_len[u], i)) or running[u] == 0:
                    done_np[u] = True
                    changed = True
            tp = self._tick("host_bookkeeping", tp, dev)
            if done_np.all():
                break
            if changed:
                active = torch.from_numpy(~done_np).to(dev).unsqueeze(1)
                score = torch.where(active, score, torch.full_like(score, _NEG_INF))
                if fused:
                    active_u8 = active.view(-1).to(torch.uint8)
        self.last_steps = i + 1

        # n-best lists: ONE upload of all token sequences (padded) and one tensor of all scores;
        # every Hypothesis holds views of them (yseq on the input's device like the reference's)
        order = [sorted(ended[u], key=lambda h: h["score"], reverse=True) for u in range(U)]
        flat = [h for nb in order for h in nb]
        out: List[List[Hypothesis]] = [[] for _ in range(U)]
        if flat:
            lmax = max(len(h["yseq"]) for h in flat)
            ypad = np.full((len(flat), lmax), self.eos, dtype=np.int64)
            vals = np.empty((len(flat), 1 + len(names)), dtype=np.float64)
            for r, h in enumerate(flat):
                ypad[r, : len(h["yseq"])] = h["yseq"]
                vals[r, 0] = h["score"]
                for q, k in enumerate(names):
                    vals[r, 1 + q] = h["scores"][k]
            y_all = torch.from_numpy(ypad).to(dev)
            v_all = torch.from_numpy(vals).to(xs.dtype)
            r = 0
            for u in range(U):
                for h in order[u]:
                    out[u].append(Hypothesis(yseq=y_all[r, : len(h["yseq"])], score=v_all[r, 0],
                                             scores={k: v_all[r, 1 + q] for q, k in enumerate(names)},
                                             states=h["states"]))
                    r += 1
        return out

    @staticmethod
    def _end_detect(best: float, best_by_len: Dict[int, float], i: int, M: int = 3) -> bool:
        """``end_detect`` (e2e_asr_common.py:18-49) on running maxima: ``best`` over all ended
        hypotheses and ``best_by_len[n]`` over those of length n."""
        if not best_by_len:
            return False
        count = 0
        for m in range(M):
            v = best_by_len.get(i - m)
            if v is not None and v - best < D_END:
                count += 1
        return count == M
