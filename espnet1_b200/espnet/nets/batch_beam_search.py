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
from collections.abc import Mapping
from typing import Any, Dict, List

import numpy as np
import torch

from .beam_search import Hypothesis
from .e2e_asr_common import D_END, end_detect  # noqa: F401
from .scorer_interface import PartialScorerInterface, ScorerInterface

_NEG_INF = float("-inf")


def _is_partial(v) -> bool:
    return isinstance(v, PartialScorerInterface) or hasattr(v, "batch_score_partial")


class _Ended:
    """Ended hypotheses of U utterances (batch_beam_search.py:285-347), kept as one numpy block
    per decoding position instead of one Python record per hypothesis: at the last position
    every running hypothesis ends at once (U x beam of them), and per-record dicts, float()
    conversions and tensor views were a third of a decode's host time."""

    def __init__(self, U: int, names: List[str], eos: int):
        self.U, self.names, self.eos = U, names, eos
        self.u: List[np.ndarray] = []        # utterance of every ended hypothesis, per block
        self.val: List[np.ndarray] = []      # total score
        self.sc: List[np.ndarray] = []       # (K, m) per-scorer scores
        self.rows: List[np.ndarray] = []     # token rows, one 1-D array per hypothesis
        self.states: List[Dict[str, Any]] = []
        self.best_all = np.full(U, -np.inf)
        self.best_len: List[Dict[int, float]] = [dict() for _ in range(U)]

    def add(self, us, js, pack, y_rows, is_last, states) -> None:
        """``us, js``: utterance / slot of the hypotheses that end at this position; ``pack`` the
        position's host record; ``y_rows`` (m, L) their prefixes; ``states`` one dict each."""
        vals = pack[0][us, js].astype(np.float64)
        self.u.append(us)
        self.val.append(vals)
        self.sc.append(pack[3:, us, js].astype(np.float64))
        L = y_rows.shape[1]
        last = is_last[us]
        if last.any():  # eos appended in the final loop (:318-335)
            y_rows = np.concatenate((y_rows, np.full((len(us), 1), self.eos, dtype=y_rows.dtype)), 1)
            self.rows.extend(r[: L + 1] if e else r[:L] for r, e in zip(y_rows, last))
        else:
            self.rows.extend(y_rows)
        self.states.extend(states)
        np.maximum.at(self.best_all, us, vals)
        top = np.full(self.U, -np.inf)
        np.maximum.at(top, us, vals)
        for u in np.unique(us):
            n = L + 1 if is_last[u] else L
            d = self.best_len[u]
            d[n] = max(d.get(n, -np.inf), float(top[u]))

    def nbest(self, dev, dtype) -> List[List[Hypothesis]]:
        """Per utterance, sorted by descending score (stable, like ``sorted(..., reverse=True)``).
        ONE upload of all token sequences and one tensor of all scores; every Hypothesis holds
        views of them (yseq on the input's device like the reference's)."""
        out: List[List[Hypothesis]] = [[] for _ in range(self.U)]
        if not self.rows:
            return out
        us, vals = np.concatenate(self.u), np.concatenate(self.val)
        sc = np.concatenate(self.sc, axis=1)
        order = np.lexsort((np.arange(len(us)), -vals, us))
        rows = [self.rows[r] for r in order]
        y_all = torch.from_numpy(np.concatenate(rows)).to(dev)
        ys = torch.split(y_all, [len(r) for r in rows])
        # score / scores as Python floats (the Hypothesis annotation's Union[float, Tensor]): a
        # 0-dim tensor per value costs ~1 us each, U x beam x (1 + scorers) of them per decode
        v_tot = vals[order].tolist()
        v_sc = sc[:, order].T.tolist()
        names, states, make = self.names, self.states, Hypothesis._make  # (yseq, score, scores, states)
        for q, (r, u) in enumerate(zip(order.tolist(), us[order].tolist())):
            out[u].append(make((ys[q], v_tot[q], dict(zip(names, v_sc[q])), states[r])))
        return out


class _SlotStates(Mapping):
    """``Hypothesis.states`` of a hypothesis that ended in the position-call path: full scorers
    map to None, the CTC scorer to the (r (T,2), s (V,), f_min, f_max) tuple its
    ``select_state`` would give - built on first access (two tensor views per hypothesis that
    nobody reads after a hypothesis ended were a tenth of a decode's host time)."""

    __slots__ = ("_keys", "_part", "_r", "_s", "_slot", "_val")

    def __init__(self, keys, part, r, s, slot):
        self._keys, self._part, self._r, self._s, self._slot, self._val = keys, part, r, s, slot, None

    def __getitem__(self, k):
        if k == self._part:
            if self._val is None:
                self._val = (self._r[:, :, self._slot], self._s[self._slot], 0, 0)
            return self._val
        if k in self._keys:
            return None
        raise KeyError(k)

    def __iter__(self):
        return iter(self._keys)

    def __len__(self):
        return len(self._keys)


class _PinnedPool:
    """Pinned host buffers kept across decodes: the per-position record and the staging of a few
    small uploads.  Allocating pinned memory per decode now and then costs tens of milliseconds
    (cudaHostAlloc when the host allocator's cached block is still tied to an unfinished event)."""

    def __init__(self):
        self._buf: Dict[Any, torch.Tensor] = {}

    def __call__(self, name, shape, dtype) -> torch.Tensor:
        key = (name, tuple(shape), dtype)
        if key not in self._buf:
            self._buf[key] = torch.empty(*shape, dtype=dtype, pin_memory=torch.cuda.is_available())
        return self._buf[key]

    def upload(self, name, values, dtype, dev) -> torch.Tensor:
        """A few host values as a device tensor without blocking: a pageable host-to-device copy
        waits until the GPU has drained everything queued before it (here: the CTC projection)."""
        t = torch.tensor(values, dtype=dtype)
        if dev.type != "cuda":
            return t.to(dev)
        stage = self(name, t.shape, dtype)
        stage.copy_(t)
        return stage.to(dev, non_blocking=True)


class _PositionPlan:
    """Buffers of one decode for ``b200ctc_beam_position`` (include/b200ctc.h): everything that
    follows the full scorers in a position - pre-beam, CTC prefix scores, pruning, CTC state
    selection, the step record's copy to pinned memory - is ONE library call on these buffers.
    The generic path below makes the same library calls one by one through the scorer classes,
    with ~20 allocations and four argument marshallings per position."""

    def __init__(self, impl, U, W, V, K_full, weights, snum, prebeam_key, eos, yseq0, stride, maxlen, pinned):
        import ctypes

        from ..._lib import BeamPlan

        dev, T, n, K = impl.device, impl.input_length, U * W, K_full + 1
        f32 = dict(dtype=torch.float32, device=dev)
        i64 = dict(dtype=torch.int64, device=dev)
        sw = snum if snum > 0 else V
        t = self.t = {  # keeps the tensors alive as long as the plan
            "lp": impl._lp, "end_frames": impl._end_frames_dev,
            "lp_t": impl._time_contiguous() if snum > 0 else None,
            # CTC state of the slots: [i & 1] is read by position i, [(i + 1) & 1] written
            "r_state0": torch.empty(T, 2, n, **f32), "s_state0": torch.empty(n, V, **f32),
            "r_state1": torch.empty(T, 2, n, **f32), "s_state1": torch.empty(n, V, **f32),
            "r": torch.empty(T, 2, n, sw, **f32), "log_psi": torch.empty(n, V, **f32),
            "ctc_scores": torch.empty(n, V, **f32),
            "idmap": torch.empty(n, V, **i64) if snum > 0 else None,
            "rsum": torch.empty(T, n, **f32),
            "part_ids": torch.empty(n, snum, **i64) if snum > 0 else None,
            "run0": torch.full((U, W), _NEG_INF, **f32), "run1": torch.empty(U, W, **f32),
            "cum0": torch.zeros(K, n, **f32), "cum1": torch.empty(K, n, **f32),
            "y0": torch.empty(n, stride, **i64), "y1": torch.empty(n, stride, **i64),
            "active": torch.ones(U, dtype=torch.uint8, device=dev),
            "maxlen": pinned.upload("maxlen", maxlen, torch.int32, dev),
            "out_val": torch.empty(n, **f32), "out_flat": torch.empty(n, **i64),
            "pack": torch.empty(3 + K, U, W, **f32),
            # two records: position i+1 may be issued while the host still reads position i's
            "pack_host": pinned("pack", (2, 3 + K, U, W), torch.float32),
        }
        t["run0"][:, 0] = 0.0
        t["y0"][:, : yseq0.shape[1]] = yseq0
        c = self.c = BeamPlan()
        c.U, c.W, c.V, c.T, c.snum, c.K_full, c.prebeam_key = U, W, V, T, snum, K_full, prebeam_key
        c.blank, c.eos, c.logzero = impl.blank, eos, impl.logzero
        for q, w in enumerate(weights):
            c.weights[q] = float(w)
        for name in ("lp", "lp_t", "end_frames", "r", "log_psi", "ctc_scores", "idmap",
                     "rsum", "part_ids", "active", "maxlen", "out_val", "out_flat", "pack"):
            setattr(c, name, None if t[name] is None else t[name].data_ptr())
        c.lp_t_ld = t["lp_t"].shape[2] if snum > 0 else T
        for b in (0, 1):
            c.run_score[b], c.cum[b], c.yseq[b] = (t[f"run{b}"].data_ptr(), t[f"cum{b}"].data_ptr(),
                                                   t[f"y{b}"].data_ptr())
            c.r_state[b], c.s_state[b] = t[f"r_state{b}"].data_ptr(), t[f"s_state{b}"].data_ptr()
        c.yseq_stride = stride
        self.ref = ctypes.byref(c)
        self.fptrs = (ctypes.c_void_p * max(K_full, 1))()
        self.fstrides = (ctypes.c_int64 * (2 * max(K_full, 1)))()
        self.pack_np = t["pack_host"].numpy()
        self.pack_ptr = [t["pack_host"][b].data_ptr() for b in (0, 1)]
        self.ybuf = (t["y0"], t["y1"])
        self.run = (t["run0"], t["run1"])
        self.stream = torch.cuda.current_stream(dev)
        self.done = (torch.cuda.Event(), torch.cuda.Event())  # position i's record has arrived
        self.U, self.W = U, W

    def launch(self, mats, step_i: int, cur_len: int) -> None:
        """Issue position ``step_i`` on this step's full-scorer matrices (asynchronous)."""
        from ..._lib import check, lib

        fp, fs = self.fptrs, self.fstrides
        for q, m in enumerate(mats):
            if m.dtype is not torch.float32 or m.dim() != 2 or not m.is_cuda:
                raise TypeError("beam score matrices must be 2-D fp32 CUDA tensors")
            fp[q] = m.data_ptr()
            fs[2 * q], fs[2 * q + 1] = m.stride()
        self.c.pack_host = self.pack_ptr[step_i & 1]
        check(lib().b200ctc_beam_position(self.ref, fp, fs, step_i, cur_len, self.stream.cuda_stream),
              "beam_position")
        self.done[step_i & 1].record(self.stream)

    def wait(self, step_i: int) -> np.ndarray:
        """The host record of position ``step_i`` (at most one younger position is in flight)."""
        self.done[step_i & 1].synchronize()
        return self.pack_np[step_i & 1]

    def deactivate(self, done_np: np.ndarray, next_step: int) -> None:
        """Utterances that finished take no further part: their slots' running scores become -inf
        (what the reference achieves by dropping their hypotheses from the running list).
        ``next_step``: the next position still to be issued - its input buffer is masked."""
        act = torch.from_numpy(~done_np).to(self.t["active"].device)
        self.t["active"].copy_(act)
        self.run[next_step & 1].masked_fill_(~act.unsqueeze(1), _NEG_INF)

    def ended_states(self, step_i: int):
        """A COPY of the CTC state of all slots after position ``step_i``.  That buffer is only
        READ by position step_i + 1 (which may already be in flight) and rewritten by step_i + 2,
        which is not issued before this copy is queued."""
        b = (step_i + 1) & 1
        return self.t[f"r_state{b}"].clone(), self.t[f"s_state{b}"].clone()


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
        # with the CUDA CTC scorer as the one partial scorer, run everything after the full scorers
        # as ONE library call per position (b200ctc_beam_position); False = the scorer-by-scorer path
        self.fuse_position = True
        self._pinned = _PinnedPool()
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
        if hasattr(d, "batch_select_state"):
            raise ValueError("a full scorer with device-resident state (batch_select_state) that reads the "
                             "encoder output must offer batch_score_padded(ys, states, xs, xlens) to decode "
                             "utterances of different lengths together")
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
        dev = xs.device
        if dev.type == "cuda" and dev.index is not None and dev.index != torch.cuda.current_device():
            with torch.cuda.device(dev):  # the library launches on the current device
                return self._decode_on_device(xs, lens, maxlenratio)
        return self._decode_on_device(xs, lens, maxlenratio)

    def _decode_on_device(self, xs, lens, maxlenratio):
        U, W, V, dev = xs.shape[0], self.beam_size, self.n_vocab, xs.device
        n = U * W
        t_start = time.perf_counter()
        maxlen = [self._maxlen(T, maxlenratio) for T in lens]
        primer = [self.sos] if self.hyp_primer is None else list(self.hyp_primer)
        P0 = len(primer)

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
        yseq = self._pinned.upload("primer", primer, torch.int64, dev).repeat(n, 1)
        names = list(self.full_scorers) + list(self.part_scorers)  # accumulation order (:217-241)
        # a scorer that ignores the encoder output does not need the (n, T, D) copy of it
        need_x = any(not getattr(d, "ignores_x", False) for d in self.full_scorers.values())
        xs_rep = xs[:1].expand(n, *xs.shape[1:]) if (U == 1 or not need_x) else xs.repeat_interleave(W, dim=0)
        # fused pruning kernel (csrc/beam.cu): CUDA, fp32, segment and winner counts within its limits
        fused = (dev.type == "cuda" and xs.dtype == torch.float32 and W * V <= 55296 and V <= 55296 and W <= 64
                 and (not self.do_pre_beam or self.pre_beam_size <= 64)
                 and not os.environ.get("B200CTC_BEAM_TORCH"))
        plan = None
        if fused and self.fuse_position and len(self.part_scorers) == 1 and len(names) <= 8:
            from .scorers.ctc import CTCPrefixScorer

            (kp, dp), = self.part_scorers.items()
            if type(dp) is CTCPrefixScorer and dp.impl.dtype == torch.float32 and dp.impl.margin == 0:
                fkeys = list(self.full_scorers)
                snum = self.pre_beam_size if self.do_pre_beam else 0
                key = -1 if (not self.do_pre_beam or self.pre_beam_score_key == "full") else fkeys.index(
                    self.pre_beam_score_key)
                plan = _PositionPlan(dp.impl, U, W, V, len(fkeys), [self.weights[k] for k in names], snum, key,
                                     self.eos, yseq, P0 + max(maxlen) + 1, maxlen, self._pinned)
                yseq = plan.ybuf[0][:, :P0]
        if plan is None:
            score = torch.full((U, W), _NEG_INF, dtype=xs.dtype, device=dev)
            score[:, 0] = 0.0
            sc = {k: torch.zeros(n, dtype=xs.dtype, device=dev) for k in names}
            slot0 = (torch.arange(U, device=dev) * W).unsqueeze(1)        # first slot of each utterance
            maxlen_t = torch.tensor(maxlen, device=dev).unsqueeze(1)
            active = torch.ones(U, 1, dtype=torch.bool, device=dev)      # utterance still decoding
        if fused and plan is None:
            from ... import functional as F

            ybuf = [torch.empty(n, P0 + max(maxlen) + 1, dtype=torch.int64, device=dev) for _ in range(2)]
            ybuf[0][:, :P0] = yseq
            cum = torch.zeros(len(names), n, dtype=torch.float32, device=dev)
            active_u8 = torch.ones(U, dtype=torch.uint8, device=dev)
            maxlen_i32 = torch.tensor(maxlen, dtype=torch.int32, device=dev)
            yseq = ybuf[0][:, :P0]
        ended = _Ended(U, names, self.eos)
        done_np = np.zeros(U, dtype=bool)
        maxlen_np = np.asarray(maxlen)
        eos_f = np.float32(self.eos)
        stateful = None  # full scorers that returned states (decided at the first position)
        # Full scorers with DEVICE-resident state (an extension of the scorer interface, like the
        # CTC scorer's batched-state API): ``batch_select_state(states, prev_rows)`` with
        # ``prev_rows`` an (n,) int64 device tensor of the surviving slots' source rows returns the
        # states of all n slots - no list of per-hypothesis states, no host read between positions.
        dev_state = [k for k, d in self.full_scorers.items() if hasattr(d, "batch_select_state")]
        slot0_flat = torch.arange(U, device=dev).repeat_interleave(W) * W  # first row of each slot's utterance

        def score_full(i_pos, ys):
            """full scorers of one position (batch_beam_search.py:217-241), fp32 for the kernels"""
            out_sc: Dict[str, torch.Tensor] = {}
            out_st: Dict[str, Any] = {}
            for k, d in self.full_scorers.items():
                out_sc[k], out_st[k] = self._score_full(d, ys, full_states[k], xs, xs_rep, lens, W)
                if fused and out_sc[k].dtype != torch.float32:
                    out_sc[k] = out_sc[k].float()  # the pruning kernel reads fp32 matrices
            return out_sc, out_st

        i = -1
        issued = -1   # last position handed to the GPU (position-call path)
        kept = {}     # position -> (step_sc, new_full): keeps a position's score matrices alive
        run_ahead = False
        n_pos = max(maxlen)
        tp = self._tick("init", t_start, dev) if self.prof is not None else 0.0
        for i in range(n_pos):
            tp = self._tick("host_other", time.perf_counter(), dev) if self.prof is not None else 0.0
            # ---- scoring (batch_beam_search.py:217-241), same accumulation order
            w = None if fused else torch.zeros(n, V, dtype=xs.dtype, device=dev)
            if plan is None:
                step_sc, new_full = score_full(i, yseq)
                if not fused:
                    for k in self.full_scorers:
                        w += self.weights[k] * step_sc[k]
                tp = self._tick("full_scorers", tp, dev)
            else:
                # ---- everything else of the position: one library call.  Nothing a position
                # computes depends on the host unless a full scorer carries per-hypothesis state
                # (selected on the host from the back-pointers): otherwise position i+1 is issued
                # BEFORE the host waits for position i, the GPU runs positions back to back and
                # the host's bookkeeping trails it.  (An utterance found finished is masked up to
                # two positions late; its extra hypotheses are ignored through `done_np`.)
                while issued < min(i + (1 if run_ahead else 0), n_pos - 1):
                    j = issued + 1
                    sc_j, st_j = score_full(j, plan.ybuf[j & 1][:, : P0 + j])
                    plan.launch([sc_j[k] for k in self.full_scorers], j, P0 + j)
                    kept[j] = (sc_j, st_j)
                    kept.pop(j - 2, None)
                    issued = j
                    if dev_state:  # rows the new slots came from, on the device (queued behind position j)
                        prev_rows = torch.div(plan.t["out_flat"], V, rounding_mode="floor") + slot0_flat
                        for k in dev_state:
                            if st_j[k] is not None:
                                full_states[k] = self.full_scorers[k].batch_select_state(st_j[k], prev_rows)
                    if j == 0:
                        run_ahead = all(v is None or k in dev_state for k, v in st_j.items())
                step_sc, new_full = kept[i]
                cur_len = P0 + i
                pack = plan.wait(i)
                yseq = plan.ybuf[(i + 1) & 1][:, : cur_len + 1]
                tp = self._tick("position_call", tp, dev)
            if plan is None:
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
                        step_sc[k], new_part[k] = d.batch_score_partial(yseq, part_ids, part_states[k],
                                                                        xs[0, : lens[0]])
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
                if dev_state:
                    prev_rows = torch.div(top.reshape(-1), V, rounding_mode="floor") + slot0_flat
                    for k in dev_state:
                        if new_full[k] is not None:
                            full_states[k] = self.full_scorers[k].batch_select_state(new_full[k], prev_rows)
                tp = self._tick("beam_topk_gathers", tp, dev)
                pack = pack_d.cpu().numpy()  # the one device->host read of the step
                tp = self._tick("d2h_read", tp, dev)
            h_val = pack[0]
            if stateful is None:
                stateful = [k for k in self.full_scorers if new_full[k] is not None and k not in dev_state]
            if stateful or not all(batched.values()):
                h_prev = pack[1].astype(np.int64)
                h_tok = pack[2].astype(np.int64)
            for k in stateful:  # a stateless scorer's select_state(None, i) is None: nothing to select
                d, st = self.full_scorers[k], new_full[k]
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
            ends = valid & ((pack[2] == eos_f) | is_last[:, None]) & ~done_np[:, None]
            running = (valid & ~ends).sum(1)
            ended_args = None
            if ends.any():  # the device reads an ended hypothesis needs: its prefix and CTC state
                us, js = np.nonzero(ends)
                slots = us * W + js
                y_rows = yseq.cpu().numpy()[slots]
                if plan is not None:
                    rc, sc_ = plan.ended_states(i)
                    states = [_SlotStates(names, kp, rc, sc_, int(q)) for q in slots]
                else:
                    states = []
                    for u, j in zip(us, js):
                        st = {k: None for k in self.full_scorers}
                        for k, d in self.part_scorers.items():
                            st[k] = (d.slot_state(part_states[k], int(u) * W + int(j)) if batched[k]
                                     else part_states[k][int(j)])
                        states.append(st)
                ended_args = (us, js, pack, y_rows, is_last, states)
            if ended_args is not None:
                ended.add(*ended_args)
            if maxlenratio == 0.0:
                newly = np.array([not done_np[u] and (running[u] == 0 or self._end_detect(
                    ended.best_all[u], ended.best_len[u], i)) for u in range(U)])
            else:
                newly = ~done_np & (running == 0)
            changed = bool(newly.any())
            done_np |= newly
            tp = self._tick("host_bookkeeping", tp, dev)
            if done_np.all():
                break
            if changed:
                if plan is not None:
                    plan.deactivate(done_np, issued + 1)
                else:
                    active = torch.from_numpy(~done_np).to(dev).unsqueeze(1)
                    score = torch.where(active, score, torch.full_like(score, _NEG_INF))
                    if fused:
                        active_u8 = active.view(-1).to(torch.uint8)
        self.last_steps = i + 1
        tp = time.perf_counter()
        out = ended.nbest(dev, xs.dtype)
        if plan is not None and issued > i:  # a position issued ahead still writes the pinned record
            plan.stream.synchronize()
        self._tick("nbest", tp, dev)
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
