"""Drop-in for ``espnet.nets.scorers.ctc.CTCPrefixScorer`` (B200 CUDA).

Mirrors /root/reference/espnet/nets/scorers/ctc.py:10-157.  It stays a
``BatchPartialScorerInterface`` (not an nn.Module), so the unmodified
``BeamSearch`` / ``BatchBeamSearch`` file it under ``part_scorers``
(beam_search.py:81-84).  Both the batched methods and the per-hypothesis ones
(which the reference serves with its numpy scorer on the CPU, ``:26-38,65-85``)
run on the CUDA prefix-scoring kernels.
"""
from __future__ import annotations

import torch

from ...._lib import check, lib, on_tensor_device, ptr, stream_ptr
from ..ctc_prefix_score import CTCPrefixScoreTH
from ..scorer_interface import BatchPartialScorerInterface


class CTCPrefixScorer(BatchPartialScorerInterface):
    """Decoder interface wrapper for CTCPrefixScoreTH."""

    def __init__(self, ctc: torch.nn.Module, eos: int):
        """Initialize class.

        Args:
            ctc (torch.nn.Module): The CTC implementation (anything with
                ``log_softmax(hs_pad)``, e.g. :class:`espnet1_b200.espnet2.asr.ctc.CTC`)
            eos (int): The end-of-sequence id.
        """
        self.ctc = ctc
        self.eos = eos
        self.impl = None

    def _make_impl(self, x: torch.Tensor):
        logp = self.ctc.log_softmax(x.unsqueeze(0)).detach()  # batch_size = 1 (:96)
        xlen = torch.tensor([logp.size(1)])
        self.impl = CTCPrefixScoreTH(logp, xlen, 0, self.eos)

    # ---- per-hypothesis API (BeamSearch) --------------------------------
    @on_tensor_device
    def init_state(self, x: torch.Tensor):
        """Initial state for non-batched decoding: (prev_score, r (T,2))."""
        self._make_impl(x)
        # the (T,2) lattice of the empty prefix is the scorer's initial r_prev
        r0 = torch.empty(self.impl.input_length, 2, 1, dtype=torch.float32, device=x.device)
        check(lib().b200ctc_prefix_init_state(ptr(self.impl._lp), 1, self.impl.input_length,
                                              self.impl.odim, 1, 0, self.impl.logzero, ptr(r0),
                                              stream_ptr(x.device)), "prefix_init_state")
        return 0, r0[:, :, 0]

    def select_state(self, state, i, new_id=None):
        """Select state with relative ids in the main beam search (:40-63)."""
        if type(state) == tuple:
            if len(state) == 2:  # per-hypothesis path
                sc, st = state
                return sc[i], st[i]
            else:  # batched path (need new_id > 0)
                r, log_psi, f_min, f_max, scoring_idmap = state
                s = log_psi[i, new_id].expand(log_psi.size(1))
                if scoring_idmap is not None:
                    return r[:, :, i, scoring_idmap[i, new_id]], s, f_min, f_max
                else:
                    return r[:, :, i, new_id], s, f_min, f_max
        return None if state is None else state[i]

    def score_partial(self, y, ids, state, x):
        """Score new tokens for ONE hypothesis (:65-85).

        Returns (scores (len(ids),), (presub_scores, r (len(ids), T, 2)))."""
        prev_score, r_prev = state
        ids_d = ids.to(device=x.device, dtype=torch.int64).view(1, -1)
        _, st = self.impl(y.view(1, -1), (r_prev.unsqueeze(2), 0.0, 0, 1), ids_d)
        r, log_psi = st[0], st[1]
        presub = log_psi[0, ids_d[0]]
        tscore = (presub - prev_score).to(device=x.device, dtype=x.dtype)
        return tscore, (presub, r[:, :, 0, :].permute(2, 0, 1))

    # ---- batched API (BatchBeamSearch) ----------------------------------
    def batch_init_state(self, x: torch.Tensor):
        """Get an initial state for decoding (:87-99)."""
        self._make_impl(x)
        return None

    def batch_score_partial(self, y, ids, state, x):
        """Score new tokens for all running hypotheses (:101-126)."""
        batch_state = (
            (
                torch.stack([s[0] for s in state], dim=2),
                torch.stack([s[1] for s in state]),
                state[0][2],
                state[0][3],
            )
            if state[0] is not None
            else None
        )
        return self.impl(y, batch_state, ids)

    # ---- batched-state API (used by ..batch_beam_search) ----
    # The state never becomes n_beam Python tuples: it stays the (r (T,2,n_bh), s (n_bh,V),
    # f_min, f_max) tuple CTCPrefixScoreTH itself consumes (ctc_prefix_score.py:68-96), for
    # U utterances x beam_size slots, and pruning is ONE index_select_state gather
    # (ctc_prefix_score.py:189-220) instead of n_beam select_state slices (:40-63).
    def batch_init_state_multi(self, xs: torch.Tensor, xlens):
        """Initial state for U utterances ``xs (U, Tmax, D)`` decoded together."""
        logp = self.ctc.log_softmax(xs).detach()
        self.impl = CTCPrefixScoreTH(logp, torch.as_tensor(xlens), 0, self.eos)
        return None

    def batch_score_partial_batched(self, y, ids, state):
        """``batch_score_partial`` on the batched state; returns (scores (n_bh,V), state5)."""
        return self.impl(y, state, ids)

    def batch_select_state(self, state, best_ids):
        """Prune with flat ids ``best_ids (U, beam)`` over (slots x vocabulary)."""
        return self.impl.index_select_state(state, best_ids)

    def slot_state(self, state, i):
        """The per-hypothesis view ``select_state`` would have returned for slot ``i``."""
        r, s, f_min, f_max = state
        return r[:, :, i], s[i], f_min, f_max

    # ---- streaming (:128-157) -------------------------------------------
    def extend_prob(self, x: torch.Tensor):
        """Extend probs for streaming decoding (Eq. 14 of arXiv:2006.14941)."""
        logp = self.ctc.log_softmax(x.unsqueeze(0)).detach()
        self.impl.extend_prob(logp)

    def extend_state(self, state):
        """Extend states for streaming decoding."""
        return [self.impl.extend_state(s) for s in state]
