"""Scorer-interface adapter around the oracle prefix scorers.

TEST INFRASTRUCTURE ONLY (see oracle/__init__).  Restates ``CTCPrefixScorer``
(espnet/nets/scorers/ctc.py:10-157, relative to /root/reference) on top of
``oracle.prefix_torch.PrefixOracle`` and ``oracle.prefix_numpy`` so the beam
harness can be pinned against the reference's golden n-best lists on CPU.
"""
from __future__ import annotations

import numpy as np
import torch

from . import prefix_numpy
from .prefix_torch import PrefixOracle


class OracleCTCPrefixScorer:
    def __init__(self, ctc, eos):
        self.ctc, self.eos, self.impl, self._logp = ctc, eos, None, None

    # -- non-batched path: numpy scorer (scorers/ctc.py:26-38,65-85) -------
    def init_state(self, x):
        self._logp = self.ctc.log_softmax(x.unsqueeze(0)).detach().squeeze(0).cpu().numpy()
        return 0, prefix_numpy.initial_state(self._logp, 0)

    def score_partial(self, y, ids, state, x):
        prev_score, r_prev = state
        psi, new_r = prefix_numpy.score(self._logp, y.cpu().numpy(), ids.cpu().numpy(),
                                        r_prev, 0, self.eos)
        return torch.as_tensor(psi - prev_score, device=x.device, dtype=x.dtype), (psi, new_r)

    def final_score(self, state):
        return 0.0

    # -- batched path (scorers/ctc.py:87-126) ------------------------------
    def batch_init_state(self, x):
        logp = self.ctc.log_softmax(x.unsqueeze(0))
        self.impl = PrefixOracle(logp, torch.tensor([logp.size(1)]), 0, self.eos)
        return None

    def batch_score_partial(self, y, ids, state, x):
        if state[0] is not None:
            batch_state = (torch.stack([s[0] for s in state], dim=2),
                           torch.stack([s[1] for s in state]), state[0][2], state[0][3])
        else:
            batch_state = None
        return self.impl.score(y, batch_state, ids)

    def select_state(self, state, i, new_id=None):
        """scorers/ctc.py:40-63."""
        if type(state) == tuple:
            if len(state) == 2:
                return state[0][i], state[1][i]
            r, log_psi, f_min, f_max, idmap = state
            s = log_psi[i, new_id].expand(log_psi.size(1))
            col = idmap[i, new_id] if idmap is not None else new_id
            return r[:, :, i, col], s, f_min, f_max
        return None if state is None else state[i]

    # -- batched-state API of the product scorer (espnet1_b200 .../scorers/ctc.py), so the
    #    host logic of the device-resident beam search can be checked on CPU --------------
    def batch_init_state_multi(self, xs, xlens):
        self.impl = PrefixOracle(self.ctc.log_softmax(xs), torch.as_tensor(xlens), 0, self.eos)
        return None

    def batch_score_partial_batched(self, y, ids, state):
        return self.impl.score(y, state, ids)

    def batch_select_state(self, state, best_ids):
        return self.impl.select(state, best_ids)

    def slot_state(self, state, i):
        r, s, f_min, f_max = state
        return r[:, :, i], s[i], f_min, f_max

    def extend_prob(self, x):
        self.impl.extend_prob(self.ctc.log_softmax(x.unsqueeze(0)))

    def extend_state(self, state):
        return [self.impl.extend_state(s) for s in state]


class TableScorer:
    """Deterministic stand-in for an attention decoder: next-token log-probs
    depend on the last token only (a bigram table).  Mirrors the role of the
    tiny random decoders in test/test_beam_search.py:90-119."""

    def __init__(self, table: torch.Tensor):
        self.table = table  # (V, V) log-probs

    def init_state(self, x):
        return None

    def score(self, y, state, x):
        return self.table[int(y[-1])].to(dtype=x.dtype, device=x.device), None

    def batch_score(self, ys, states, xs):
        return self.table.to(dtype=xs.dtype, device=xs.device)[ys[:, -1]], None

    def final_score(self, state):
        return 0.0


class LengthBonusPort:
    """espnet/nets/scorers/length_bonus.py:9-61: constant 1.0 per token."""

    def __init__(self, n):
        self.n = n

    def init_state(self, x):
        return None

    def score(self, y, state, x):
        return torch.tensor([1.0], device=x.device, dtype=x.dtype).expand(self.n), None

    def batch_score(self, ys, states, xs):
        return torch.tensor([1.0], device=xs.device, dtype=xs.dtype).expand(ys.shape[0], self.n), None

    def final_score(self, state):
        return 0.0
