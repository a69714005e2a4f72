"""Drop-in for ``espnet.nets.scorers.length_bonus.LengthBonus``
(/root/reference/espnet/nets/scorers/length_bonus.py:9-61): a constant 1.0 for
every next token; stateless."""
from __future__ import annotations

import torch

from ..scorer_interface import BatchScorerInterface


class LengthBonus(BatchScorerInterface):
    """Length bonus in beam search."""

    def __init__(self, n_vocab: int):
        self.n = n_vocab

    def score(self, y, state, x):
        return torch.tensor([1.0], device=x.device, dtype=x.dtype).expand(self.n), None

    def batch_score(self, ys, states, xs):
        return torch.tensor([1.0], device=xs.device, dtype=xs.dtype).expand(ys.shape[0], self.n), None
