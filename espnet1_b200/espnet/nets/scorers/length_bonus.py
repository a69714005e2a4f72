"""Drop-in for ``espnet.nets.scorers.length_bonus.LengthBonus``
(/root/reference/espnet/nets/scorers/length_bonus.py:9-61): a constant 1.0 for
every next token; stateless."""
from __future__ import annotations

import torch

from ..scorer_interface import BatchScorerInterface


class LengthBonus(BatchScorerInterface):
    """Length bonus in beam search."""

    ignores_x = True  # BatchBeamSearch.forward_batch: no need to hide padding frames from it

    def __init__(self, n_vocab: int):
        self.n = n_vocab
        self._one = {}  # (device, dtype) -> the constant, made once (the reference uploads it every call)

    def _const(self, ref: torch.Tensor) -> torch.Tensor:
        key = (ref.device, ref.dtype)
        if key not in self._one:
            self._one[key] = torch.tensor([1.0], device=ref.device, dtype=ref.dtype)
        return self._one[key]

    def score(self, y, state, x):
        return self._const(x).expand(self.n), None

    def batch_score(self, ys, states, xs):
        return self._const(xs).expand(ys.shape[0], self.n), None
