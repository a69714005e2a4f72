"""Scorer interface used by the beam-search callers.

When the real ESPnet is importable its classes are re-exported, so that
``isinstance(scorer, PartialScorerInterface)`` checks inside the unmodified
``espnet.nets.beam_search.BeamSearch`` (beam_search.py:76-84) keep working.
Otherwise minimal stand-ins with the same method contract
(/root/reference/espnet/nets/scorer_interface.py:9-186) are defined.
"""
from __future__ import annotations

try:  # pragma: no cover - depends on the environment
    from espnet.nets.scorer_interface import (  # type: ignore
        BatchPartialScorerInterface,
        BatchScorerInterface,
        PartialScorerInterface,
        ScorerInterface,
    )
except Exception:  # ESPnet itself is not installed: same contract, local definition

    class ScorerInterface:
        def init_state(self, x):
            return None

        def select_state(self, state, i, new_id=None):
            return None if state is None else state[i]

        def score(self, y, state, x):
            raise NotImplementedError

        def final_score(self, state):
            return 0.0

    class BatchScorerInterface(ScorerInterface):
        def batch_init_state(self, x):
            return self.init_state(x)

        def batch_score(self, ys, states, xs):
            raise NotImplementedError

    class PartialScorerInterface(ScorerInterface):
        def score_partial(self, y, next_tokens, state, x):
            raise NotImplementedError

    class BatchPartialScorerInterface(BatchScorerInterface, PartialScorerInterface):
        def batch_score_partial(self, ys, next_tokens, states, xs):
            raise NotImplementedError
