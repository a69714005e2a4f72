"""``Hypothesis`` record of the beam searches
(/root/reference/espnet/nets/beam_search.py:20-35): same fields, same
``asdict()``.  When the real ESPnet is importable its class is re-exported so
that results of the two implementations compare as the same type."""
from __future__ import annotations

from typing import Any, Dict, NamedTuple, Union

import torch

try:  # pragma: no cover - depends on the environment
    from espnet.nets.beam_search import Hypothesis  # type: ignore
except Exception:

    class Hypothesis(NamedTuple):
        """Hypothesis data type."""

        yseq: torch.Tensor
        score: Union[float, torch.Tensor] = 0
        scores: Dict[str, Union[float, torch.Tensor]] = dict()
        states: Dict[str, Any] = dict()

        def asdict(self) -> dict:
            """Convert data to JSON-friendly dict."""
            return self._replace(
                yseq=self.yseq.tolist(),
                score=float(self.score),
                scores={k: float(v) for k, v in self.scores.items()},
            )._asdict()
