"""End detection of the label-synchronous beam search.

Mirrors /root/reference/espnet/nets/e2e_asr_common.py:18-49 (Eq. 50 of Watanabe
et al. 2017, "Hybrid CTC/Attention Architecture for End-to-End Speech
Recognition"): decoding stops once, for each of the last ``M`` hypothesis
lengths, the best hypothesis that ended at that length is more than ``|D_end|``
below the best ended hypothesis overall.
"""
from __future__ import annotations

import math

D_END = math.log(1 * math.exp(-10))


def end_detect(ended_hyps, i, M=3, D_end=D_END):
    """:param ended_hyps: dicts with ``"score"`` and ``"yseq"`` (a sized sequence)
    :param int i: current output position
    """
    if len(ended_hyps) == 0:
        return False
    best = max(float(h["score"]) for h in ended_hyps)
    count = 0
    for m in range(M):
        same = [float(h["score"]) for h in ended_hyps if len(h["yseq"]) == i - m]
        if same and max(same) - best < D_end:
            count += 1
    return count == M
