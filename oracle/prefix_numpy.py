"""Single-hypothesis CTC prefix scorer in numpy.

TEST INFRASTRUCTURE ONLY (see oracle/__init__).  Restates ``CTCPrefixScore``
(espnet/nets/ctc_prefix_score.py:272-358, relative to /root/reference), the
reference's own second implementation that its equivalence tests compare the
batched scorer with (test/test_batch_beam_search.py:184-188).
"""
from __future__ import annotations

import numpy as np

LOGZERO = -10000000000.0


def initial_state(x: np.ndarray, blank: int) -> np.ndarray:
    """``:289-301``: (T,2) lattice of the empty prefix; column 1 = blank run."""
    T = x.shape[0]
    r = np.full((T, 2), LOGZERO, dtype=np.float32)
    r[:, 1] = np.cumsum(x[:, blank].astype(np.float32), dtype=np.float32)
    # the reference accumulates sequentially in fp32; cumsum does the same
    return r


def score(x: np.ndarray, y, cs, r_prev: np.ndarray, blank: int, eos: int):
    """``__call__`` (``:303-358``).  x (T,V) log-probs, y prefix (with sos),
    cs candidate ids, r_prev (T,2).  Returns (log_psi (K,), r (K,T,2))."""
    T = x.shape[0]
    cs = np.asarray(cs)
    K = len(cs)
    out_len = len(y) - 1
    xs = x[:, cs]
    r = np.empty((T, 2, K), dtype=np.float32)
    if out_len == 0:
        r[0, 0] = xs[0]
        r[0, 1] = LOGZERO
    else:
        r[out_len - 1] = LOGZERO
    r_sum = np.logaddexp(r_prev[:, 0], r_prev[:, 1])
    last = int(y[-1])
    log_phi = np.repeat(r_sum[:, None], K, axis=1).astype(np.float32)
    if out_len > 0:
        log_phi[:, cs == last] = r_prev[:, 1][:, None]
    start = max(out_len, 1)
    log_psi = r[start - 1, 0].copy()
    for t in range(start, T):
        r[t, 0] = np.logaddexp(r[t - 1, 0], log_phi[t - 1]) + xs[t]
        r[t, 1] = np.logaddexp(r[t - 1, 0], r[t - 1, 1]) + x[t, blank]
        log_psi = np.logaddexp(log_psi, log_phi[t - 1] + xs[t])
    log_psi[cs == eos] = r_sum[-1]
    log_psi[cs == blank] = LOGZERO
    return log_psi, np.moveaxis(r, 2, 0)
