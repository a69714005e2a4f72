"""Shared helpers: golden loading and prefix-trace replay."""
import json
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LOGZERO = -10000000000.0


def load_npz(name):
    with np.load(os.path.join(GOLDEN, name)) as z:
        return {k: z[k] for k in z.files}


def load_json(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


PREFIX_CASES = ["pre_m0", "pre_m5", "full_m0", "full_m5", "b1_w3"]


def replay_prefix_trace(g, tag, make_scorer, device="cpu", rtol=1e-6, atol=1e-6,
                        follow_golden=True):
    """Re-drive a recorded CTCPrefixScoreTH trace (tests/golden/make_golden.py
    ``prefix_trace``) through ``make_scorer(lpz, xlens, blank, eos, margin)``
    and compare every recorded tensor.  The scorer object must expose the
    reference's ``__call__``/``index_select_state`` protocol."""
    B, T, V, W, snum, margin, eos, steps = [int(v) for v in g[f"{tag}__cfg"]]
    lpz = torch.tensor(g[f"{tag}__lpz"], device=device)
    xlens = torch.tensor(g[f"{tag}__xlens"])
    scorer = make_scorer(lpz, xlens, 0, eos, margin)
    # the constructor pads the caller's tensor in place (ctc_prefix_score.py:47-50)
    np.testing.assert_array_equal(lpz.cpu().numpy(), g[f"{tag}__lpz_padded"])
    state = None
    for i in range(steps):
        p = f"{tag}__s{i}__"
        y = g[p + "y"].tolist()
        part = torch.tensor(g[p + "part_ids"], device=device) if snum else None
        att = torch.tensor(g[p + "att_w"], device=device) if margin > 0 else None
        sc, state5 = scorer(y, state, part, att)
        r, log_psi, f_min, f_max, idmap = state5
        assert [int(f_min), int(f_max)] == g[p + "fminmax"].tolist(), (tag, i)
        close(sc, g[p + "scores"], rtol, atol, f"{tag} step {i} scores")
        close(log_psi, g[p + "log_psi"], rtol, atol, f"{tag} step {i} log_psi")
        close(r, g[p + "r"], rtol, atol, f"{tag} step {i} r")
        if snum:
            np.testing.assert_array_equal(idmap.cpu().numpy(), g[p + "idmap"])
        else:
            assert idmap is None
        best = torch.tensor(g[p + "best_ids"], device=device)
        state = scorer.index_select_state(state5, best)
        close(state[0], g[p + "sel_r"], rtol, atol, f"{tag} step {i} sel_r")
        close(state[1], g[p + "sel_s"], rtol, atol, f"{tag} step {i} sel_s")
        assert (int(state[2]), int(state[3])) == (int(f_min), int(f_max))


def close(a, b, rtol, atol, msg=""):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol, err_msg=msg)
