"""Batched CTC prefix scoring restated with torch CPU ops.

TEST INFRASTRUCTURE / CPU BASELINE ONLY (see oracle/__init__).  Restates
``CTCPrefixScoreTH`` (espnet/nets/ctc_prefix_score.py:11-269, citations
relative to /root/reference): Algorithm 2 of Watanabe et al. 2017 vectorised
over (hypotheses x candidate tokens) as in Seki et al. 2019.  The Python loops
of the reference over hypotheses are expressed with index tensors here; the
floating-point operations (pairwise ``logsumexp`` + add per frame, one
``logsumexp`` over the window for psi) are the same ones in the same order, so
on CPU the results agree with the reference to the last bit (pinned by
tests/test_oracle_golden.py).
"""
from __future__ import annotations

import torch

LOGZERO = -10000000000.0  # ctc_prefix_score.py:33


class PrefixOracle:
    """State container mirroring the reference constructor (``:22-66``)."""

    def __init__(self, x, xlens, blank, eos, margin=0):
        self.logzero = LOGZERO
        self.blank, self.eos, self.margin = blank, eos, margin
        self.batch, self.input_length, self.odim = x.shape
        self.dtype, self.device = x.dtype, x.device
        xlens = torch.as_tensor(xlens)
        # frames past each utterance's end: every label impossible, blank certain
        # (``:47-50``) -- done IN PLACE on the caller's tensor like the reference.
        for b in range(self.batch):
            n = int(xlens[b])
            if n < self.input_length:
                x[b, n:, :] = self.logzero
                x[b, n:, blank] = 0
        self.lp = x.transpose(0, 1)  # (T, B, V) view
        self.lp_blank = self.lp[:, :, blank]  # (T, B)
        self.end_frames = xlens - 1
        self.scoring_num = 0
        if margin > 0:
            self.frame_ids = torch.arange(self.input_length, dtype=self.dtype)

    # ------------------------------------------------------------------
    def initial_r(self, n_hyps):
        """``:83-91``: r^n = logzero, r^b = running sum of blank log-probs."""
        T, B = self.input_length, self.batch
        r = torch.full((T, 2, B, n_hyps), self.logzero, dtype=self.dtype)
        r[:, 1] = torch.cumsum(self.lp_blank, 0).unsqueeze(2)
        return r.view(T, 2, B * n_hyps)

    def score(self, y, state, scoring_ids=None, att_w=None):
        """``__call__`` (``:68-187``).  Returns (scores, new_state5)."""
        T, B, V = self.input_length, self.batch, self.odim
        out_len = len(y[0]) - 1
        last = torch.as_tensor([int(yi[-1]) for yi in y], dtype=torch.long)
        n_bh = last.numel()
        n_hyps = n_bh // B
        self.scoring_num = scoring_ids.size(-1) if scoring_ids is not None else 0
        if state is None:
            r_prev, s_prev, f_min_prev, f_max_prev = self.initial_r(n_hyps), 0.0, 0, 1
        else:
            r_prev, s_prev, f_min_prev, f_max_prev = state
        utt = torch.arange(n_bh) // n_hyps  # utterance of each hypothesis
        hyp = torch.arange(n_bh)

        if self.scoring_num > 0:
            snum = self.scoring_num
            idmap = torch.full((n_bh, V), -1, dtype=torch.long)
            idmap[hyp.view(-1, 1), scoring_ids] = torch.arange(snum)
            x_lab = self.lp[:, utt.view(-1, 1), scoring_ids]  # (T, n_bh, snum)
        else:
            snum, idmap, scoring_ids = V, None, None
            x_lab = self.lp[:, utt, :]  # (T, n_bh, V)
        x_blk = self.lp_blank[:, utt].unsqueeze(2).expand(T, n_bh, snum)

        r = torch.full((T, 2, n_bh, snum), self.logzero, dtype=self.dtype)
        if out_len == 0:
            r[0, 0] = x_lab[0]
        r_sum = torch.logsumexp(r_prev, 1)  # (T, n_bh)
        log_phi = r_sum.unsqueeze(2).repeat(1, 1, snum)
        # a candidate equal to the hypothesis' last label may only follow a blank
        if idmap is not None:
            pos = idmap[hyp, last]
            ok = pos >= 0
            log_phi[:, hyp[ok], pos[ok]] = r_prev[:, 1, hyp[ok]]
        else:
            log_phi[:, hyp, last] = r_prev[:, 1, hyp]

        if att_w is not None and self.margin > 0:  # ``:144-149`` one window for all
            f_arg = torch.matmul(att_w, self.frame_ids)
            f_min = max(int(f_arg.min()), f_min_prev)
            f_max = max(int(f_arg.max()), f_max_prev)
            start = min(f_max_prev, max(f_min - self.margin, out_len, 1))
            end = min(f_max + self.margin, T)
        else:
            f_min = f_max = 0
            start, end = max(out_len, 1), T

        for t in range(start, end):  # ``:156-161``
            rn, rb = r[t - 1, 0], r[t - 1, 1]
            new_n = torch.logsumexp(torch.stack([rn, log_phi[t - 1]]), 0) + x_lab[t]
            new_b = torch.logsumexp(torch.stack([rn, rb]), 0) + x_blk[t]
            r[t, 0], r[t, 1] = new_n, new_b

        phi_shift = torch.cat((log_phi[0:1], log_phi[:-1]), 0) + x_lab  # ``:164``
        psi_cand = torch.logsumexp(
            torch.cat((phi_shift[start:end], r[start - 1, 0].unsqueeze(0)), 0), 0
        )
        if idmap is not None:
            log_psi = torch.full((n_bh, V), self.logzero, dtype=self.dtype)
            log_psi[hyp.view(-1, 1), scoring_ids] = psi_cand
        else:
            log_psi = psi_cand
        log_psi[hyp, self.eos] = r_sum[self.end_frames[utt], hyp]  # ``:181-182``
        log_psi[:, self.blank] = self.logzero  # ``:185``
        return log_psi - s_prev, (r, log_psi, f_min, f_max, idmap)

    def select(self, state, best_ids):
        """``index_select_state`` (``:189-220``)."""
        r, s, f_min, f_max, idmap = state
        B, V = self.batch, self.odim
        n_bh = s.shape[0]
        n_hyps = n_bh // B
        base = torch.arange(B).view(-1, 1)
        flat = (best_ids + base * (n_hyps * V)).view(-1)
        s_new = s.reshape(-1)[flat].view(-1, 1).repeat(1, V)
        hyp = (best_ids // V + base * n_hyps).view(-1)
        lab = torch.fmod(best_ids, V).view(-1)
        if idmap is not None:
            col = idmap[hyp, lab].clone()
            col[col == -1] = 0
        else:
            col = lab
        r_new = r[:, :, hyp, col]
        return r_new, s_new, f_min, f_max

    def extend_prob(self, x):
        """``extend_prob`` (``:222-242``), batch == 1 streaming."""
        if self.input_length < x.shape[1]:
            old = self.lp
            new = x.transpose(0, 1).clone()
            new[: old.shape[0]] = old
            self.lp = new
            self.lp_blank = self.lp[:, :, self.blank]
            self.input_length = x.size(1)
            self.end_frames = torch.as_tensor([x.size(1)]) - 1

    def extend_state(self, state):
        """``extend_state`` (``:244-269``): continue the blank running sum."""
        if state is None:
            return None
        r_prev, s_prev, f_min_prev, f_max_prev = state
        r_new = torch.full((self.input_length, 2), self.logzero, dtype=self.dtype)
        start = max(r_prev.shape[0], 1)
        r_new[0:start] = r_prev
        for t in range(start, self.input_length):
            r_new[t, 1] = r_new[t - 1, 1] + self.lp[t, 0, self.blank]
        return r_new, s_prev, f_min_prev, f_max_prev
