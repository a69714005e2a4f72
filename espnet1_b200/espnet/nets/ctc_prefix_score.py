"""Drop-in for ``espnet.nets.ctc_prefix_score.CTCPrefixScoreTH`` (B200 CUDA).

Mirrors /root/reference/espnet/nets/ctc_prefix_score.py:11-269: constructor
``(x, xlens, blank, eos, margin=0)`` that pads the caller's ``x`` in place,
``__call__(y, state, scoring_ids=None, att_w=None) -> (scores, state5)``,
``index_select_state``, ``extend_prob``, ``extend_state`` and the public
attributes callers read.  The per-frame Python loop of the reference
(``:156-161``) is ONE kernel launch here; there is no CPU fallback.
"""
from __future__ import annotations

import torch

from ..._lib import check, lib, on_tensor_device, ptr, require_cuda, stream_ptr


class CTCPrefixScoreTH(object):
    """Batch processing of CTCPrefixScore (Watanabe et al. 2017, Algorithm 2;
    Seki et al. 2019) on the GPU."""

    @on_tensor_device
    def __init__(self, x, xlens, blank, eos, margin=0, per_utt_window=False):
        """Construct CTC prefix scorer

        :param torch.Tensor x: input label posterior sequences (B, T, O)
        :param torch.Tensor xlens: input lengths (B,)
        :param int blank: blank label id
        :param int eos: end-of-sequence id
        :param int margin: margin parameter for windowing (0 means no windowing)
        :param bool per_utt_window: EXTENSION (not in the reference): with
            ``margin > 0`` take the attention window per utterance instead of one window over
            the whole batch (ctc_prefix_score.py:144-153) - each utterance is then scored
            exactly as if it were decoded alone.  The window never leaves the device (no
            read-back); the state's ``f_min`` slot then holds a (B, 2) int32 device tensor
            ``{f_min, f_max}`` and the ``f_max`` slot a view of its second column.
        """
        require_cuda(x, "CTCPrefixScoreTH x")
        self.logzero = -10000000000.0
        self.blank = blank
        self.eos = eos
        self.batch = x.size(0)
        self.input_length = x.size(1)
        self.odim = x.size(2)
        self.dtype = x.dtype
        self.device = x.device
        xl = torch.as_tensor(xlens)
        if xl.is_cuda:
            self._xlens_dev = xl.to(device=self.device, dtype=torch.int64).contiguous()
        else:  # from pinned memory, not blocking: a pageable upload would wait for the GPU to drain
            self._xlens_dev = xl.to(torch.int64).pin_memory().to(self.device, non_blocking=True)
        self.end_frames = xl - 1  # same object kind as the reference (:55)
        self._end_frames_dev = self._xlens_dev - 1
        self._set_lp(x, pad=True)
        self.margin = margin
        self.per_utt_window = bool(per_utt_window)
        if margin > 0:
            self.frame_ids = torch.arange(self.input_length, dtype=self.dtype, device=self.device)
        self.idx_bh = None
        self.idx_b = torch.arange(self.batch, device=self.device)
        self.idx_bo = (self.idx_b * self.odim).unsqueeze(1)
        self.scoring_num = 0

    # -- log-prob storage -------------------------------------------------
    @on_tensor_device
    def _set_lp(self, x, pad):
        B, T, V = x.shape
        inplace = x.dtype == torch.float32 and x.is_contiguous()
        lp = x if inplace else x.to(torch.float32).contiguous()
        if pad:
            check(lib().b200ctc_prefix_pad(ptr(lp), ptr(self._xlens_dev), B, T, V, self.blank,
                                           self.logzero, stream_ptr(self.device)), "prefix_pad")
            if not inplace:  # keep the reference's side effect on the caller's tensor (:47-50)
                x.copy_(lp.to(x.dtype))
        self._lp = lp
        self._lp_store = None  # streaming (extend_prob): (1, Tcap, V) buffer that _lp is a view of
        # (B, V, ld) time-contiguous copy, made on the first pre-beam step; frames [0, _lp_t_T)
        # of it are valid.  ld > T once extend_prob has given it headroom.
        self._lp_t = None
        self._lp_t_T = 0

    @on_tensor_device
    def _time_contiguous(self):
        B, T, V = self._lp.shape
        if self._lp_t is None or self._lp_t.shape[2] < T:
            old, t_old = self._lp_t, self._lp_t_T
            # streaming grows by chunks: geometric headroom makes the appends O(new frames)
            # (a time stride that is a multiple of 16 frames lets the scoring kernel fetch whole
            # groups of frames with 16-byte loads up to the end of a row)
            ld = T if self._lp_store is None else max(T, int(1.5 * T) + 32)
            ld = -(-ld // 16) * 16
            self._lp_t = torch.empty(B, V, ld, dtype=torch.float32, device=self.device)
            if old is not None and t_old > 0:
                self._lp_t[:, :, :t_old] = old[:, :, :t_old]
            else:
                self._lp_t_T = 0
        if self._lp_t_T < T:  # transpose only the frames that are new
            check(lib().b200ctc_prefix_transpose(ptr(self._lp), ptr(self._lp_t), B, T, V,
                                                 self._lp_t_T, self._lp_t.shape[2],
                                                 stream_ptr(self.device)), "prefix_transpose")
            self._lp_t_T = T
        return self._lp_t

    @property
    def x(self):
        """(2, T, B, O) stack [labels, blank broadcast] as the reference exposes (:52-54);
        materialised only when somebody asks for it."""
        xn = self._lp.transpose(0, 1)
        xb = xn[:, :, self.blank].unsqueeze(2).expand(-1, -1, self.odim)
        return torch.stack([xn, xb]).to(self.dtype)

    # -- scoring ----------------------------------------------------------
    @on_tensor_device
    def __call__(self, y, state, scoring_ids=None, att_w=None):
        """Compute CTC prefix scores for next labels

        :param list y: prefix label sequences
        :param tuple state: previous CTC state
        :param torch.Tensor scoring_ids: candidate ids per hypothesis (BW, snum)
        :param torch.Tensor att_w: attention weights to decide CTC window
        :return new_state, ctc_local_scores (BW, O)
        """
        dev, T, V = self.device, self.input_length, self.odim
        if isinstance(y, torch.Tensor):
            output_length = y.size(1) - 1
            last_ids = y[:, -1].to(device=dev, dtype=torch.int64).contiguous()
        else:
            output_length = len(y[0]) - 1
            last_ids = torch.as_tensor([int(yi[-1]) for yi in y], dtype=torch.int64).to(dev)
        n_bh = last_ids.numel()
        n_hyps = n_bh // self.batch
        self.scoring_num = scoring_ids.size(-1) if scoring_ids is not None else 0
        st = stream_ptr(dev)
        if state is None:
            r_prev = torch.empty(T, 2, n_bh, dtype=torch.float32, device=dev)
            check(lib().b200ctc_prefix_init_state(ptr(self._lp), self.batch, T, V, n_hyps,
                                                  self.blank, self.logzero, ptr(r_prev), st),
                  "prefix_init_state")
            s_prev, f_min_prev, f_max_prev = None, 0, 1
        else:
            r_prev, s_prev, f_min_prev, f_max_prev = state
            r_prev = r_prev.to(torch.float32).contiguous()
            if isinstance(s_prev, torch.Tensor):
                s_prev = s_prev.to(torch.float32).expand(n_bh, V).contiguous()
            else:
                s_prev = None if float(s_prev) == 0.0 else torch.full(
                    (n_bh, V), float(s_prev), dtype=torch.float32, device=dev)

        if self.scoring_num > 0:
            snum = self.scoring_num
            ids = scoring_ids.to(device=dev, dtype=torch.int64).contiguous()
            idmap = torch.empty(n_bh, V, dtype=torch.int64, device=dev)
        else:
            snum, ids, idmap = V, None, None

        win = None
        if att_w is not None and self.margin > 0 and self.per_utt_window:
            # per-utterance windows, computed and consumed on the device
            aw = att_w.to(torch.float32).contiguous()
            if isinstance(f_min_prev, torch.Tensor):
                fprev = f_min_prev.to(device=dev, dtype=torch.int32).contiguous()
            elif state is None:
                fprev = None
            else:
                fprev = torch.tensor([[int(f_min_prev), int(f_max_prev)]] * self.batch,
                                     dtype=torch.int32, device=dev)
            fnew = torch.empty(self.batch, 2, dtype=torch.int32, device=dev)
            win = torch.empty(self.batch, 2, dtype=torch.int32, device=dev)
            check(lib().b200ctc_prefix_window_utt(ptr(aw), self.batch, n_hyps, T, self.margin,
                                                  output_length, ptr(fprev), ptr(fnew), ptr(win), st),
                  "prefix_window_utt")
            f_min, f_max = fnew, fnew[:, 1]
            start, end = 1, T  # ignored: the kernel reads `win`
        # window (:144-153); one global window for the whole batch like the reference
        elif att_w is not None and self.margin > 0:
            mm = torch.empty(2, dtype=torch.float32, device=dev)
            aw = att_w.to(torch.float32).contiguous()
            check(lib().b200ctc_prefix_window(ptr(aw), n_bh, T, ptr(mm), st), "prefix_window")
            # the reference syncs here too (:146-147); a cached pinned landing buffer + a stream
            # wait instead of .cpu() (pageable allocation + blocking copy per step)
            if getattr(self, "_mm_host", None) is None:
                self._mm_host = torch.empty(2, dtype=torch.float32, pin_memory=True)
            self._mm_host.copy_(mm, non_blocking=True)
            torch.cuda.current_stream(dev).synchronize()
            lo, hi = self._mm_host.tolist()
            f_min = max(int(lo), f_min_prev)
            f_max = max(int(hi), f_max_prev)
            start = min(f_max_prev, max(f_min - self.margin, output_length, 1))
            end = min(f_max + self.margin, T)
        else:
            f_min = f_max = 0
            start = max(output_length, 1)
            end = T

        r = torch.empty(T, 2, n_bh, snum, dtype=torch.float32, device=dev)
        log_psi = torch.empty(n_bh, V, dtype=torch.float32, device=dev)
        scores = torch.empty(n_bh, V, dtype=torch.float32, device=dev)
        rsum = torch.empty(T, n_bh, dtype=torch.float32, device=dev)
        lp_t = self._time_contiguous() if ids is not None else None
        lp_t_ld = lp_t.shape[2] if lp_t is not None else T
        check(lib().b200ctc_prefix_step(ptr(self._lp), ptr(lp_t), ptr(r_prev), ptr(s_prev), ptr(last_ids),
                                        ptr(ids), ptr(self._end_frames_dev), self.batch, n_hyps,
                                        T, V, snum, output_length, start, end, self.blank,
                                        self.eos, self.logzero, ptr(r), ptr(log_psi), ptr(scores),
                                        ptr(idmap), ptr(rsum), lp_t_ld, ptr(win), st), "prefix_step")
        if self.dtype != torch.float32:
            scores, r, log_psi = scores.to(self.dtype), r.to(self.dtype), log_psi.to(self.dtype)
        return scores, (r, log_psi, f_min, f_max, idmap)

    @on_tensor_device
    def index_select_state(self, state, best_ids):
        """Select CTC states according to best ids

        :param state    : CTC state
        :param best_ids : index numbers selected by beam pruning (B, W)
        :return selected_state
        """
        r, s, f_min, f_max, scoring_idmap = state
        dev, V = self.device, self.odim
        T, _, n_bh, snum = r.shape
        n_hyps = n_bh // self.batch
        best = best_ids.to(device=dev, dtype=torch.int64).contiguous()
        W = best.size(1)
        rc = r.to(torch.float32).contiguous()
        sc = s.to(torch.float32).contiguous()
        r_new = torch.empty(T, 2, self.batch * W, dtype=torch.float32, device=dev)
        s_new = torch.empty(self.batch * W, V, dtype=torch.float32, device=dev)
        check(lib().b200ctc_prefix_select(ptr(rc), ptr(sc), ptr(scoring_idmap), ptr(best),
                                          self.batch, n_hyps, W, T, V, snum, ptr(r_new),
                                          ptr(s_new), stream_ptr(dev)), "prefix_select")
        if self.dtype != torch.float32:
            r_new, s_new = r_new.to(self.dtype), s_new.to(self.dtype)
        return r_new, s_new, f_min, f_max

    # -- streaming (:222-269), batch == 1 ---------------------------------
    @on_tensor_device
    def extend_prob(self, x):
        """Extend CTC prob.

        :param torch.Tensor x: input label posterior sequences (B, T, O)
        """
        if self.input_length < x.shape[1]:
            old = self._lp
            T_old, T_new = old.shape[1], x.shape[1]
            if self.batch == 1:
                # streaming (the reference uses extend_prob with batch 1, scorers/ctc.py:128-146):
                # both copies live in buffers with headroom and only the NEW frames are written -
                # the already-scored prefix (:238) stays where it is, (B,T,V) and (B,V,T) alike
                if self._lp_store is None or self._lp_store.shape[1] < T_new:
                    cap = max(T_new, int(1.5 * T_new) + 32)
                    store = torch.empty(1, cap, self.odim, dtype=torch.float32, device=self.device)
                    store[:, :T_old] = old
                    self._lp_store = store
                self._lp_store[:, T_old:T_new] = x[:, T_old:T_new].to(torch.float32)
                self._lp = self._lp_store[:, :T_new]  # contiguous for batch 1
            else:
                new = x.to(torch.float32).contiguous().clone()
                new[:, :T_old] = old
                self._lp = new
                self._lp_t, self._lp_t_T = None, 0
            self.input_length = x.size(1)
            self.end_frames = torch.as_tensor([x.size(1)]) - 1
            self._xlens_dev = torch.full((self.batch,), x.size(1), dtype=torch.int64,
                                         device=self.device)
            self._end_frames_dev = self._xlens_dev - 1
            if self.margin > 0:
                self.frame_ids = torch.arange(self.input_length, dtype=self.dtype,
                                              device=self.device)

    @on_tensor_device
    def extend_state(self, state):
        """Compute CTC prefix state for the extended input (blank running sum).

        :param state    : CTC state
        :return ctc_state
        """
        if state is None:
            return state
        r_prev, s_prev, f_min_prev, f_max_prev = state
        T = self.input_length
        r_new = torch.full((T, 2), self.logzero, dtype=r_prev.dtype, device=self.device)
        start = max(r_prev.shape[0], 1)
        r_new[0:start] = r_prev
        if start < T:
            tail = self._lp[0, start:, self.blank].to(r_prev.dtype)
            r_new[start:, 1] = r_new[start - 1, 1] + torch.cumsum(tail, 0)
        return (r_new, s_prev, f_min_prev, f_max_prev)
