"""Drop-in for the legacy ``espnet.nets.pytorch_backend.ctc.CTC`` / ``ctc_for``.

Mirrors /root/reference/espnet/nets/pytorch_backend/ctc.py:12-269: 3-argument
``forward(hs_pad, hlens, ys_pad)`` with ``-1``-padded labels, ``self.loss`` /
``self.probs`` side attributes, ``zero_infinity=True`` always (``:38-40``),
``/B`` then ``.sum()`` when ``reduce`` (``:62-64,129-131``), ``forced_align``.
"""
from __future__ import annotations

import logging

import torch

from .... import functional as B


class CTC(torch.nn.Module):
    """CTC module

    :param int odim: dimension of outputs
    :param int eprojs: number of encoder projection units
    :param float dropout_rate: dropout rate (0.0 ~ 1.0)
    :param str ctc_type: builtin
    :param bool reduce: reduce the CTC loss into a scalar
    """

    def __init__(self, odim, eprojs, dropout_rate, ctc_type="builtin", reduce=True):
        super().__init__()
        self.dropout_rate = dropout_rate
        self.loss = None
        self.ctc_lo = torch.nn.Linear(eprojs, odim)
        self.dropout = torch.nn.Dropout(dropout_rate)
        self.probs = None  # for visualization
        # torch >= 1.7 always maps to the builtin loss in the reference (:31-35)
        if ctc_type not in ("builtin", "cudnnctc", "gtnctc", "warpctc"):
            raise ValueError('ctc_type must be "builtin" or "gtnctc": {}'.format(ctc_type))
        self.ctc_type = "builtin"
        if ctc_type != self.ctc_type:
            logging.warning(f"CTC was set to {self.ctc_type} (B200 CUDA path).")
        self.ignore_id = -1
        self.reduce = reduce

    def loss_fn(self, th_pred, th_target, th_ilen, th_olen, max_tgt_len=None):
        """(T,B,V) activations -> loss / B; summed over the batch iff reduce
        (the reference folds that sum into CTCLoss(reduction="sum"), :37-40)."""
        logits = th_pred.transpose(0, 1)
        size = th_pred.size(1)
        if self.reduce:  # sum()/B comes out of the loss kernels themselves
            return B.ctc_loss(logits, th_target, th_ilen, th_olen, blank=0, zero_infinity=True,
                              max_tgt_len=max_tgt_len, reduce_scale=B._f32_recip(size))
        nll = B.ctc_loss(logits, th_target, th_ilen, th_olen, blank=0, zero_infinity=True,
                         max_tgt_len=max_tgt_len)
        return nll / size

    def forward(self, hs_pad, hlens, ys_pad):
        """CTC forward

        :param torch.Tensor hs_pad: batch of padded hidden state sequences (B, Tmax, D)
        :param torch.Tensor hlens: batch of lengths of hidden state sequences (B)
        :param torch.Tensor ys_pad: batch of padded character id sequence tensor (B, Lmax)
        :return: ctc loss value
        :rtype: torch.Tensor
        """
        keep = ys_pad != self.ignore_id  # parse padded ys (:84)
        olens = keep.sum(dim=1)
        ys_true = ys_pad[keep]
        ys_hat = self._logits(self.dropout(hs_pad)).transpose(0, 1)
        hlens = torch.as_tensor(hlens).long()
        self.loss = self.loss_fn(ys_hat, ys_true, hlens, olens, max_tgt_len=ys_pad.size(1))
        self.loss = self.loss.to(dtype=hs_pad.dtype)
        if self.reduce:
            self.loss = self.loss.sum()
        return self.loss

    def _logits(self, hs):
        """``ctc_lo(hs)`` for the loss.  In fp32 cuBLASLt adds the Linear's bias in a separate pass
        over the (B, T, odim) activations and autograd sums the gradient's columns for d bias in
        another one (0.9 ms of 6.4 ms for forward + backward at B=64, T=500, 256 -> 5000 -
        tools/exp/linear_ones.py - against 0.33 ms for the whole CTC loss step): the bias rides
        along as a 257th input column instead, ``[hs, 1] @ [W, b]^T`` - same values (bit-identical
        in that measurement), ``ctc_lo`` keeps its parameters and checkpoint keys."""
        lo = self.ctc_lo
        if (isinstance(lo, torch.nn.Linear) and lo.bias is not None and hs.is_cuda and hs.dtype == torch.float32
                and lo.weight.dtype == torch.float32 and not torch.is_autocast_enabled()):
            ones = hs.new_ones(*hs.shape[:-1], 1)
            return torch.nn.functional.linear(torch.cat((hs, ones), -1), torch.cat((lo.weight, lo.bias.unsqueeze(1)), 1))
        return lo(hs)

    def _projection(self, hs_pad):
        """``ctc_lo(hs_pad)`` as (GEMM output, bias for the kernel to add) - see
        :meth:`espnet1_b200.espnet2.asr.ctc.CTC._projection`."""
        lo = self.ctc_lo
        if (isinstance(lo, torch.nn.Linear) and lo.bias is not None and hs_pad.is_cuda
                and hs_pad.dtype == lo.weight.dtype == lo.bias.dtype and not torch.is_autocast_enabled()):
            return torch.nn.functional.linear(hs_pad, lo.weight), lo.bias
        return lo(hs_pad), None

    def softmax(self, hs_pad):
        """softmax of frame activations (B, Tmax, eprojs) -> (B, Tmax, odim)"""
        self.probs = B.softmax(*self._projection(hs_pad))
        return self.probs

    def log_softmax(self, hs_pad):
        """log_softmax of frame activations (B, Tmax, eprojs) -> (B, Tmax, odim)"""
        return B.log_softmax(*self._projection(hs_pad))

    def argmax(self, hs_pad):
        """argmax of frame activations (B, Tmax, eprojs) -> (B, Tmax)"""
        return B.argmax(*self._projection(hs_pad))

    def forced_align(self, h, y, blank_id=0):
        """forced alignment (:163-226).

        :param torch.Tensor h: hidden state sequence, 2d tensor (T, D) or (1, T, D)
        :param torch.Tensor y: id sequence tensor 1d tensor (L)
        :param int blank_id: blank symbol index
        :return: best alignment results
        :rtype: list
        """
        lpz = self.log_softmax(h)
        lpz = lpz.squeeze(0)
        y = torch.as_tensor(y, device=lpz.device)
        return B.forced_align(lpz, y, blank_id).tolist()


def ctc_for(args, odim, reduce=True):
    """Returns the CTC module(s) for the given args and output dimension (:229-269)."""
    num_encs = getattr(args, "num_encs", 1)
    if num_encs == 1:
        return CTC(odim, args.eprojs, args.dropout_rate, ctc_type=args.ctc_type, reduce=reduce)
    elif num_encs >= 1:
        ctcs_list = torch.nn.ModuleList()
        n = 1 if args.share_ctc else num_encs
        for idx in range(n):
            ctcs_list.append(CTC(odim, args.eprojs, args.dropout_rate[idx],
                                 ctc_type=args.ctc_type, reduce=reduce))
        return ctcs_list
    else:
        raise ValueError(
            "Number of encoders needs to be more than one. {}".format(num_encs))
