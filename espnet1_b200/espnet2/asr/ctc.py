"""Drop-in for ``espnet2.asr.ctc.CTC`` on top of the B200 CUDA library.

Mirrors /root/reference/espnet2/asr/ctc.py:6-127: same constructor signature
(the ``--ctc_conf`` surface, ``:19-28``), same attributes (``ctc_lo`` keeps the
checkpoint keys ``ctc.ctc_lo.{weight,bias}``, ``ctc_type``, ``reduce``,
``dropout_rate``), same methods and result conventions.  ``ctc_lo`` stays a
cuBLAS GEMM (dense contraction, out of the hot-path scope); everything after
the logits runs in libb200ctc.so and there is no CPU fallback.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from ... import functional as B


class CTC(torch.nn.Module):
    """CTC module (B200-native).

    Args:
        odim: dimension of outputs
        encoder_output_size: number of encoder projection units
        dropout_rate: dropout rate (0.0 ~ 1.0)
        ctc_type: builtin or gtnctc
        reduce: reduce the CTC loss into a scalar
        ignore_nan_grad: same as zero_infinity (kept for backward compatibility)
        zero_infinity: zero infinite losses and the associated gradients
    """

    def __init__(
        self,
        odim: int,
        encoder_output_size: int,
        dropout_rate: float = 0.0,
        ctc_type: str = "builtin",
        reduce: bool = True,
        ignore_nan_grad: bool = None,
        zero_infinity: bool = True,
    ):
        super().__init__()
        eprojs = encoder_output_size
        self.dropout_rate = dropout_rate
        self.ctc_lo = torch.nn.Linear(eprojs, odim)
        self.ctc_type = ctc_type
        if ignore_nan_grad is not None:  # reference :35-36
            zero_infinity = ignore_nan_grad
        self.zero_infinity = bool(zero_infinity)

        if self.ctc_type == "builtin":
            self.ctc_loss = self._builtin_ctc_loss
        elif self.ctc_type == "gtnctc":
            # reference :43-46 delegates to the optional third-party `gtn` WFST
            # library on CPU threads; that backend is outside the CUDA hot path.
            raise NotImplementedError(
                'ctc_type="gtnctc" needs the third-party gtn library and is not part of the '
                'B200 CTC path; use ctc_type="builtin"')
        else:
            raise ValueError(f'ctc_type must be "builtin" or "gtnctc": {self.ctc_type}')

        self.reduce = reduce
        # batch-sharded training: set to a sharded.PeerLossExchange and every loss_fn call also
        # leaves the global batch-weighted loss of all ranks in `loss_exchange.out`
        # (the all-reduce of trainer.py:604-613, fused into the loss kernel)
        self.loss_exchange = None
        # The gradient w.r.t. the logits is produced during the forward call, pre-scaled by the
        # upstream gradient this loss is EXPECTED to receive: 1/B times `upstream_scale`.  Set it to
        # the factor the caller multiplies the CTC loss with before backward() - `ctc_weight` in
        # espnet2/asr/espnet_model.py (`loss = ctc_weight * loss_ctc + ...`), an AMP loss scale -
        # and backward stays a no-op check; any other upstream gradient is still exact, it only
        # costs one rescaling pass over the (B,T,V) gradient.
        self.upstream_scale = 1.0

    def _grad_hint(self, per_utt_scale: float):
        if self.upstream_scale == 1.0:
            return None  # functional default: per_utt_scale itself
        import numpy as np

        # fp32 product, exactly what autograd hands to backward (fl32(weight) * fl32(1/B))
        return float(np.float32(self.upstream_scale) * np.float32(per_utt_scale))

    # same call convention as torch.nn.CTCLoss(reduction="none") but fed with
    # batch-major LOGITS: log-softmax is fused into the loss kernels.
    def _builtin_ctc_loss(self, logits_btv, th_target, th_ilen, th_olen, max_tgt_len=None,
                          reduce_scale=None):
        per_utt = reduce_scale if reduce_scale is not None else B._f32_recip(logits_btv.shape[0])
        return B.ctc_loss(logits_btv, th_target, th_ilen, th_olen, blank=0,
                          zero_infinity=self.zero_infinity, max_tgt_len=max_tgt_len,
                          reduce_scale=reduce_scale, grad_hint=self._grad_hint(per_utt),
                          peer=self.loss_exchange)

    def loss_fn(self, th_pred, th_target, th_ilen, th_olen, max_tgt_len=None) -> torch.Tensor:
        """th_pred is (T, B, V) like the reference (:52); a transposed view of a
        batch-major tensor (what ``forward`` passes) costs no copy."""
        if self.ctc_type != "builtin":
            raise NotImplementedError
        logits = th_pred.transpose(0, 1)  # (B, T, V); contiguous when th_pred was a view
        size = th_pred.size(1)
        if self.reduce:
            # batch-size average (:58-60): sum()/size comes out of the loss kernels
            return self.ctc_loss(logits, th_target, th_ilen, th_olen, max_tgt_len,
                                 reduce_scale=B._f32_recip(size))
        return self.ctc_loss(logits, th_target, th_ilen, th_olen, max_tgt_len) / size

    def forward(self, hs_pad, hlens, ys_pad, ys_lens):
        """Calculate CTC loss.

        Args:
            hs_pad: batch of padded hidden state sequences (B, Tmax, D)
            hlens: batch of lengths of hidden state sequences (B)
            ys_pad: batch of padded character id sequence tensor (B, Lmax)
            ys_lens: batch of lengths of character sequence (B)
        """
        # NOTE: F.dropout defaults to training=True exactly like the reference (:82)
        ys_hat = self._logits(F.dropout(hs_pad, p=self.dropout_rate))
        ys_hat = ys_hat.transpose(0, 1)  # (B, T, V) -> (T, B, V) view
        # (B, L) -> (sum L,) without the reference's per-utterance Python loop (:91)
        Lmax = ys_pad.size(1)
        lens_dev = torch.as_tensor(ys_lens).to(device=ys_pad.device)
        keep = torch.arange(Lmax, device=ys_pad.device).unsqueeze(0) < lens_dev.unsqueeze(1)
        ys_true = ys_pad[keep]
        loss = self.loss_fn(ys_hat, ys_true, hlens, ys_lens, max_tgt_len=Lmax).to(
            device=hs_pad.device, dtype=hs_pad.dtype
        )
        return loss

    def forward_multi(self, hs_pads, hlens, ys_pad, ys_lens, upstream_scales=None):
        """Losses of several encoder outputs against the SAME labels: the intermediate-CTC
        pattern of ``espnet2/asr/espnet_model.py:243-286`` (one ``self.ctc(...)`` call per
        tapped layer plus the final one).  Returns ``[forward(hs, ...) for hs in hs_pads]``
        (same values and gradients) from ONE batched pass over heads x utterances:
        the hidden states are stacked (32 MB per head at C2, the logits would be 640 MB), ``ctc_lo``
        runs as one GEMM, and the loss kernels see a (H*B, T, V) batch - one label prep, one
        emission gather, one lattice launch with the heads' lattices co-resident on the SMs, one
        gradient row pass.  ``upstream_scales[h]``: the factor the caller multiplies head h's loss
        with before backward (ESPnet weights the final and the intermediate heads differently);
        with it backward is a no-op check per utterance, without it (or with other factors)
        gradients are still exact, the mismatching heads pay one rescaling pass."""
        H = len(hs_pads)
        # (with a sharded-loss exchange attached every head must make its own exchange call, in
        #  the same order on every rank: one `forward` per head, as the reference issues them)
        if H <= 1 or not hs_pads[0].is_cuda or self.ctc_type != "builtin" or self.loss_exchange is not None:
            return [self.forward(hs, hlens, ys_pad, ys_lens) for hs in hs_pads]
        import numpy as np

        Bn = hs_pads[0].size(0)
        dev = hs_pads[0].device
        hs = torch.cat(list(hs_pads), dim=0)  # (H*B, T, D)
        ys_hat = self._logits(F.dropout(hs, p=self.dropout_rate))  # one GEMM, independent masks per head
        Lmax = ys_pad.size(1)
        lens_dev = torch.as_tensor(ys_lens).to(device=ys_pad.device)
        keep = torch.arange(Lmax, device=ys_pad.device).unsqueeze(0) < lens_dev.unsqueeze(1)
        ys_true = ys_pad[keep].repeat(H)
        il = torch.as_tensor(hlens).to(device=dev, dtype=torch.int64).repeat(H)
        tl = lens_dev.to(device=dev, dtype=torch.int64).repeat(H)
        inv = B._f32_recip(Bn)
        scales = [self.upstream_scale] * H if upstream_scales is None else list(upstream_scales)
        hints = torch.tensor([float(np.float32(sc) * np.float32(inv)) for sc in scales],
                             dtype=torch.float32).repeat_interleave(Bn).to(dev, non_blocking=True)
        nll = B.ctc_loss(ys_hat, ys_true, il, tl, blank=0, zero_infinity=self.zero_infinity,
                         max_tgt_len=Lmax, grad_hint=hints)
        per_head = nll.view(H, Bn)
        if self.reduce:
            losses = [per_head[h].sum() * inv for h in range(H)]
        else:
            losses = [per_head[h] / Bn for h in range(H)]
        return [l.to(device=dev, dtype=hs_pads[0].dtype) for l in losses]

    def forward_multi_streams(self, hs_pads, hlens, ys_pad, ys_lens):
        """Round-1 form of :meth:`forward_multi`, kept for A/B measurement (tools/bench_multihead.py):
        one ``forward`` per head, issued on alternating side streams so that the latency-bound
        lattice of one head runs under the HBM-bound row pass of another."""
        if len(hs_pads) <= 1 or not hs_pads[0].is_cuda:
            return [self.forward(hs, hlens, ys_pad, ys_lens) for hs in hs_pads]
        dev = hs_pads[0].device
        cur = torch.cuda.current_stream(dev)
        if getattr(self, "_side_streams", None) is None or self._side_streams[0].device != dev:
            self._side_streams = [torch.cuda.Stream(dev), torch.cuda.Stream(dev)]
        ready = cur.record_event()
        out = []
        for i, hs in enumerate(hs_pads):
            side = self._side_streams[i % 2]
            side.wait_event(ready)
            with torch.cuda.stream(side):
                out.append(self.forward(hs, hlens, ys_pad, ys_lens))
        for side in self._side_streams[: len(hs_pads)]:
            cur.wait_stream(side)
        return out

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
            return F.linear(torch.cat((hs, ones), -1), torch.cat((lo.weight, lo.bias.unsqueeze(1)), 1))
        return lo(hs)

    def _projection(self, hs_pad):
        """``ctc_lo(hs_pad)`` as (GEMM output, bias to add): where the soft-max kernel can take the
        bias (same dtype as the GEMM output, CUDA) the GEMM runs without it - cuBLASLt adds the
        bias of this shape in a separate pass over the (B, T, odim) output, a third of the
        projection's time - and the soft-max's one read of each row absorbs it."""
        lo = self.ctc_lo
        if (isinstance(lo, torch.nn.Linear) and lo.bias is not None and hs_pad.is_cuda
                and hs_pad.dtype == lo.weight.dtype == lo.bias.dtype and not torch.is_autocast_enabled()):
            return F.linear(hs_pad, lo.weight), lo.bias
        return lo(hs_pad), None

    def softmax(self, hs_pad):
        """softmax of frame activations: (B, Tmax, eprojs) -> (B, Tmax, odim)."""
        x, bias = self._projection(hs_pad)
        return B.softmax(x, bias)

    def log_softmax(self, hs_pad):
        """log_softmax of frame activations: (B, Tmax, eprojs) -> (B, Tmax, odim)."""
        x, bias = self._projection(hs_pad)
        return B.log_softmax(x, bias)

    def argmax(self, hs_pad):
        """argmax of frame activations: (B, Tmax, eprojs) -> (B, Tmax) int64."""
        x, bias = self._projection(hs_pad)
        return B.argmax(x, bias)
