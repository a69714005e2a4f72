#!/usr/bin/env python3
"""Beam-search glue timing at C3 widths: beam 10, V=5000, T=500,
joint CTC + bigram-table decoder + length bonus, pre-beam 15 ("full").
  engine_batch32 : BatchBeamSearch.forward_batch, 32 utterances together
  engine_single  : BatchBeamSearch.forward, one utterance at a time
  listglue_single: the reference's per-hypothesis glue (oracle/beam_harness.py restates
                   batch_beam_search.py:206-283) around the SAME CUDA scorer
Fixed number of positions (maxlenratio < 0) so that the three do equal work."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from espnet1_b200 import functional as F  # noqa: E402
from espnet1_b200.espnet.nets.batch_beam_search import BatchBeamSearch  # noqa: E402
from espnet1_b200.espnet.nets.scorers.ctc import CTCPrefixScorer  # noqa: E402
from espnet1_b200.espnet.nets.scorers.length_bonus import LengthBonus  # noqa: E402
from oracle import beam_harness  # noqa: E402


class TableScorer:
    """Stand-in attention decoder of the benchmark (defined here: the measured engine path uses
    nothing from oracle/): next-token log-probs depend on the last token only (a bigram table),
    the role of the tiny random decoders in the reference's test/test_beam_search.py:90-119."""

    def __init__(self, table):
        self.table = table

    def init_state(self, x):
        return None

    def batch_score(self, ys, states, xs):
        return self.table[ys[:, -1]], None

    def score(self, y, state, x):
        return self.table[int(y[-1])], None

    def final_score(self, state):
        return 0.0


D = 256  # encoder output size: x (T, D) -> ctc_lo -> (T, V), as espnet2/asr/ctc.py:109-117


def run(dev="cuda:0", POS=30, cpu_port=True):
    """Returns the timing dict; imported by bench.py (``beam_search`` block)."""
    from espnet1_b200.espnet2.asr.ctc import CTC
    from oracle.scorer_port import LengthBonusPort, OracleCTCPrefixScorer

    U, W, V, T = 32, 10, 5000, 500
    g = torch.Generator().manual_seed(5)
    xs_h = torch.randn(U, T, D, generator=g)
    xs = xs_h.to(dev)
    lens = [T] * U
    table_h = torch.log_softmax(torch.randn(V, V, generator=g) * 2, dim=1)
    table = table_h.to(dev)
    eos = V - 1
    weights = {"ctc": 0.3, "decoder": 0.7, "length_bonus": 0.2}
    ctc = CTC(V, D)
    with torch.no_grad():
        ctc.ctc_lo.weight.copy_(torch.randn(V, D, generator=g) * 0.5)  # peaky-ish posteriors
        ctc.ctc_lo.bias.zero_()
    w_h, b_h = ctc.ctc_lo.weight.detach().clone(), ctc.ctc_lo.bias.detach().clone()
    ctc = ctc.to(dev)

    class CpuCTC:  # the reference's CPU arithmetic for ctc.log_softmax (espnet2/asr/ctc.py:109-117)
        def log_softmax(self, x):
            return torch.log_softmax(torch.nn.functional.linear(x, w_h, b_h), dim=-1)

    def scorers():
        return {"ctc": CTCPrefixScorer(ctc, eos), "decoder": TableScorer(table), "length_bonus": LengthBonus(V)}

    bs = BatchBeamSearch(scorers(), weights, W, V, eos, eos, pre_beam_score_key="full")
    out = {"config": f"beam {W}, V={V}, T={T}, D={D}, pre-beam 15 (full), ctc 0.3 + bigram-table decoder 0.7 "
                     f"+ length bonus 0.2, {POS} positions per utterance (maxlenratio=-{POS})"}
    cases = [
        ("engine_batch32", lambda: bs.forward_batch(xs.clone(), lens, maxlenratio=-POS), U),
        ("engine_single", lambda: [bs(xs[u].clone(), maxlenratio=-POS) for u in range(4)], 4),
        ("listglue_single", lambda: [beam_harness.batch_beam_search(
            xs[u].clone(), scorers(), weights, W, V, eos, eos, pre_beam_score_key="full",
            maxlenratio=-POS) for u in range(4)], 4),
    ]
    if cpu_port:
        torch.set_num_threads(os.cpu_count() or 1)
        cases.append(("cpu_port_single", lambda: [beam_harness.batch_beam_search(
            xs_h[0].clone(), {"ctc": OracleCTCPrefixScorer(CpuCTC(), eos), "decoder": TableScorer(table_h),
                              "length_bonus": LengthBonusPort(V)},
            weights, W, V, eos, eos, pre_beam_score_key="full", maxlenratio=-POS)], 1))
    only = os.environ.get("CASES")  # e.g. CASES=engine_batch32 under ncu
    if only:
        cases = [c for c in cases if c[0] in only.split(",")]
    with torch.no_grad():
        for name, fn, n_utt in cases:
            if not name.startswith("cpu"):
                fn()
            torch.cuda.synchronize()
            if bs.prof is not None:
                bs.prof.clear()
            # GPU cases: 5 timed runs, the median is reported and all samples kept (one sample
            # now and then carries a pinned-memory or device allocation of tens of ms)
            samples = []
            for _ in range(1 if name.startswith("cpu") else 5):
                t0 = time.perf_counter()
                res = fn()
                torch.cuda.synchronize()
                samples.append(time.perf_counter() - t0)
            dt = sorted(samples)[len(samples) // 2]
            first = res[0][0]
            out[name] = {"utterances": n_utt, "positions": POS, "s": dt, "samples_s": [round(v, 6) for v in samples],
                         "ms_per_position": 1e3 * dt / POS / (1 if "batch" in name else n_utt),
                         "utt_positions_per_s": n_utt * POS / dt,
                         "best0": [int(v) for v in (first.yseq if hasattr(first, "yseq") else first["yseq"])][:8]}
            if bs.prof is not None and name.startswith("engine"):
                out[name]["phase_ms_per_position"] = {
                    k: round(1e3 * v / len(samples) / POS / (1 if "batch" in name else n_utt), 4)
                    for k, v in bs.prof.items()}
    if "engine_batch32" in out:
        # marginal cost of a position: two decodes of 2 x POS and POS positions, set-up and n-best
        # construction cancel (they are the same in both)
        with torch.no_grad():
            ts = []
            for pos in (POS, 2 * POS):
                best = 1e9
                for _ in range(3):
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    bs.forward_batch(xs.clone(), lens, maxlenratio=-pos)
                    torch.cuda.synchronize()
                    best = min(best, time.perf_counter() - t0)
                ts.append(best)
        out["engine_batch32"]["marginal_ms_per_position"] = 1e3 * (ts[1] - ts[0]) / POS
    out["same_best_hypothesis"] = len({tuple(v["best0"]) for k, v in out.items() if isinstance(v, dict)}) == 1
    return out


def main():
    print(json.dumps(run(POS=int(os.environ.get("POS", "30")))))


if __name__ == "__main__":
    main()
