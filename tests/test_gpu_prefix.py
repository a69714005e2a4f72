"""GPU parity of the prefix scorer: golden traces of the real reference,
the oracle at larger sizes, and n-best identity through the beam-search caller."""
import numpy as np
import pytest
import torch

from oracle import beam_harness
from oracle.prefix_torch import PrefixOracle
from oracle.scorer_port import LengthBonusPort, TableScorer

from _util import LOGZERO, PREFIX_CASES, close, load_json, load_npz, replay_prefix_trace

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
# same class of tolerance as test/test_batch_beam_search.py:184-188 (rtol 1e-6 on
# hypothesis scores); per-frame lattice values get a little absolute slack
RTOL, ATOL = 2e-6, 2e-5


def _th():
    from espnet1_b200.espnet.nets.ctc_prefix_score import CTCPrefixScoreTH

    return CTCPrefixScoreTH


@pytest.mark.parametrize("tag", PREFIX_CASES)
def test_golden_traces(tag):
    g = load_npz("prefix_th.npz")
    replay_prefix_trace(g, tag, lambda *a: _th()(*a), device=DEV, rtol=RTOL, atol=ATOL)


def _drive(scorer, select, B, W, V, T, snum, margin, steps, seed, device):
    """recognize_beam_batch-style loop (rnn/decoders.py:811-937); returns traces."""
    g = torch.Generator().manual_seed(seed)
    eos = V - 1
    yseq = [[eos] for _ in range(B * W)]
    vscores = torch.zeros(B, W)
    state, out = None, []
    pad_b = (torch.arange(B) * W).view(-1, 1)
    t = torch.arange(T, dtype=torch.float32)
    for i in range(steps):
        local = torch.randn(B * W, V, generator=g)
        local[:, 0] = LOGZERO
        part = torch.topk(local, snum, dim=-1)[1] if snum else None
        att = None
        if margin > 0:
            c = T * (i + 1) / (steps + 2.0) + (torch.rand(B * W, generator=g) * 6 - 3)
            att = torch.softmax(-((t[None] - c[:, None]) ** 2) / 18.0, dim=1)
        sc, st5 = scorer(yseq, state, None if part is None else part.to(device),
                         None if att is None else att.to(device))
        sc_c = sc.detach().cpu()
        out.append((sc_c, st5[1].detach().cpu(), int(st5[2]), int(st5[3])))
        tot = (local + 0.5 * sc_c).view(B, W, V)
        if i == 0:
            tot[:, 1:, :] = LOGZERO
        vs = vscores.view(B, W, 1).repeat(1, 1, V)
        vs[:, :, eos] = LOGZERO
        vscores, best = torch.topk((vs + tot).view(B, -1), W, 1)
        tok = torch.fmod(best, V).view(-1).tolist()
        src = (best // V + pad_b).view(-1).tolist()
        yseq = [yseq[s][:] + [k] for s, k in zip(src, tok)]
        state = select(st5, best.to(device))
    return out, yseq


@pytest.mark.parametrize("B,W,V,T,snum,margin", [
    (4, 5, 500, 120, 8, 0),
    (4, 5, 500, 120, 8, 10),
    (2, 3, 300, 90, None, 0),
    (2, 3, 300, 90, None, 12),
    (32, 10, 5000, 100, 15, 20),   # C3 widths at a shorter T (CPU oracle stays fast)
    # shapes around the time-parallel (scan) kernel's edges: one candidate, 16 / 17 / 33 candidates
    # (one, two, three CTAs per hypothesis), T not a multiple of 32, windows that cut chunks short,
    # and T < 64 (the serial kernel keeps those)
    (3, 4, 200, 65, 1, 0),
    (3, 4, 200, 333, 16, 0),
    (3, 4, 200, 333, 17, 9),
    (2, 3, 120, 97, 33, 4),
    (2, 3, 120, 40, 5, 0),
    (2, 3, 120, 640, 7, 30),
])
def test_vs_oracle_multi_step(B, W, V, T, snum, margin):
    g = torch.Generator().manual_seed(1000 + V + (snum or 0) + margin)
    lpz = torch.log_softmax(torch.randn(B, T, V, generator=g) * 2, dim=2)
    xlens = torch.randint(T // 2, T + 1, (B,), generator=g)
    xlens[0] = T
    eos = V - 1
    ora = PrefixOracle(lpz.clone(), xlens, 0, eos, margin)
    cu = _th()(lpz.clone().to(DEV), xlens, 0, eos, margin)
    ref, yref = _drive(lambda *a: ora.score(*a), ora.select, B, W, V, T, snum, margin, 6, 7, "cpu")
    got, ygot = _drive(cu, cu.index_select_state, B, W, V, T, snum, margin, 6, 7, DEV)
    assert yref == ygot  # identical surviving token sequences
    if T <= 128:
        for (s0, p0, a0, b0), (s1, p1, a1, b1) in zip(ref, got):
            assert (a0, b0) == (a1, b1)
            close(s1, s0.numpy(), RTOL, ATOL)
            close(p1, p0.numpy(), RTOL, ATOL)
        return
    # Long inputs: hundreds of fp32 additions to values around -500 (ulp 3e-5) leave BOTH fp32
    # implementations ~1e-3 from exact arithmetic, and the time-parallel kernel adds in another
    # order than the reference's frame loop - so the yardstick is the same recursion in fp64:
    # the CUDA result must be as close to it as the reference's own fp32 arithmetic is.
    ora64 = PrefixOracle(lpz.double().clone(), xlens, 0, eos, margin)
    if margin > 0:  # the window arithmetic stays the reference's fp32 (same windows in all three)
        ora64.frame_ids = ora64.frame_ids.float()
    tru, ytru = _drive(lambda *a: ora64.score(*a), ora64.select, B, W, V, T, snum, margin, 6, 7, "cpu")
    assert ytru == ygot
    for (s0, p0, a0, b0), (s1, p1, a1, b1), (s2, p2, _, _) in zip(ref, got, tru):
        assert (a0, b0) == (a1, b1)
        for x32, xcu, x64 in ((s0, s1, s2), (p0, p1, p2)):
            live = x64 > -1e9  # log-zero entries: multiples of -1e10, compared separately
            e_ref = (x32.double() - x64)[live].abs().max().item() if live.any() else 0.0
            e_cu = (xcu.double() - x64)[live].abs().max().item() if live.any() else 0.0
            assert e_cu <= max(2.0 * e_ref, 2e-4), (e_cu, e_ref)
            close(xcu[~live], x32[~live].numpy(), 1e-6, 0)


class _StubCTC:
    def log_softmax(self, x):
        from espnet1_b200 import functional as F

        return F.log_softmax(x)


@pytest.mark.parametrize("name", ["ctc_only_v5", "joint_v5", "joint_v40", "ctc_only_v40"])
def test_beam_search_nbest_identical_to_reference(name):
    """Token sequences identical and scores rtol 1e-6-class vs the n-best lists the
    REAL reference BeamSearch / BatchBeamSearch produced (tests/golden)."""
    from espnet1_b200.espnet.nets.scorers.ctc import CTCPrefixScorer

    ref = load_json("beam_nbest.json")[name]
    arr = load_npz("beam_inputs.npz")
    x = torch.tensor(arr[f"{name}__x"], device=DEV)
    V, eos = ref["V"], ref["eos"]
    for fn, tag in ((beam_harness.batch_beam_search, "batch"), (beam_harness.beam_search, "beam")):
        scorers = {"ctc": CTCPrefixScorer(_StubCTC(), eos), "length_bonus": LengthBonusPort(V)}
        if "decoder" in ref["weights"]:
            scorers["decoder"] = TableScorer(torch.tensor(arr[f"{name}__table"], device=DEV))
        with torch.no_grad():
            nbest = fn(x.clone(), scorers, ref["weights"], ref["beam_size"], V, eos, eos,
                       pre_beam_score_key=ref["pre_beam_key"])
        assert [h["yseq"] for h in nbest] == [h["yseq"] for h in ref[tag]], (name, tag)
        np.testing.assert_allclose([h["score"] for h in nbest], [h["score"] for h in ref[tag]],
                                   rtol=2e-6)


def test_streaming_extend_vs_reference_fixture():
    from espnet1_b200.espnet.nets.scorers.ctc import CTCPrefixScorer

    g = load_npz("prefix_stream.npz")
    T1, T2, V, W, snum, eos = [int(v) for v in g["cfg"]]
    x_full = torch.tensor(g["x_full"], device=DEV)
    sc = CTCPrefixScorer(_StubCTC(), eos)
    sc.batch_init_state(x_full[:T1].clone())
    y = torch.full((W, 1), eos, dtype=torch.long, device=DEV)
    ids = torch.tensor(g["ids0"], device=DEV)
    scores, st = sc.batch_score_partial(y, ids, [None] * W, x_full[:T1])
    close(scores, g["scores0"], RTOL, ATOL)
    states = [sc.select_state(st, i, int(ids[i, 0])) for i in range(W)]
    sc.extend_prob(x_full.clone())
    states = sc.extend_state(states)
    close(torch.stack([s[0] for s in states]), g["ext_r"], RTOL, ATOL)
    y = torch.cat((y, torch.tensor(g["toks"], device=DEV).view(-1, 1)), 1)
    scores, st = sc.batch_score_partial(y, torch.tensor(g["ids1"], device=DEV), states, x_full)
    close(scores, g["scores1"], RTOL, ATOL)
    close(st[0], g["r1"], RTOL, ATOL)


def test_streaming_extend_in_several_chunks_vs_oracle():
    """Online decoding (batch_beam_search_online.py:306-309) extends the posteriors chunk by chunk.
    `extend_prob` keeps both layouts in buffers with headroom and writes only the NEW frames
    ((B,T,V) rows and the time-contiguous (B,V,T) columns): four growing lengths - the second
    fits the first buffer, the later ones regrow it - against the oracle scorer at every stage."""
    from espnet1_b200.espnet.nets.scorers.ctc import CTCPrefixScorer
    from oracle.scorer_port import OracleCTCPrefixScorer

    class _CpuCTC:
        def log_softmax(self, x):
            return torch.log_softmax(x, dim=-1)

    g = torch.Generator().manual_seed(21)
    V, W, snum, eos = 60, 4, 7, 59
    lens = [40, 55, 130, 400]
    x_full = torch.randn(lens[-1], V, generator=g) * 2
    cu = CTCPrefixScorer(_StubCTC(), eos)
    ora = OracleCTCPrefixScorer(_CpuCTC(), eos)
    cu.batch_init_state(x_full[: lens[0]].to(DEV))
    ora.batch_init_state(x_full[: lens[0]].clone())
    y = torch.full((W, 1), eos, dtype=torch.long)
    st_c, st_o = [None] * W, [None] * W
    for stage, T in enumerate(lens):
        if stage > 0:
            cu.extend_prob(x_full[:T].to(DEV))
            ora.extend_prob(x_full[:T].clone())
            st_c, st_o = cu.extend_state(st_c), ora.extend_state(st_o)
            for a, b in zip(st_c, st_o):
                close(a[0], b[0].numpy(), RTOL, ATOL, f"extended r, stage {stage}")
            assert cu.impl.input_length == T and tuple(cu.impl._lp.shape) == (1, T, V)
        ids = torch.stack([torch.randperm(V - 2, generator=g)[:snum] + 1 for _ in range(W)])
        sc_c, s5c = cu.batch_score_partial(y.to(DEV), ids.to(DEV), st_c, x_full[:T].to(DEV))
        sc_o, s5o = ora.batch_score_partial(y, ids, st_o, x_full[:T])
        close(sc_c, sc_o.numpy(), RTOL, ATOL, f"scores, stage {stage}")
        tok = ids[:, stage % snum]
        st_c = [cu.select_state(s5c, i, int(tok[i])) for i in range(W)]
        st_o = [ora.select_state(s5o, i, int(tok[i])) for i in range(W)]
        y = torch.cat((y, tok.view(-1, 1)), 1)
    # the time-contiguous copy was appended to, not rebuilt: it has headroom and holds all frames
    lp_t = cu.impl._lp_t
    assert lp_t is not None and lp_t.shape[2] >= lens[-1] and cu.impl._lp_t_T == lens[-1]
    np.testing.assert_array_equal(lp_t[0, :, : lens[-1]].cpu().numpy(),
                                  cu.impl._lp[0].transpose(0, 1).cpu().numpy())


def test_per_utterance_windows_equal_decoding_each_utterance_alone():
    """`per_utt_window=True` (extension): the batched scorer must give every utterance
    the scores, states and f_min / f_max that the REFERENCE-semantics scorer gives when that
    utterance is scored alone (a batch-wide window over a batch of one IS the per-utterance
    window) - here the oracle with batch 1; 5 carried steps, ragged lengths, attention centres
    that differ a lot between utterances (so the batch-wide window is much wider)."""
    B, W, V, T, snum, margin, steps = 4, 3, 80, 120, 6, 7, 5
    g = torch.Generator().manual_seed(77)
    lpz = torch.log_softmax(torch.randn(B, T, V, generator=g) * 2, dim=2)
    xlens = torch.tensor([120, 90, 120, 64])
    eos = V - 1
    cu = _th()(lpz.clone().to(DEV), xlens, 0, eos, margin, per_utt_window=True)
    alone = [PrefixOracle(lpz[b:b + 1].clone(), xlens[b:b + 1], 0, eos, margin) for b in range(B)]
    t = torch.arange(T, dtype=torch.float32)
    yseq = [[eos] for _ in range(B * W)]
    st_c, st_o = None, [None] * B
    speed = torch.tensor([3.0, 9.0, 15.0, 6.0])
    for i in range(steps):
        c = (speed.repeat_interleave(W) * (i + 1) + torch.rand(B * W, generator=g) * 4 - 2).clamp(1, T - 2)
        att = torch.softmax(-((t[None] - c[:, None]) ** 2) / 8.0, dim=1)
        part = torch.stack([torch.randperm(V - 2, generator=g)[:snum] + 1 for _ in range(B * W)])
        sc, st5 = cu(yseq, st_c, part.to(DEV), att.to(DEV))
        fwin = st5[2].cpu()
        best = []
        for b in range(B):
            sl = slice(b * W, (b + 1) * W)
            so, s5o = alone[b].score(yseq[sl], st_o[b], part[sl], att[sl])
            close(sc[sl], so.numpy(), RTOL, ATOL, f"scores utt {b} step {i}")
            close(st5[1][sl], s5o[1].numpy(), RTOL, ATOL, f"log_psi utt {b} step {i}")
            assert fwin[b].tolist() == [int(s5o[2]), int(s5o[3])], (b, i)
            bi = torch.topk(so.view(1, -1), W, 1)[1]
            best.append(bi)
            st_o[b] = alone[b].select(s5o, bi)
        best = torch.cat(best, 0)
        st_c = cu.index_select_state(st5, best.to(DEV))
        tok = torch.fmod(best, V).view(-1).tolist()
        src = (best // V + (torch.arange(B) * W).view(-1, 1)).view(-1).tolist()
        yseq = [yseq[s_][:] + [k] for s_, k in zip(src, tok)]
    assert len({tuple(v) for v in fwin.tolist()}) > 1  # the windows really differ per utterance


def test_public_attributes_and_x_view():
    lpz = torch.log_softmax(torch.randn(2, 6, 5), dim=2).to(DEV)
    s = _th()(lpz, torch.tensor([6, 4]), 0, 4, margin=3)
    assert (s.batch, s.input_length, s.odim, s.blank, s.eos, s.margin) == (2, 6, 5, 0, 4, 3)
    assert s.logzero == -10000000000.0 and s.dtype == torch.float32
    assert s.end_frames.tolist() == [5, 3]
    assert tuple(s.x.shape) == (2, 6, 2, 5)
    assert float(lpz[1, 5, 0]) == 0.0 and float(lpz[1, 5, 1]) == -10000000000.0  # padded in place
