"""Pin the CPU oracle against fixtures produced by the real reference.

The reference ships no golden vectors for the CTC path; tests/golden/*.npz are
outputs of the unmodified reference modules (tests/golden/make_golden.py).
"""
import numpy as np
import pytest
import torch

from oracle import beam_harness, ctc_numpy, espnet_cpu, prefix_numpy
from oracle.prefix_torch import PrefixOracle
from oracle.scorer_port import LengthBonusPort, OracleCTCPrefixScorer, TableScorer

from _util import PREFIX_CASES, close, load_json, load_npz, replay_prefix_trace


@pytest.fixture(scope="module")
def gm():
    return load_npz("ctc_module.npz")


def _targets(g):
    return [g["ys_pad"][i, : g["ys_lens"][i]] for i in range(len(g["ys_lens"]))]


def test_torch_port_matches_espnet2_module(gm):
    hs = torch.tensor(gm["hs"])
    w, b = torch.tensor(gm["weight"]), torch.tensor(gm["bias"])
    for reduce, tag in ((True, "reduce"), (False, "noreduce")):
        x = hs.clone().requires_grad_(True)
        loss = espnet_cpu.ctc_forward(x, torch.tensor(gm["hlens"]), torch.tensor(gm["ys_pad"]),
                                      torch.tensor(gm["ys_lens"]), w, b, reduce=reduce)
        close(loss, gm[f"loss_{tag}"], 1e-6, 1e-6)
        B = hs.shape[0]
        (loss if reduce else (loss * torch.arange(1, B + 1)).sum()).backward()
        close(x.grad, gm[f"ghs_{tag}"], 1e-5, 1e-6)


def test_torch_port_matches_legacy_module(gm):
    hs = torch.tensor(gm["hs"])
    w, b = torch.tensor(gm["weight"]), torch.tensor(gm["bias"])
    for reduce, tag in ((True, "reduce"), (False, "noreduce")):
        loss = espnet_cpu.legacy_ctc_forward(hs, torch.tensor(gm["hlens"]),
                                             torch.tensor(gm["ys_pad"]), w, b, reduce=reduce)
        close(loss, gm[f"legacy_loss_{tag}"], 1e-6, 1e-6)


def test_numpy_ctc_matches_reference_logit_grad(gm):
    nll, grad = ctc_numpy.ctc_loss_and_grad(gm["logits"], _targets(gm), gm["hlens"],
                                            grad_out=1.0 / len(gm["hlens"]))
    B = len(nll)
    close(nll / B, gm["loss_noreduce"], 1e-5, 1e-6)
    close(ctc_numpy.espnet2_loss(nll, True), gm["loss_reduce"], 1e-5, 1e-6)
    assert nll[3] == 0.0  # infeasible utterance zeroed (zero_infinity)
    close(grad, gm["glogits_reduce"], 0, 1e-5)
    assert np.all(grad[3] == 0) and np.all(grad[1, gm["hlens"][1]:] == 0)
    nll_inf, _ = ctc_numpy.ctc_loss_and_grad(gm["logits"], _targets(gm), gm["hlens"],
                                             zero_infinity=False)
    assert np.isinf(nll_inf[3]) and np.isinf(gm["loss_noreduce_keepinf"][3])


def test_numpy_softmax_family(gm):
    close(ctc_numpy.log_softmax(gm["logits"]), gm["log_softmax"], 1e-5, 1e-6)
    close(ctc_numpy.softmax(gm["logits"]), gm["softmax"], 1e-5, 1e-7)
    np.testing.assert_array_equal(ctc_numpy.argmax(gm["logits"]), gm["argmax"])


def test_argmax_ties_and_nan():
    x = np.array([[1.0, 3.0, 3.0, 2.0], [np.nan, 5.0, np.nan, 1.0], [-np.inf] * 4],
                 dtype=np.float32)
    np.testing.assert_array_equal(ctc_numpy.argmax(x), torch.argmax(torch.tensor(x), -1).numpy())


def test_forced_align(gm):
    lp = ctc_numpy.log_softmax(gm["logits"][0])
    ali = ctc_numpy.viterbi_align(lp, _targets(gm)[0])
    np.testing.assert_array_equal(np.asarray(ali), gm["forced_align_utt0"])


def test_forced_align_and_greedy_collapse_vs_reference_fixture():
    """tests/golden/align_greedy.npz: five alignments of the reference's ``forced_align`` and the
    greedy collapse its own ``calculate_cer_ctc`` produced (make_golden.py ``make_align_greedy``)."""
    g = load_npz("align_greedy.npz")
    for i in range(int(g["align_n"])):
        ali = ctc_numpy.viterbi_align(g[f"align{i}__lp"], g[f"align{i}__y"])
        np.testing.assert_array_equal(np.asarray(ali), g[f"align{i}__ali"], err_msg=f"case {i}")
    am = ctc_numpy.argmax(g["greedy__logits"])
    for b in range(am.shape[0]):
        want = g["greedy__ids"][b, : int(g["greedy__lens"][b])].tolist()
        assert ctc_numpy.greedy_collapse(am[b]) == want, b


def test_numpy_ctc_c1():
    g = load_npz("ctc_c1.npz")
    off = np.concatenate(([0], np.cumsum(g["tgt_lens"])))
    tg = [g["targets"][off[i]: off[i + 1]] for i in range(len(g["tgt_lens"]))]
    nll, grad = ctc_numpy.ctc_loss_and_grad(g["logits"], tg, g["in_lens"],
                                            grad_out=1.0 / len(tg))
    close(nll, g["nll"], 1e-5, 0)
    # both sides are fp32 log-space lattices with |alpha+beta| ~ 5e2 (ulp 6e-5):
    # the reference's own rounding noise in the gradient is ~1e-5 absolute.
    close(grad, g["grad"], 0, 5e-5)
    n64, g64 = ctc_numpy.ctc_loss_and_grad(g["logits"], tg, g["in_lens"],
                                           grad_out=1.0 / len(tg), dtype=np.float64)
    close(n64, g["nll"], 1e-5, 0)
    close(g64, g["grad"], 0, 5e-5)
    # torch-CPU port at the logits boundary (the timed CPU baseline)
    n2, g2 = espnet_cpu.logits_loss_and_grad(torch.tensor(g["logits"]), torch.tensor(g["targets"]),
                                             torch.tensor(g["in_lens"]), torch.tensor(g["tgt_lens"]))
    close(n2, g["nll"], 1e-6, 0)
    close(g2, g["grad"], 0, 1e-7)


@pytest.mark.parametrize("tag", PREFIX_CASES)
def test_prefix_oracle_traces(tag):
    g = load_npz("prefix_th.npz")

    def make(lpz, xlens, blank, eos, margin):
        o = PrefixOracle(lpz, xlens, blank, eos, margin)
        o.__class__ = _Callable
        return o

    replay_prefix_trace(g, tag, make, rtol=1e-6, atol=1e-6)


class _Callable(PrefixOracle):
    def __call__(self, y, state, ids=None, att_w=None):
        return self.score(y, state, ids, att_w)

    def index_select_state(self, state, best):
        return self.select(state, best)


def test_prefix_numpy_trace():
    g = load_npz("prefix_np.npz")
    x = g["x"]
    V = x.shape[1]
    r = prefix_numpy.initial_state(x, 0)
    close(r, g["r0"], 1e-6, 1e-6)
    for i in range(5):
        y, cs = g[f"s{i}__y"], g[f"s{i}__cs"]
        psi, rs = prefix_numpy.score(x, y, cs, r, 0, V - 1)
        close(psi, g[f"s{i}__psi"], 1e-6, 1e-6)
        vf = int(g[f"s{i}__valid_from"])
        close(rs[:, vf:], g[f"s{i}__r"], 1e-6, 1e-6)
        k = int(np.argmax(np.where((cs == 0) | (cs == V - 1), -np.inf, psi)))
        r = rs[k]


class _StubCTC:
    def log_softmax(self, x):
        return torch.log_softmax(x, dim=-1)


@pytest.mark.parametrize("name", ["ctc_only_v5", "joint_v5", "joint_v40", "ctc_only_v40"])
def test_beam_harness_matches_reference_nbest(name):
    ref = load_json("beam_nbest.json")[name]
    arr = load_npz("beam_inputs.npz")
    x = torch.tensor(arr[f"{name}__x"])
    V, eos = ref["V"], ref["eos"]
    for fn, tag in ((beam_harness.beam_search, "beam"), (beam_harness.batch_beam_search, "batch")):
        scorers = {"ctc": OracleCTCPrefixScorer(_StubCTC(), eos), "length_bonus": LengthBonusPort(V)}
        if "decoder" in ref["weights"]:
            scorers["decoder"] = TableScorer(torch.tensor(arr[f"{name}__table"]))
        with torch.no_grad():
            nbest = fn(x.clone(), scorers, ref["weights"], ref["beam_size"], V, eos, eos,
                       pre_beam_score_key=ref["pre_beam_key"])
        assert [h["yseq"] for h in nbest] == [h["yseq"] for h in ref[tag]], (name, tag)
        np.testing.assert_allclose([h["score"] for h in nbest], [h["score"] for h in ref[tag]],
                                   rtol=1e-6)


def test_streaming_extend():
    g = load_npz("prefix_stream.npz")
    T1, T2, V, W, snum, eos = [int(v) for v in g["cfg"]]
    x_full = torch.tensor(g["x_full"])
    sc = OracleCTCPrefixScorer(_StubCTC(), eos)
    sc.batch_init_state(x_full[:T1].clone())
    y = torch.full((W, 1), eos, dtype=torch.long)
    ids = torch.tensor(g["ids0"])
    scores, st = sc.batch_score_partial(y, ids, [None] * W, x_full[:T1])
    close(scores, g["scores0"], 1e-6, 1e-6)
    states = [sc.select_state(st, i, int(ids[i, 0])) for i in range(W)]
    sc.extend_prob(x_full.clone())
    states = sc.extend_state(states)
    close(torch.stack([s[0] for s in states]), g["ext_r"], 1e-6, 1e-6)
    y = torch.cat((y, torch.tensor(g["toks"]).view(-1, 1)), 1)
    scores, st = sc.batch_score_partial(y, torch.tensor(g["ids1"]), states, x_full)
    close(scores, g["scores1"], 1e-6, 1e-6)
    close(st[0], g["r1"], 1e-6, 1e-6)
