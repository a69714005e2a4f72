"""GPU parity: CUDA CTC path (through the C ABI) vs the oracle and the golden fixtures.

Tolerances (BASELINE.json): loss 1e-5 relative, gradient 1e-4 absolute in fp32,
integer outputs identical.
"""
import numpy as np
import pytest
import torch

from _util import close, load_npz

pytestmark = pytest.mark.gpu

LOSS_RTOL = 1e-5
GRAD_ATOL = 1e-4


def _dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def gm():
    return load_npz("ctc_module.npz")


def _mk_espnet2(gm, **kw):
    from espnet1_b200.espnet2.asr.ctc import CTC

    V, D = gm["weight"].shape
    m = CTC(V, D, **kw).to(_dev())
    with torch.no_grad():
        m.ctc_lo.weight.copy_(torch.tensor(gm["weight"]))
        m.ctc_lo.bias.copy_(torch.tensor(gm["bias"]))
    return m


def test_espnet2_module_matches_reference_fixture(gm):
    d = _dev()
    m = _mk_espnet2(gm)
    hs = torch.tensor(gm["hs"], device=d)
    hl, yp, yl = (torch.tensor(gm[k], device=d) for k in ("hlens", "ys_pad", "ys_lens"))
    B = hs.shape[0]
    for reduce, tag in ((True, "reduce"), (False, "noreduce")):
        m.reduce = reduce
        m.zero_grad()
        x = hs.clone().requires_grad_(True)
        loss = m(x, hl, yp, yl)
        close(loss, gm[f"loss_{tag}"], LOSS_RTOL, 1e-6)
        (loss if reduce else (loss * torch.arange(1, B + 1, device=d)).sum()).backward()
        close(x.grad, gm[f"ghs_{tag}"], 0, GRAD_ATOL)
        close(m.ctc_lo.weight.grad, gm[f"gw_{tag}"], 1e-4, GRAD_ATOL)
        close(m.ctc_lo.bias.grad, gm[f"gb_{tag}"], 1e-4, GRAD_ATOL)


def test_espnet2_zero_infinity_false_keeps_inf(gm):
    d = _dev()
    m = _mk_espnet2(gm, zero_infinity=False, reduce=False)
    loss = m(torch.tensor(gm["hs"], device=d), torch.tensor(gm["hlens"], device=d),
             torch.tensor(gm["ys_pad"], device=d), torch.tensor(gm["ys_lens"], device=d))
    ref = gm["loss_noreduce_keepinf"]
    assert np.isinf(ref[3]) and torch.isinf(loss[3])
    keep = np.isfinite(ref)
    close(loss.detach().cpu().numpy()[keep], ref[keep], LOSS_RTOL, 1e-6)
    m2 = _mk_espnet2(gm, ignore_nan_grad=True, zero_infinity=False)
    assert m2.zero_infinity is True  # ignore_nan_grad overrides (ctc.py:35-36)


def test_ctc_type_handling(gm):
    from espnet1_b200.espnet2.asr.ctc import CTC

    with pytest.raises(ValueError):
        CTC(5, 4, ctc_type="nope")
    with pytest.raises(NotImplementedError):
        CTC(5, 4, ctc_type="gtnctc")
    m = CTC(5, 4)
    assert sorted(k for k, _ in m.named_parameters()) == ["ctc_lo.bias", "ctc_lo.weight"]


def test_cpu_tensor_fails_loudly(gm):
    m = _mk_espnet2(gm).cpu()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.tensor(gm["hs"]), torch.tensor(gm["hlens"]), torch.tensor(gm["ys_pad"]),
          torch.tensor(gm["ys_lens"]))


def test_legacy_module_matches_reference_fixture(gm):
    from espnet1_b200.espnet.nets.pytorch_backend.ctc import CTC

    d = _dev()
    V, D = gm["weight"].shape
    for reduce, tag in ((True, "reduce"), (False, "noreduce")):
        m = CTC(V, D, 0.0, reduce=reduce).to(d)
        with torch.no_grad():
            m.ctc_lo.weight.copy_(torch.tensor(gm["weight"]))
            m.ctc_lo.bias.copy_(torch.tensor(gm["bias"]))
        x = torch.tensor(gm["hs"], device=d).requires_grad_(True)
        loss = m(x, torch.tensor(gm["hlens"], device=d), torch.tensor(gm["ys_pad"], device=d))
        close(loss, gm[f"legacy_loss_{tag}"], LOSS_RTOL, 1e-6)
        loss.sum().backward()
        close(x.grad, gm[f"legacy_ghs_{tag}"], 0, GRAD_ATOL)
    ali = m.forced_align(torch.tensor(gm["hs"][0:1], device=d),
                         gm["ys_pad"][0][: gm["ys_lens"][0]].tolist())
    np.testing.assert_array_equal(np.asarray(ali), gm["forced_align_utt0"])


def test_softmax_family_vs_fixture(gm):
    d = _dev()
    m = _mk_espnet2(gm)
    hs = torch.tensor(gm["hs"], device=d)
    with torch.no_grad():
        close(m.log_softmax(hs), gm["log_softmax"], 1e-5, 1e-6)
        close(m.softmax(hs), gm["softmax"], 1e-5, 1e-7)
        np.testing.assert_array_equal(m.argmax(hs).cpu().numpy(), gm["argmax"])


@pytest.mark.parametrize("dtype,V", [(torch.float32, 5000), (torch.float32, 37), (torch.bfloat16, 5000),
                                     (torch.float16, 1024)])
def test_softmax_with_projection_bias(dtype, V):
    """`b200ctc_softmax_fwd_bias`: (log_)softmax(x + bias) with the bias absorbed by the kernel,
    forward and backward (incl. d bias) against torch autograd in fp32 on the same inputs; and
    `CTC.log_softmax` (GEMM without bias + this) against the plain module arithmetic."""
    from espnet1_b200 import functional as F
    from espnet1_b200.espnet2.asr.ctc import CTC

    d = _dev()
    g = torch.Generator().manual_seed(8)
    x = (torch.randn(7, 33, V, generator=g) * 3).to(dtype).to(d)
    b = torch.randn(V, generator=g).to(dtype).to(d)
    tol = 1e-5 if dtype == torch.float32 else (2e-2 if dtype == torch.bfloat16 else 3e-3)
    for fn, ref in ((F.log_softmax, torch.log_softmax), (F.softmax, torch.softmax)):
        xa = x.clone().requires_grad_(True)
        ba = b.clone().requires_grad_(True)
        y = fn(xa, ba)
        xr = x.float().clone().requires_grad_(True)
        br = b.float().clone().requires_grad_(True)
        yr = ref(xr + br, -1)
        close(y.float(), yr.detach().cpu().numpy(), tol, tol)
        w = torch.randn(y.shape, generator=g).to(d)
        (y.float() * w).sum().backward()
        (yr * w).sum().backward()
        close(xa.grad.float(), xr.grad.cpu().numpy(), 10 * tol, 10 * tol)
        close(ba.grad.float(), br.grad.cpu().numpy(), 50 * tol, 50 * tol * max(1.0, float(br.grad.abs().max())))
    # argmax of x + bias: the fp32 sum's first maximum (ties cannot be told from torch's for 16-bit
    # inputs summed in another precision, so those compare against the fp32 sum, which is what the kernel forms)
    want = (x.float() + b.float()).argmax(-1)
    assert torch.equal(F.argmax(x, b), want)
    if dtype == torch.float32:
        m = CTC(V, 16).to(d)
        hs = torch.randn(3, 20, 16, generator=g).to(d)
        with torch.no_grad():
            assert torch.equal(m.argmax(hs), m.ctc_lo(hs).argmax(-1))
            close(m.log_softmax(hs), torch.log_softmax(m.ctc_lo(hs), -1).cpu().numpy(), 1e-5, 2e-6)
            close(m.softmax(hs), torch.softmax(m.ctc_lo(hs), -1).cpu().numpy(), 1e-5, 1e-7)


@pytest.mark.parametrize("T,L,V", [(120, 40, 50), (600, 200, 300), (2000, 200, 260), (1100, 520, 600), (40, 1, 9)])
def test_forced_align_sizes_vs_oracle(T, L, V):
    """Forced alignment against the numpy restatement of the reference's `forced_align` (pinned to
    the reference's own outputs in test_oracle_golden) at sizes that take every path of the kernel:
    packed back-pointers in shared memory, in the global workspace (T=2000: they no longer fit),
    more states than threads (L=520: 1041 states), and a single label."""
    from espnet1_b200 import functional as F
    from oracle import ctc_numpy

    d = _dev()
    g = torch.Generator().manual_seed(T + L)
    lp = torch.log_softmax(torch.randn(T, V, generator=g) * 2, -1)
    y = torch.randint(1, V, (L,), generator=g)
    if L > 4:
        y[3] = y[2]  # a repeated label (no skip transition across it)
    want = ctc_numpy.viterbi_align(lp.numpy(), y.numpy(), 0)
    got = F.forced_align(lp.to(d), y.to(d))
    np.testing.assert_array_equal(got.cpu().numpy(), np.asarray(want))


def test_training_projection_with_the_bias_as_an_input_column():
    """`CTC.forward` runs `ctc_lo` as one GEMM over [hs, 1] x [W, b]^T in fp32 (no separate bias
    pass): loss and the gradients of hs, ctc_lo.weight and ctc_lo.bias equal those of the plain
    `ctc_lo(hs)` formulation; under autocast / for 16-bit weights the plain Linear is used."""
    from espnet1_b200 import functional as F
    from espnet1_b200.espnet2.asr.ctc import CTC

    d = _dev()
    g = torch.Generator().manual_seed(17)
    Bn, T, D, V = 5, 40, 24, 50
    m = CTC(V, D).to(d)
    hs = torch.randn(Bn, T, D, generator=g).to(d).requires_grad_(True)
    hl = torch.tensor([40, 33, 40, 21, 37], device=d)
    yl = torch.tensor([6, 3, 9, 2, 5], device=d)
    ys = torch.randint(1, V, (Bn, 9), generator=g).to(d)
    loss = m(hs, hl, ys, yl)
    loss.backward()
    got = (float(loss.detach()), hs.grad.clone(), m.ctc_lo.weight.grad.clone(), m.ctc_lo.bias.grad.clone())
    hs.grad = None
    m.zero_grad()
    ys_true = torch.cat([ys[i, : int(yl[i])] for i in range(Bn)])
    ref = F.ctc_loss(m.ctc_lo(hs), ys_true, hl, yl, reduce_scale=F._f32_recip(Bn))
    ref.backward()
    assert abs(got[0] - float(ref.detach())) <= 1e-6 * abs(float(ref.detach()))
    close(got[1], hs.grad.cpu().numpy(), 1e-5, 1e-6)
    close(got[2], m.ctc_lo.weight.grad.cpu().numpy(), 1e-5, 1e-6)
    close(got[3], m.ctc_lo.bias.grad.cpu().numpy(), 1e-5, 1e-6)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        assert m._logits(hs.detach()).dtype == torch.bfloat16  # the plain Linear under autocast


def test_c1_logits_boundary_vs_fixture_and_oracle():
    from espnet1_b200 import functional as F
    from oracle import ctc_numpy

    g = load_npz("ctc_c1.npz")
    d = _dev()
    x = torch.tensor(g["logits"], device=d).requires_grad_(True)
    nll = F.ctc_loss(x, torch.tensor(g["targets"]), torch.tensor(g["in_lens"]),
                     torch.tensor(g["tgt_lens"]))
    close(nll, g["nll"], LOSS_RTOL, 0)
    (nll.sum() / x.shape[0]).backward()
    close(x.grad, g["grad"], 0, GRAD_ATOL)
    # attribution against the fp64 oracle: the CUDA lattice is MORE accurate
    # than the reference's fp32 log-space one
    off = np.concatenate(([0], np.cumsum(g["tgt_lens"])))
    tg = [g["targets"][off[i]: off[i + 1]] for i in range(len(g["tgt_lens"]))]
    n64, g64 = ctc_numpy.ctc_loss_and_grad(g["logits"], tg, g["in_lens"],
                                           grad_out=1.0 / len(tg), dtype=np.float64)
    err_ours = np.abs(x.grad.cpu().numpy() - g64).max()
    err_ref = np.abs(g["grad"] - g64).max()
    assert err_ours <= max(2 * err_ref, 2e-6), (err_ours, err_ref)
    close(nll.detach().cpu().numpy().astype(np.float64), n64, 2e-6, 0)


def test_forward_multi_equals_sequential_calls():
    """Intermediate-CTC heads (espnet_model.py:243-286) as ONE batched pass over heads x utterances
    give the losses and gradients of one ``forward`` call per head - with equal and with different
    loss weights per head (with and without the matching ``upstream_scales``)."""
    import torch

    from espnet1_b200.espnet2.asr.ctc import CTC

    dev = "cuda:0"
    g = torch.Generator().manual_seed(9)
    B, T, D, V = 6, 70, 16, 40
    hs = [torch.randn(B, T, D, generator=g).to(dev).requires_grad_(True) for _ in range(3)]
    hl = torch.tensor([70, 64, 70, 33, 70, 50], device=dev)
    yl = torch.tensor([9, 4, 0, 12, 7, 5], device=dev)
    ys = torch.randint(1, V, (B, 12), generator=g).to(dev)
    ctc = CTC(V, D).to(dev)
    for wts, scales in (([1.0, 1.0, 1.0], None), ([0.35, 0.15, 0.5], [0.35, 0.15, 0.5]),
                        ([0.35, 0.15, 0.5], None)):
        for h in hs:
            h.grad = None
        ctc.zero_grad(set_to_none=True)
        want = [ctc(h, hl, ys, yl) for h in hs]
        sum(w * l for w, l in zip(wts, want)).backward()
        gw = [h.grad.clone() for h in hs] + [ctc.ctc_lo.weight.grad.clone(), ctc.ctc_lo.bias.grad.clone()]
        for h in hs:
            h.grad = None
        ctc.zero_grad(set_to_none=True)
        got = ctc.forward_multi(hs, hl, ys, yl, upstream_scales=scales)
        sum(w * l for w, l in zip(wts, got)).backward()
        torch.cuda.synchronize()
        for a, b_ in zip(got, want):
            assert abs(float(a.detach()) - float(b_.detach())) <= 2e-6 * abs(float(b_.detach()))
        for a, h in zip(gw[:3], hs):
            torch.testing.assert_close(h.grad, a, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(ctc.ctc_lo.weight.grad, gw[3], rtol=1e-5, atol=2e-6)
        torch.testing.assert_close(ctc.ctc_lo.bias.grad, gw[4], rtol=1e-5, atol=2e-6)
    assert len(ctc.forward_multi(hs[:1], hl, ys, yl)) == 1
    ctc.reduce = False
    a = ctc.forward_multi(hs, hl, ys, yl)
    b_ = [ctc(h, hl, ys, yl) for h in hs]
    for x, y in zip(a, b_):
        torch.testing.assert_close(x, y, rtol=2e-6, atol=1e-6)
    # the round-1 stream-alternation form stays bit-identical to sequential calls
    ctc.reduce = True
    c = ctc.forward_multi_streams(hs, hl, ys, yl)
    torch.cuda.synchronize()
    assert [float(v.detach()) for v in c] == [float(v.detach()) for v in want]


def test_upstream_scale_hint_matches_weighted_loss():
    """`loss = ctc_weight * ctc(...)` (espnet2/asr/espnet_model.py): with `upstream_scale` set to the
    weight the pre-scaled gradient needs no rescaling pass, without it the rescale makes it exact -
    both must give the reference's gradient."""
    from oracle import espnet_cpu

    d = _dev()
    g = torch.Generator().manual_seed(3)
    Bn, T, D, V, w = 5, 60, 12, 40, 0.3
    hs = torch.randn(Bn, T, D, generator=g)
    hl = torch.tensor([60, 50, 60, 41, 33])
    yl = torch.tensor([7, 3, 9, 1, 5])
    ys = torch.randint(1, V, (Bn, 9), generator=g)
    from espnet1_b200.espnet2.asr.ctc import CTC

    m = CTC(V, D).to(d)
    xr = hs.clone().requires_grad_(True)
    ref = w * espnet_cpu.ctc_forward(xr, hl, ys, yl, m.ctc_lo.weight.detach().cpu(), m.ctc_lo.bias.detach().cpu())
    ref.backward()
    for scale in (1.0, w):
        m.upstream_scale = scale
        x = hs.to(d).requires_grad_(True)
        loss = w * m(x, hl.to(d), ys.to(d), yl.to(d))
        loss.backward()
        close(loss, ref.detach().numpy(), LOSS_RTOL, 1e-6)
        close(x.grad, xr.grad.numpy(), 0, GRAD_ATOL)


def test_forced_align_and_greedy_decode_vs_reference_fixture():
    """The CUDA forced alignment (workspace and static-scratch entry points) and greedy decode
    against outputs of the REAL reference (tests/golden/align_greedy.npz): five alignment shapes
    incl. repeated labels and a target nearly as long as the input; the collapse of
    ``ErrorCalculator.calculate_cer_ctc``.  Integer-identical."""
    from espnet1_b200 import _lib
    from espnet1_b200 import functional as F

    d = _dev()
    g = load_npz("align_greedy.npz")
    for i in range(int(g["align_n"])):
        lp = torch.tensor(g[f"align{i}__lp"], device=d)
        y = torch.tensor(g[f"align{i}__y"], device=d)
        ali = F.forced_align(lp, y)
        np.testing.assert_array_equal(ali.cpu().numpy(), g[f"align{i}__ali"], err_msg=f"case {i}")
        out = torch.empty(lp.shape[0], dtype=torch.int64, device=d)
        lpc = lp.contiguous()
        _lib.check(_lib.lib().b200ctc_forced_align_static(lpc.data_ptr(), y.data_ptr(), lp.shape[0], lp.shape[1],
                                                          y.numel(), 0, out.data_ptr(), _lib.stream_ptr(lp.device)),
                   "forced_align_static")
        np.testing.assert_array_equal(out.cpu().numpy(), g[f"align{i}__ali"], err_msg=f"static, case {i}")
    ids, lens = F.greedy_decode(torch.tensor(g["greedy__logits"], device=d))
    np.testing.assert_array_equal(lens.cpu().numpy(), g["greedy__lens"])
    ids = ids.cpu().numpy()
    for b, n in enumerate(g["greedy__lens"]):
        np.testing.assert_array_equal(ids[b, :n], g["greedy__ids"][b, :n])
        assert (ids[b, n:] == -1).all()
