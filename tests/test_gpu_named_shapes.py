"""GPU parity at the shapes BASELINE.json NAMES ("parity gaps"): C4 at V=5000,
C5 per-rank and single-GPU batch sizes, C3 at T=500 (pre-beam and full vocabulary), soft-max
backward at V=5000, and the attention-window boundary.  The checker is the torch-CPU port of the
reference path (oracle.espnet_cpu / oracle.prefix_torch); where the CPU cost would exceed about a
minute the CPU side runs on a SLICE of the utterances (CTC utterances are independent) while the
GPU runs the whole named shape - each test says which."""
import numpy as np
import pytest
import torch

from oracle import espnet_cpu
from oracle.prefix_torch import PrefixOracle

from _util import close
from test_gpu_ctc_random import make_case
from test_gpu_prefix import ATOL, RTOL, _drive, _th

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cuda_loss_grad(x, tg, il, tl, lmax):
    from espnet1_b200 import functional as F

    x = x.requires_grad_(True)
    nll = F.ctc_loss(x, tg, il, tl, max_tgt_len=lmax)
    (nll.sum() / x.shape[0]).backward()
    return nll.detach(), x.grad.detach()


def test_c4_exact_shape_v5000_vs_fp64_truth():
    """C4 as named: B=16, T=4000, V=5000, L in [600,1200] on the GPU (split lattice + full-row
    emission pass + WIDE row pass together).  CPU side: the first 4 utterances, in fp64 AND fp32.
    At NLL ~ 3e4 torch's own fp32 lattice is ~1e-2 off per gradient element, so the yardstick is
    the fp64 run of the same torch path; the contract's 1e-4 absolute is asserted against it."""
    B, T, V, nb = 16, 4000, 5000, 4
    logits, tg, il, tl = make_case(B, T, V, 600, 1200, seed=4321, ragged=False)
    nll, g = _cuda_loss_grad(logits.to(DEV), tg, il, tl, 1200)
    nll, g = nll.cpu(), g[:nb].cpu()
    ntg = int(tl[:nb].sum())
    n64, g64 = espnet_cpu.logits_loss_and_grad(logits[:nb].double(), tg[:ntg], il[:nb], tl[:nb])
    n32, g32 = espnet_cpu.logits_loss_and_grad(logits[:nb], tg[:ntg], il[:nb], tl[:nb])
    g64, g32 = g64 * (nb / B), g32 * (nb / B)  # both sides carry 1/B of THEIR batch
    np.testing.assert_allclose(nll[:nb].numpy(), n64.numpy(), rtol=1e-5)
    err_ours = (g.double() - g64).abs().max().item()
    err_ref = (g32.double() - g64).abs().max().item()
    print(f"C4 V=5000: max |grad - fp64| ours {err_ours:.3e}, torch-fp32 {err_ref:.3e}")
    # measured on B200: ours 3.1e-05, torch-fp32 3.7e-03
    assert err_ours < 1e-4 and err_ours * 50 < err_ref, (err_ours, err_ref)
    assert torch.isfinite(nll).all() and (nll > 0).all()


@pytest.mark.parametrize("B,lo", [(128, 0), (1024, 480)])
def test_c5_batch_sizes_vs_cpu_slice(B, lo):
    """C5 shapes: 128 utterances (one rank's shard at N=8; 256 lattice CTAs > 148 SMs, the
    multi-wave regime) and all 1024 on one GPU (N=1).  Logits are generated on the device; the CPU
    port checks a 64-utterance slice [lo, lo+64) with the contract's tolerances (gradient compared
    at the 1/64 scale of a C2 batch)."""
    T, V, lmin, lmax, ns = 500, 5000, 75, 150, 64
    gd = torch.Generator(device=DEV).manual_seed(1234 + B)
    x = torch.randn(B, T, V, generator=gd, device=DEV) * 2.0
    g = torch.Generator().manual_seed(99 + B)
    tl = torch.randint(lmin, lmax + 1, (B,), generator=g)
    il = torch.randint(300, T + 1, (B,), generator=g)
    il[lo] = T
    tg = torch.randint(1, V, (int(tl.sum()),), generator=g)
    xs = x[lo:lo + ns].cpu()
    nll, gr = _cuda_loss_grad(x, tg, il, tl, lmax)
    nll, gs = nll.cpu(), (gr[lo:lo + ns] * (B / ns)).cpu()
    del x, gr
    torch.cuda.empty_cache()
    o0, o1 = int(tl[:lo].sum()), int(tl[:lo + ns].sum())
    rn, rg = espnet_cpu.logits_loss_and_grad(xs, tg[o0:o1], il[lo:lo + ns], tl[lo:lo + ns])
    np.testing.assert_allclose(nll[lo:lo + ns].numpy(), rn.numpy(), rtol=1e-5)
    np.testing.assert_allclose(gs.numpy(), rg.numpy(), rtol=0, atol=1e-4)
    assert torch.isfinite(nll).all()
    for b in range(lo, lo + ns, 7):
        assert torch.all(gs[b - lo, il[b]:] == 0)


@pytest.mark.parametrize("B,snum,margin,steps", [
    (32, 15, 0, 4),     # C3(i)  as named: batch 32, beam 10, V=5000, T=500, pre-beam 15
    (32, 15, 20, 4),    # C3(ii) windowed margin
    (8, None, 20, 2),   # C3(iii) full vocabulary + margin on 8 of the 32 utterances (the CPU
                        #         oracle's state is T*2*n_bh*V*4 = 1.6 GB per step at B=8)
])
def test_c3_prefix_scoring_at_t500(B, snum, margin, steps):
    W, V, T = 10, 5000, 500
    g = torch.Generator().manual_seed(31 + (snum or 0) + margin)
    lpz = torch.log_softmax(torch.randn(B, T, V, generator=g) * 2, dim=2)
    xlens = torch.randint(250, T + 1, (B,), generator=g)
    xlens[0] = T
    eos = V - 1
    ora = PrefixOracle(lpz.clone(), xlens, 0, eos, margin)
    cu = _th()(lpz.clone().to(DEV), xlens, 0, eos, margin)
    ref, yref = _drive(lambda *a: ora.score(*a), ora.select, B, W, V, T, snum, margin, steps, 7, "cpu")
    got, ygot = _drive(cu, cu.index_select_state, B, W, V, T, snum, margin, steps, 7, DEV)
    assert yref == ygot  # identical surviving token sequences after `steps` carried steps
    for (s0, p0, a0, b0), (s1, p1, a1, b1) in zip(ref, got):
        assert (a0, b0) == (a1, b1)  # f_min / f_max of the window
        close(s1, s0.numpy(), RTOL, ATOL)
        close(p1, p0.numpy(), RTOL, ATOL)


@pytest.mark.parametrize("dtype,rtol,atol", [(torch.float32, 1e-5, 1e-6), (torch.bfloat16, 2e-2, 2e-2)])
@pytest.mark.parametrize("mode", ["softmax", "log_softmax"])
def test_softmax_backward_vs_torch_autograd_v5000(mode, dtype, rtol, atol):
    """`ctc.softmax` runs INSIDE training encoders (self-conditioned CTC,
    espnet2/asr/encoder/conformer_encoder.py:360): its backward (b200ctc_softmax_bwd) against
    torch autograd on the same rounded inputs, V=5000, with a non-trivial upstream gradient."""
    from espnet1_b200 import functional as F

    g = torch.Generator().manual_seed(5)
    x = (torch.randn(6, 37, 5000, generator=g) * 2).to(dtype)
    up = torch.randn(6, 37, 5000, generator=g).to(dtype)
    xc = x.to(DEV).requires_grad_(True)
    y = getattr(F, mode)(xc)
    y.backward(up.to(DEV))
    xr = x.float().requires_grad_(True)
    yr = getattr(torch, mode)(xr, dim=-1)
    yr.backward(up.float())
    close(y.float(), yr.detach().numpy(), rtol, atol, f"{mode} forward")
    # backward: dx = dy - exp(y) * sum(dy) (softmax: y * (dy - sum(dy * y))).  sum(dy) runs over 5000
    # N(0,1) terms (|sum| up to a few hundred): fp32 summation ORDER alone moves it by ~1e-5, and
    # where dy and exp(y)*sum nearly cancel that is the whole error - hence the absolute 2e-5.
    close(xc.grad.float(), xr.grad.numpy(), rtol, 2e-5 if dtype == torch.float32 else 4 * atol, f"{mode} backward")
    assert xc.grad.dtype == dtype


def _window(att):
    from espnet1_b200 import _lib
    from espnet1_b200._lib import check, ptr, stream_ptr

    a = att.to(DEV).contiguous()
    mm = torch.empty(2, dtype=torch.float32, device=DEV)
    check(_lib.lib().b200ctc_prefix_window(ptr(a), a.shape[0], a.shape[1], ptr(mm), stream_ptr(a.device)), "window")
    lo, hi = mm.cpu().tolist()
    return int(lo), int(hi)


def test_window_boundary_exact_cases():
    """`f_arg = att_w @ frame_ids; int(f_arg.min()), int(f_arg.max())`
    (espnet/nets/ctc_prefix_score.py:145-147).  A value one fp32 step below an integer must
    truncate DOWN, a value exactly on it must stay.  These weights are dyadic, so every product
    and partial sum is exact in fp32 in ANY summation order (BLAS or ours, FMA or not): the
    reference's answer is order-independent and the comparison is strict."""
    T = 500
    frames = torch.arange(T, dtype=torch.float32)
    rows, want = [], []
    # smallest step below k that keeps k - eps, (1 - eps) * k and eps * (k - 1) exact in fp32
    for k, e0 in ((3, 2.0 ** -20), (15, 2.0 ** -20), (200, 2.0 ** -16), (498, 2.0 ** -15)):
        assert float(torch.tensor(k - e0, dtype=torch.float32)) == k - e0
        for eps, floor in ((0.0, k), (e0, k - 1), (2.0 ** -12, k - 1)):
            w = torch.zeros(T)
            w[k - 1], w[k] = eps, 1.0 - eps  # sum = k - eps, exactly representable
            rows.append(w)
            want.append(floor)
    att = torch.stack(rows)
    f_ref = torch.matmul(att, frames)
    assert [int(v) for v in f_ref.tolist()] == want  # the reference formula itself
    for i, w in enumerate(rows):  # one row at a time: min == max == that row
        assert _window(w[None]) == (want[i], want[i]), (i, float(f_ref[i]))
    assert _window(att) == (int(f_ref.min()), int(f_ref.max()))


def test_window_boundary_random_near_integers():
    """Rows aimed at integers +- a few fp32 ulps with 3-5 non-dyadic weights.  Where the fp64 value
    of the sum is more than 4 ulp(fp32) away from an integer no summation order can flip int();
    the kernel must agree with the reference formula there.  Closer rows are order-dependent in
    the REFERENCE itself (BLAS order is unspecified); they are counted and printed, not asserted."""
    T, n = 500, 4000
    g = torch.Generator().manual_seed(8)
    frames = torch.arange(T, dtype=torch.float32)
    att = torch.zeros(n, T, dtype=torch.float64)
    for i in range(n):
        k = int(torch.randint(3, 6, (1,), generator=g))
        idx = (torch.randperm(T - 2, generator=g)[:k] + 1).double()
        w = torch.rand(k, generator=g).double() + 0.05
        w = w / w.sum()
        c = float((w * idx).sum())
        near = float(round(c))
        ulp = 2.0 ** (int(np.floor(np.log2(max(near, 1.0)))) - 23)
        tgt = near + float(torch.randint(-60, 61, (1,), generator=g)) * ulp
        # move mass between the two outer frames so that the sum lands near `tgt`
        a, b = int(idx.argmin()), int(idx.argmax())
        d = (tgt - c) / float(idx[b] - idx[a])
        w[b] += d
        w[a] -= d
        if w.min() <= 0:
            continue
        att[i, idx.long()] = w
    att = att[att.sum(1) > 0].float()  # what the scorer receives
    f64 = torch.matmul(att.double(), frames.double())
    f32 = torch.matmul(att, frames)  # the reference formula, fp32 on CPU
    ulp = torch.pow(2.0, torch.floor(torch.log2(f64.abs().clamp_min(1.0))) - 23)
    dist = (f64 - f64.round()).abs() / ulp
    safe = dist > 8  # 3-5 terms: no summation order moves the fp32 sum by 8 ulp
    bad, flips = 0, 0
    for i in range(att.shape[0]):
        lo, hi = _window(att[i:i + 1])
        assert lo == hi
        if lo != int(f32[i]):
            if safe[i]:
                bad += 1
            else:
                flips += 1
    print(f"window probe: {att.shape[0]} rows aimed within 60 ulp of an integer; {int(safe.sum())} unambiguous "
          f"(> 8 ulp away), mismatches there: {bad}; {int((~safe).sum())} order-dependent rows, of which "
          f"{flips} differ from torch's CPU BLAS order")
    assert bad == 0
    assert int(safe.sum()) > 1000 and int((dist < 40).sum()) > 1000
