"""Batch-sharded loss: the peer-memory exchange fused into the loss kernel."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("deferred", [True, False])
def test_single_rank_exchange_is_the_local_mean(deferred):
    from espnet1_b200 import functional as F
    from espnet1_b200.sharded import PeerLossExchange

    ex = PeerLossExchange(deferred=deferred)
    g = torch.Generator().manual_seed(3)
    x = (torch.randn(5, 30, 40, generator=g) * 2).to(DEV).requires_grad_(True)
    tl = torch.tensor([3, 5, 2, 7, 4]); il = torch.full((5,), 30)
    tg = torch.randint(1, 40, (int(tl.sum()),), generator=g)
    for _ in range(3):  # the step counter / parity slots are reused call after call
        loss = F.ctc_loss(x, tg, il, tl, reduce_scale=F._f32_recip(5), peer=ex)
        assert float(ex.out[1]) == 5.0
        assert abs(float(ex.out[0]) - float(loss.detach())) <= 1e-6 * abs(float(loss.detach()))
    out = ex.exchange(torch.tensor([6.0, 3.0]))
    assert out.tolist() == [2.0, 3.0]
    assert ex.out.tolist() == [2.0, 3.0]
    # a call that cannot use the fused form (no gradient wanted: validation) makes the same
    # collective through the stand-alone kernel instead of raising
    with torch.no_grad():
        nll = F.ctc_loss(x.detach(), tg, il, tl, peer=ex)
    assert float(ex.out[1]) == 5.0
    assert abs(float(ex.out[0]) - float(nll.mean())) <= 2e-6 * abs(float(nll.mean()))
    lossd = F.ctc_loss(x, tg, il, tl, reduce_scale=F._f32_recip(5), peer=ex, grad_mode="deferred")
    assert abs(float(ex.out[0]) - float(lossd.detach())) <= 2e-6 * abs(float(lossd.detach()))
    assert not ex.timed_out()
    ex.raise_if_timed_out()
    ex.close()


def test_deferred_exchange_reads_the_latest_step():
    """Steps issued back to back WITHOUT reading `out` in between (the training loop): each
    exchange consumes its predecessor first; reading `out` afterwards gives the LAST step."""
    from espnet1_b200 import functional as F
    from espnet1_b200.sharded import PeerLossExchange

    ex = PeerLossExchange()  # deferred by default
    assert ex.deferred
    g = torch.Generator().manual_seed(4)
    losses = []
    for it in range(5):
        Bn = 3 + it
        x = (torch.randn(Bn, 25, 30, generator=g) * 2).to(DEV).requires_grad_(True)
        tl = torch.randint(1, 6, (Bn,), generator=g)
        il = torch.full((Bn,), 25)
        tg = torch.randint(1, 30, (int(tl.sum()),), generator=g)
        loss = F.ctc_loss(x, tg, il, tl, reduce_scale=F._f32_recip(Bn), peer=ex)
        loss.backward()
        losses.append((float(loss.detach()), Bn))
    want, n = losses[-1]
    got = ex.out.tolist()
    assert got[1] == float(n) and abs(got[0] - want) <= 1e-6 * abs(want)
    assert ex.out.tolist() == got  # idempotent: nothing left to consume
    ex.close()


def test_forward_multi_with_an_exchange_attached_exchanges_once_per_head():
    """Sharded training + intermediate CTC: `forward_multi` must not bypass `loss_exchange` - every
    head makes its own exchange (the batched single call has no batch reduction to exchange)."""
    from espnet1_b200.espnet2.asr.ctc import CTC
    from espnet1_b200.sharded import PeerLossExchange

    g = torch.Generator().manual_seed(6)
    Bn, T, D, V = 4, 40, 8, 30
    hs = [torch.randn(Bn, T, D, generator=g).to(DEV).requires_grad_(True) for _ in range(3)]
    hl = torch.tensor([40, 31, 40, 22], device=DEV)
    yl = torch.tensor([5, 2, 7, 3], device=DEV)
    ys = torch.randint(1, V, (Bn, 7), generator=g).to(DEV)
    m = CTC(V, D).to(DEV)
    want = [float(m(h, hl, ys, yl).detach()) for h in hs]
    m.loss_exchange = PeerLossExchange()
    got = m.forward_multi(hs, hl, ys, yl)
    sum(got).backward()
    assert [float(v.detach()) for v in got] == want
    out = m.loss_exchange.out.tolist()  # the LAST head's exchange
    assert out[1] == float(Bn) and abs(out[0] - want[-1]) <= 1e-6 * abs(want[-1])
    m.loss_exchange.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_ranks_match_the_nccl_allreduce():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29533",
           os.path.join(ROOT, "tests", "_sharded_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=240, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert out.stdout.count("ok") == 2


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_tensors_on_a_non_current_device():
    """tensors on cuda:1 while cuda:0 is the current device (a model moved with
    .to('cuda:1'), no set_device).  Every wrapper must launch on the TENSOR's device."""
    import numpy as np

    from espnet1_b200 import functional as F
    from espnet1_b200.espnet.nets.ctc_prefix_score import CTCPrefixScoreTH
    from oracle import espnet_cpu
    from oracle.prefix_torch import PrefixOracle

    torch.cuda.set_device(0)
    d1 = torch.device("cuda", 1)
    g = torch.Generator().manual_seed(12)
    logits = torch.randn(4, 50, 64, generator=g) * 2
    tl = torch.tensor([5, 9, 3, 7]); il = torch.tensor([50, 44, 50, 39])
    tg = torch.randint(1, 64, (int(tl.sum()),), generator=g)
    x = logits.to(d1).requires_grad_(True)
    nll = F.ctc_loss(x, tg, il, tl)
    (nll.sum() / 4).backward()
    rn, rg = espnet_cpu.logits_loss_and_grad(logits, tg, il, tl)
    np.testing.assert_allclose(nll.detach().cpu().numpy(), rn.numpy(), rtol=1e-5)
    np.testing.assert_allclose(x.grad.cpu().numpy(), rg.numpy(), rtol=0, atol=1e-4)
    assert x.grad.device == d1 and torch.cuda.current_device() == 0
    y = F.log_softmax(logits.to(d1))
    np.testing.assert_allclose(y.cpu().numpy(), torch.log_softmax(logits, -1).numpy(), rtol=1e-5, atol=1e-6)
    assert torch.equal(F.argmax(logits.to(d1)).cpu(), logits.argmax(-1))
    lpz = torch.log_softmax(logits, -1)
    xl = torch.tensor([50, 44, 50, 39])
    sc = CTCPrefixScoreTH(lpz.clone().to(d1), xl, 0, 63)
    ora = PrefixOracle(lpz.clone(), xl, 0, 63)
    yy = [[63] for _ in range(8)]
    ids = torch.randint(1, 63, (8, 6), generator=g)
    got, _ = sc(yy, None, ids.to(d1))
    want, _ = ora.score(yy, None, ids)
    np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=2e-6, atol=2e-5)
    assert torch.cuda.current_device() == 0
