"""Batch-sharded loss: the peer-memory exchange fused into the loss kernel."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_single_rank_exchange_is_the_local_mean():
    from espnet1_b200 import functional as F
    from espnet1_b200.sharded import PeerLossExchange

    ex = PeerLossExchange()
    g = torch.Generator().manual_seed(3)
    x = (torch.randn(5, 30, 40, generator=g) * 2).to(DEV).requires_grad_(True)
    tl = torch.tensor([3, 5, 2, 7, 4]); il = torch.full((5,), 30)
    tg = torch.randint(1, 40, (int(tl.sum()),), generator=g)
    for _ in range(3):  # the step counter / parity slots are reused call after call
        loss = F.ctc_loss(x, tg, il, tl, reduce_scale=F._f32_recip(5), peer=ex)
        assert float(ex.out[1]) == 5.0
        assert abs(float(ex.out[0]) - float(loss)) <= 1e-6 * abs(float(loss))
    out = ex.exchange(torch.tensor([6.0, 3.0]))
    assert out.tolist() == [2.0, 3.0]
    assert not ex.timed_out()
    with pytest.raises(RuntimeError):  # needs the eager gradient path
        F.ctc_loss(x.detach(), tg, il, tl, peer=ex)
    ex.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_ranks_match_the_nccl_allreduce():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29533",
           os.path.join(ROOT, "tests", "_sharded_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=240, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert out.stdout.count("ok") == 2
