"""Worker of tests/test_gpu_sharded.py: one process per GPU (torchrun), NCCL group.
The fused peer-memory loss exchange must give exactly what the reference's all-reduce
of [sum of losses, utterance count] gives (trainer.py:604-613), on every rank, for
ragged per-rank batch sizes, repeatedly."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from espnet1_b200 import functional as F
from espnet1_b200.sharded import PeerLossExchange, reduce_sharded_loss

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
world = dist.get_world_size()
ex = PeerLossExchange()
for it in range(6):
    B = 3 + rank + it  # different batch size on every rank and step
    g = torch.Generator().manual_seed(100 * it + rank)
    T, V = 40 + 3 * it, 50
    x = (torch.randn(B, T, V, generator=g) * 2).to(dev).requires_grad_(True)
    tl = torch.randint(2, 9, (B,), generator=g)
    il = torch.full((B,), T)
    tg = torch.randint(1, V, (int(tl.sum()),), generator=g)
    loss = F.ctc_loss(x, tg, il, tl, reduce_scale=F._f32_recip(B), peer=ex)
    got = ex.out.clone()
    loss.backward()
    with torch.no_grad():
        nll = F.ctc_loss(x.detach(), tg, il, tl)
    want, wt = reduce_sharded_loss(nll)
    # every rank sees the same bits
    allv = [torch.empty_like(got) for _ in range(world)]
    dist.all_gather(allv, got)
    assert all(torch.equal(v, allv[0]) for v in allv), allv
    assert abs(float(got[1]) - float(wt)) < 0.5, (got, wt)
    assert abs(float(got[0]) - float(want)) <= 2e-6 * abs(float(want)), (got, want)
# the G-rank reduced loss equals the one-process loss over the whole batch
from espnet1_b200.sharded import shard_range  # noqa: E402

g = torch.Generator().manual_seed(4242)
Bg, T, V = 4 * world + 3, 60, 80  # not divisible by the world size: ragged shards
xg = torch.randn(Bg, T, V, generator=g) * 2
tlg = torch.randint(3, 12, (Bg,), generator=g)
ilg = torch.randint(40, T + 1, (Bg,), generator=g)
offs = torch.cumsum(tlg, 0) - tlg
tgg = torch.randint(1, V, (int(tlg.sum()),), generator=g)
lo, hi = shard_range(Bg, rank, world)
xs = xg[lo:hi].to(dev).requires_grad_(True)
tgs = tgg[int(offs[lo]):int(offs[hi - 1] + tlg[hi - 1])]
F.ctc_loss(xs, tgs, ilg[lo:hi], tlg[lo:hi], reduce_scale=F._f32_recip(hi - lo), peer=ex)
with torch.no_grad():
    whole = F.ctc_loss(xg.to(dev), tgg, ilg, tlg).double().mean()
assert abs(float(ex.out[0]) - float(whole)) <= 1e-6 * abs(float(whole)), (ex.out, whole)
assert float(ex.out[1]) == Bg
# a rank without utterances joins through the stand-alone call
pair = torch.tensor([float(rank + 1), 1.0], device=dev)
out = ex.exchange(pair).clone()
assert abs(float(out[1]) - world) < 1e-6 and abs(float(out[0]) - (world + 1) / 2) < 1e-5, out
assert not ex.timed_out()
torch.cuda.synchronize()
dist.barrier()
ex.close()
dist.destroy_process_group()
print(f"rank {rank} ok", flush=True)
