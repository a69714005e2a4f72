#!/usr/bin/env python3
"""Where does the N>1 step time go?  torchrun --nproc-per-node 2 tools/mgpu_probe.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from bench import C2, synth
from espnet1_b200.espnet2.asr.ctc import CTC
from espnet1_b200.sharded import pack_local_mean, reduce_sharded_mean

rank = int(os.environ.get("RANK", 0)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
cfg = C2
B, T, V = cfg["B"], cfg["T"], cfg["V"]
lg, tg, il, tl = synth(cfg, 1234 + rank)
x = lg.to(dev).requires_grad_(True)
tg, il, tl = tg.to(dev), il.to(dev), tl.to(dev)
m = CTC(V, 8).to(dev)
wbuf = torch.zeros(2, device=dev)


def step(mode):
    x.grad = None
    loss = m.loss_fn(x.transpose(0, 1), tg, il, tl, max_tgt_len=cfg["lmax"])
    if mode == 1:
        pack_local_mean(loss, B)
    if mode >= 2:
        reduce_sharded_mean(loss, B)
    loss.backward()
    return loss


def timeit(fn, n=30):
    for _ in range(5):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    c0 = time.process_time()
    for _ in range(n):
        fn()
    cpu_issue = (time.perf_counter() - t0) / n
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3, cpu_issue * 1e3


from espnet1_b200.sharded import PeerLossExchange
ex = PeerLossExchange()
for mode, name in [(0, "no dist ops"), (1, "pack only"), (2, "pack + all_reduce"), (3, "fused peer exchange")]:
    m.loss_exchange = ex if mode == 3 else None
    ms, issue = timeit(lambda: step(mode))
    if rank == 0:
        print(f"{name:22s}: {ms:.3f} ms/step (CPU issue {issue:.3f} ms)", flush=True)

dist.destroy_process_group()
