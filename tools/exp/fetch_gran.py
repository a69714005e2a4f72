#!/usr/bin/env python3
"""Experiment: does cudaLimitMaxL2FetchGranularity (64 B default) change the sparse emission
gather (64-byte DRAM atoms for 4-byte reads) or the dense row pass?  C2, per-phase CUDA events.
  python tools/exp/fetch_gran.py [32|64|128]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import C2, synth  # noqa: E402
from espnet1_b200 import _lib  # noqa: E402
from espnet1_b200 import functional as F  # noqa: E402

gran = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda:0")
torch.zeros(1, device=dev)
rt = ctypes.CDLL("libcudart.so.12")
val = ctypes.c_size_t()
rt.cudaDeviceGetLimit(ctypes.byref(val), 5)
print("limit before:", val.value)
rc = rt.cudaDeviceSetLimit(5, ctypes.c_size_t(gran))
rt.cudaDeviceGetLimit(ctypes.byref(val), 5)
print("set", gran, "rc", rc, "limit now:", val.value)
cfg = dict(C2)
logits, tg, il, tl = synth(cfg, 1234)
x = logits.to(dev).requires_grad_(True)
tg, il, tl = tg.to(dev), il.to(dev), tl.to(dev)
L = _lib.lib()


def one():
    x.grad = None
    F.ctc_loss(x, tg, il, tl, max_tgt_len=cfg["lmax"], reduce_scale=1.0 / cfg["B"]).backward()


for _ in range(5):
    one()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(30):
    one()
e1.record()
torch.cuda.synchronize()
print("gran", gran, "ms/step", e0.elapsed_time(e1) / 30)
L.b200ctc_profile_enable(1)
for _ in range(10):
    one()
torch.cuda.synchronize()
L.b200ctc_profile_enable(0)
pm, pc = (ctypes.c_float * 8)(), (ctypes.c_int * 8)()
L.b200ctc_profile_read(pm, pc, 8)
names = ["prep", "gather", "lattice", "grad_rows", "prefix", "softmax", "rescale", "fixup"]
print("gran", gran, {n: round(pm[i] / pc[i], 4) for i, n in enumerate(names) if pc[i]})
