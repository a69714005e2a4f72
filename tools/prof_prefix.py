#!/usr/bin/env python3
"""Prefix-scoring step timing (C3 variants): wall per step and library kernel time."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from espnet1_b200 import _lib

dev = torch.device("cuda:0")
L = _lib.lib()
out = bench.bench_prefix(dev, steps=int(sys.argv[1]) if len(sys.argv) > 1 else 20)
for k, v in out.items():
    print(k, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items()})
L.b200ctc_profile_enable(1)
bench.bench_prefix(dev, steps=10)
L.b200ctc_profile_enable(0)
pm, pc = (ctypes.c_float * 8)(), (ctypes.c_int * 8)()
L.b200ctc_profile_read(pm, pc, 8)
print("prefix_step library time per call (all variants mixed): %.4f ms over %d calls" % (pm[4] / max(pc[4], 1), pc[4]))
