#!/usr/bin/env python3
"""Does F.linear without bias avoid cuBLASLt's separate epilogue kernel at the CTC projection
shape (16000 x 256 -> 5000, fp32)?  Device time of both forms."""
import torch
from torch.profiler import ProfilerActivity, profile

x = torch.randn(32, 500, 256, device="cuda")
w = torch.randn(5000, 256, device="cuda")
b = torch.randn(5000, device="cuda")
for name, fn in (("linear+bias", lambda: torch.nn.functional.linear(x, w, b)),
                 ("linear no bias", lambda: torch.nn.functional.linear(x, w)),
                 ("matmul", lambda: torch.matmul(x, w.t()))):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
    rows = [(e.key[:70], e.device_time_total / 5) for e in prof.key_averages() if e.device_time_total > 0]
    print(name, [(k, round(t, 1)) for k, t in rows])
