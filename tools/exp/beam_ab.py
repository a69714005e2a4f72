#!/usr/bin/env python3
"""A/B of the beam-position kernels' launch shapes (tuning build only:
B200CTC_NVCC_EXTRA=-DB200CTC_TUNING python -m espnet1_b200.build --force)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
VARIANTS = [
    ("default", {}),
    ("no_scan", {"B200CTC_PREFIX_SCAN": "0"}),
    ("no_time_major", {"B200CTC_BEAM_TIME_MAJOR": "0"}),
    ("no_cluster", {"B200CTC_BEAM_CLUSTER": "0"}),
    ("all_old", {"B200CTC_BEAM_CLUSTER": "0", "B200CTC_PREFIX_SMALL_NT": "128", "B200CTC_PREFIX_AHEAD": "1",
                 "B200CTC_PREFIX_VEC": "0"}),
]
for name, env in VARIANTS:
    e = dict(os.environ, CASES="engine_batch32,engine_single", **env)
    best = {}
    for _ in range(2):
        out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "bench_beam.py")], env=e,
                             capture_output=True, text=True)
        if out.returncode != 0:
            print(name, "FAILED", out.stderr[-400:])
            break
        d = json.loads(out.stdout.strip().splitlines()[-1])
        for k in ("engine_batch32", "engine_single"):
            best[k] = min(best.get(k, 1e9), d[k]["ms_per_position"])
    print(name, {k: round(v, 4) for k, v in best.items()}, flush=True)
