import cProfile, pstats, io, os, sys, json
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
import torch
import bench_beam
# plain numbers first
r = bench_beam.run(cpu_port=False)
print(json.dumps({k: v for k, v in r.items() if k.startswith("engine")}))
