import os, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import torch
from test_gpu_ctc_random import make_case
from espnet1_b200 import functional as F
dev = torch.device('cuda:0')
logits, tg, il, tl = make_case(64, 500, 5000, 75, 150, seed=1234)
x = logits.to(dev)
os.environ['B200CTC_WAVES'] = '1'
def run():
    xr = x.clone().requires_grad_(True)
    nll = F.ctc_loss(xr, tg, il, tl, max_tgt_len=150)
    (nll.sum() / 64).backward()
    return nll.detach().clone(), xr.grad.clone()
n1, g1 = run(); torch.cuda.synchronize()
side = torch.cuda.Stream()
a = torch.randn(4096, 4096, device=dev); y = torch.randn(64 * 1024 * 1024, device=dev)
bad = 0
for rep in range(60):
    with torch.cuda.stream(side):
        for _ in range(6):
            if rep % 2: b_ = a @ a
            else: y.mul_(1.0001)
    n, g = run()
    torch.cuda.synchronize()
    if not torch.equal(n, n1) or not torch.equal(g, g1):
        bad += 1
        rows = (g != g1).any(-1).nonzero()
        print('rep', rep, 'nll diff', (n != n1).nonzero().flatten().tolist()[:8], 'rows', rows.shape[0], 'utts', sorted(set(rows[:, 0].tolist()))[:8])
print('perturb done, bad =', bad)
