import torch, sys
sys.path.insert(0, ".")
from espnet1_b200 import functional as F
x = torch.randn(500, 5000, device="cuda"); lp = F.log_softmax(x)
tgt = torch.randint(1, 5000, (150,), device="cuda")
for _ in range(4): F.forced_align(lp, tgt)
torch.cuda.synchronize()
