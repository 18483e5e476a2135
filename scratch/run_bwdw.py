import math, sys, torch
sys.path.insert(0, "robot-manipulation-vq-vae_b200")
import lipvq
from lipvq import ops
N, i, o = 65536, 64, 128
g = torch.Generator().manual_seed(0)
X = torch.randn(N, i, generator=g).cuda(); dY = torch.randn(N, o, generator=g).cuda()
dW, db = torch.zeros(o, i, device="cuda"), torch.zeros(o, device="cuda")
for _ in range(3):
    ops.linear_bwd_w(dY, X, 0, dW, db)
torch.cuda.synchronize()
