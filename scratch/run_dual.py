import math, sys, torch
sys.path.insert(0, "robot-manipulation-vq-vae_b200")
import lipvq
from lipvq import ops
N, i, o = 65536, 64, 128
g = torch.Generator().manual_seed(0)
X = torch.randn(N, i, generator=g).cuda(); W = (torch.randn(o, i, generator=g) / math.sqrt(i)).cuda(); b = torch.randn(o, generator=g).cuda()
Y, dYd = torch.empty(N, o, device="cuda"), torch.empty(N, o, device="cuda")
for _ in range(3):
    ops.linear_fwd_dual(X, 0, W, b, lipvq.ACT_GELU, Y, dYd)
dY = torch.randn(N, o, generator=g).cuda()
dX = torch.empty(N, i, device="cuda")
for _ in range(3):
    ops.linear_bwd_x(dY, W, torch.randn(N, i, device="cuda"), lipvq.ACT_GIVEN, out=dX)
torch.cuda.synchronize()
