import math, sys, torch
sys.path.insert(0, "robot-manipulation-vq-vae_b200")
import lipvq
from lipvq import ops
DEV = "cuda"
for (N, i, o, act) in [(4100, 128, 12, 1), (4100, 128, 12, 2), (4100, 128, 16, 1), (4100, 64, 12, 1), (4100, 128, 12, 0)]:
    g = torch.Generator().manual_seed(N + o)
    X = torch.randn(N, i, generator=g); W = torch.randn(o, i, generator=g) / math.sqrt(i); b = torch.randn(o, generator=g)
    pre = torch.nn.functional.linear(X, W, b)
    ref = {0: pre, 1: torch.relu(pre), 2: torch.nn.functional.gelu(pre)}[act]
    Y, dYd = torch.empty(N, o, device=DEV), torch.empty(N, o, device=DEV)
    ops.linear_fwd_dual(X.to(DEV), 0, W.to(DEV), b.to(DEV), act, Y, dYd)
    Y1 = ops.linear_fwd(X.to(DEV), 0, W.to(DEV), b.to(DEV), act)
    d = (Y.cpu() - ref).abs(); d1 = (Y1.cpu() - ref).abs()
    bad = (d > 1e-5).nonzero()
    print(N, i, o, act, "dual max err", d.max().item(), "single max err", d1.max().item(), "nbad", bad.shape[0], bad[:6].tolist())
    if bad.shape[0]:
        r, c = bad[0].tolist(); print("  got", Y[r, c].item(), "ref", ref[r, c].item(), "dYd", dYd[r, c].item())
