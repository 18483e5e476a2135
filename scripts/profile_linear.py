"""Three launches each of the dual forward (64 -> 128, GELU value + derivative), the input gradient through the stored
derivative and the weight gradient at the bench's token count: the target of `ncu -k regex:linear_tc` captures.
usage: python scripts/profile_linear.py"""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
import lipvq  # noqa: E402
from lipvq import ops  # noqa: E402

N, i, o = 65536, 64, 128
g = torch.Generator().manual_seed(0)
X = torch.randn(N, i, generator=g).cuda()
W = (torch.randn(o, i, generator=g) / math.sqrt(i)).cuda()
b = torch.randn(o, generator=g).cuda()
Y, dYd = torch.empty(N, o, device="cuda"), torch.empty(N, o, device="cuda")
dY = torch.randn(N, o, generator=g).cuda()
dX = torch.empty(N, i, device="cuda")
dW, db = torch.zeros(o, i, device="cuda"), torch.zeros(o, device="cuda")
for _ in range(3):
    ops.linear_fwd_dual(X, lipvq.ACT_NONE, W, b, lipvq.ACT_GELU, Y, dYd)      # launches 0-2
for _ in range(3):
    ops.linear_bwd_x(dY, W, X, lipvq.ACT_GIVEN, out=dX)                        # launches 3-5
for _ in range(3):
    ops.linear_bwd_w(dY, X, lipvq.ACT_NONE, dW, db)                            # linear_tc_bwd_w launches 0-2
torch.cuda.synchronize()
X12 = torch.randn(N, 12, generator=g).cuda()
W12 = (torch.randn(64, 12, generator=g) / math.sqrt(12)).cuda()
b12 = torch.randn(64, generator=g).cuda()
Y64, dY64 = torch.empty(N, 64, device="cuda"), torch.empty(N, 64, device="cuda")
W128x12 = (torch.randn(12, 128, generator=g) / math.sqrt(12)).cuda()   # layer 128 -> 12: dX[N,128] = dY[N,12] @ W[12,128]
if "--time" in sys.argv:
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    mode = next((a.split("=")[1] for a in sys.argv if a.startswith("--flush=")), "write")  # write | read | none
    for name, fn in (("fwd_dual 64->128 gelu", lambda: ops.linear_fwd_dual(X, lipvq.ACT_NONE, W, b, lipvq.ACT_GELU, Y, dYd)),
                     ("bwd_x given", lambda: ops.linear_bwd_x(dY, W, X, lipvq.ACT_GIVEN, out=dX)),
                     ("fwd_dual 12->64 gelu", lambda: ops.linear_fwd_dual(X12, lipvq.ACT_NONE, W12, b12, lipvq.ACT_GELU, Y64, dY64)),
                     ("fwd 128->64 sigmoid", lambda: ops.linear_fwd(dY, lipvq.ACT_NONE, W.t().contiguous(), b[:64].contiguous(), lipvq.ACT_SIGMOID, out=dX)),
                     ("bwd_x 12->128 given", lambda: ops.linear_bwd_x(X12, W128x12, Y, lipvq.ACT_GIVEN, out=dYd)),
                     ("bwd_w", lambda: ops.linear_bwd_w(dY, X, lipvq.ACT_NONE, dW, db))):
        ds = []
        for i in range(12):
            if mode == "write":
                flush.zero_()           # leaves L2 full of DIRTY lines: the timed kernel also pays their write-back
            elif mode == "read":
                flush.view(torch.int32).sum()   # leaves L2 full of clean lines
            a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); e.record(); torch.cuda.synchronize()
            if i >= 2:
                ds.append(a.elapsed_time(e) * 1e3)
        print(f"[flush={mode}] {name}: mean {sum(ds) / len(ds):.1f} us  min {min(ds):.1f} us")
