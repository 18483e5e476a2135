"""Times the token-major linear layers of the tokenizer (forward and backward-x) on both paths: tcgen05 three-way-bf16
(linear_tc.cu) and fp32 FFMA tiles (linear.cu).  CUDA events, L2 flushed between iterations."""
import json
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
import lipvq  # noqa: E402
from lipvq import ops  # noqa: E402

N = 65536
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, iters=12):
    ts = []
    for i in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(a.elapsed_time(b) * 1e3)
    return round(statistics.median(ts), 1)


for (i, o, act_in, act_out) in [(12, 64, 0, 0), (64, 128, 2, 0), (128, 64, 2, 3), (64, 64, 0, 0), (128, 12, 2, 0)]:
    X = torch.randn(N, i, device="cuda")
    W = torch.randn(o, i, device="cuda") / i ** 0.5
    b = torch.randn(o, device="cuda")
    dY = torch.randn(N, o, device="cuda")
    Y = torch.empty(N, o, device="cuda")
    dX = torch.empty(N, i, device="cuda")
    res = {"in": i, "out": o, "act_in": act_in, "act_out": act_out,
           "hbm_floor_us": round((N * (i + o) * 4) / 6.5e6, 1)}
    for algo in ("tc", "simt"):
        os.environ["LIPVQ_LINEAR_ALGO"] = algo
        res[f"fwd_{algo}_us"] = timeit(lambda: ops.linear_fwd(X, act_in, W, b, act_out, out=Y))
        res[f"bwdx_{algo}_us"] = timeit(lambda: ops.linear_bwd_x(dY, W, X, act_in, out=dX))
        dW, db = torch.zeros(o, i, device="cuda"), torch.zeros(o, device="cuda")
        res[f"bwdw_{algo}_us"] = timeit(lambda: ops.linear_bwd_w(dY, X, act_in, dW, db))
    print(json.dumps(res), flush=True)
