"""A few nearest-code searches at one shape (for ncu / timing of the individual launches).
usage: python scripts/search_once.py [N K D mode iters]"""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
import lipvq  # noqa: E402
from lipvq import ops  # noqa: E402

N, K, D = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (65536, 512, 64)
mode = sys.argv[4] if len(sys.argv) > 4 else "t1"
iters = int(sys.argv[5]) if len(sys.argv) > 5 and sys.argv[5].isdigit() else 6
g = torch.Generator().manual_seed(0)
cb = ((torch.rand(K, D, generator=g) * 2 - 1) * math.sqrt(6.0 / D)).cuda()
z = (torch.rand(N, D, generator=g) if mode == "uniform" else torch.sigmoid(0.05 * torch.randn(N, D, generator=g))).cuda()
nt = next((int(a.split("=")[1]) for a in sys.argv if a.startswith("--ties=")), 0)
if nt:  # nt rows sit (almost) midway between two codes: the fp16 pass cannot certify them (the cliff case of the N >= 2 runs)
    rows = torch.randperm(N, generator=g)[:nt]
    a_, b_ = torch.randint(0, K, (nt,), generator=g), torch.randint(0, K, (nt,), generator=g)
    mid = 0.5 * (cb.cpu()[a_] + cb.cpu()[b_]) + 1e-7 * torch.randn(nt, D, generator=g)
    z[rows.cuda()] = mid.cuda()
ws = torch.empty(ops.search_workspace_bytes(N, D, K, lipvq.METRIC_L2, lipvq.ALGO_TCGEN05), dtype=torch.uint8, device="cuda")
idx = torch.empty(N, dtype=torch.int64, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
rot = "--rotate" in sys.argv
if rot:  # inputs larger than L2: 16 different z buffers (268 MB) used in turn, no flush
    zs = [z.clone() + 0 * i for i in range(16)]
prep = "--prepared" in sys.argv  # codebook prepared once, then one launch per search (lipvq_nn_search_prepared_f32)
if prep:
    ops.nn_search_prepare(cb, ws, lipvq.METRIC_L2, center=z)
for i in range(iters):
    if rot:
        z = zs[i % 16]
    elif "--readflush" in sys.argv:
        flush.sum()  # evicts L2 by READING 256 MB: no dirty lines left behind
    else:
        flush.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    if prep:
        ops.nn_search_prepared(z, cb, ws, lipvq.METRIC_L2, out=idx)
    else:
        ops.nn_search(z, cb, lipvq.METRIC_L2, lipvq.ALGO_TCGEN05, out=idx, workspace=ws)
    b.record()
    torch.cuda.synchronize()
    print(f"iter {i}: {a.elapsed_time(b) * 1e3:.1f} us")
if prep:
    _, st = ops.nn_search_prepared(z, cb, ws, lipvq.METRIC_L2, want_stats=True)
else:
    _, st = ops.nn_search(z, cb, lipvq.METRIC_L2, lipvq.ALGO_TCGEN05, want_stats=True, workspace=ws)
print("stats", st.tolist())
