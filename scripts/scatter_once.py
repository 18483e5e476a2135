"""Times lipvq_ema_stats_f32 and lipvq_gather_ste_loss_bwd_f32 at the bench shape (developer helper; ncu target).
usage: python scripts/scatter_once.py [N K D] [--collapsed]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
import lipvq  # noqa: E402
from lipvq import ops  # noqa: E402

a = [v for v in sys.argv[1:] if not v.startswith("--")]
N, K, D = (int(v) for v in a[:3]) if len(a) >= 3 else (65536, 512, 64)
g = torch.Generator().manual_seed(0)
z = torch.rand(N, D, generator=g).cuda()
idx = torch.randint(0, 5 if "--collapsed" in sys.argv else K, (N,), generator=g).cuda()
zq = torch.rand(N, D, generator=g).cuda()
up = torch.randn(N, D, generator=g).cuda()
dz = torch.empty_like(z)
counts, sums, dcb = torch.zeros(K, device="cuda"), torch.zeros(K, D, device="cuda"), torch.zeros(K, D, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ws = torch.empty(ops.code_sort_workspace_bytes(N, D, K), dtype=torch.uint8, device="cuda")
for name, fn in (("code_sort", lambda: ops.code_sort(idx, D, K, ws)),
                 ("ema_stats_sorted", lambda: ops.ema_stats_sorted(z, ws, counts, sums)),
                 ("gather_bwd_sorted", lambda: ops.gather_bwd_sorted(z, zq, up, None, 0.25, 0.25, False, z, K, ws, dz, dcb)),
                 ("ema_stats", lambda: ops.ema_stats(z, idx, counts, sums)),
                 ("gather_bwd", lambda: ops.gather_bwd(z, zq, idx, up, None, 0.25, 0.25, False, z, K, dz, dcb))):
    ds = []
    for i in range(8):
        flush.zero_()
        a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a_.record(); fn(); b_.record(); torch.cuda.synchronize()
        if i >= 2:
            ds.append(a_.elapsed_time(b_) * 1e3)
    print(f"{name}: mean {sum(ds) / len(ds):.1f} us  min {min(ds):.1f} us")
