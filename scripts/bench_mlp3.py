"""Encoder forward (12 -> 64 -> 128 -> 64, GELU GELU sigmoid) at N tokens: one fused launch vs three linear launches."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
import lipvq as L
from lipvq import ops
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
dev = "cuda"
g = torch.Generator().manual_seed(0)
dims = (12, 64, 128, 64)
xs = [(torch.rand(N, dims[0], generator=g) * 2 - 1).to(dev) for _ in range(8)]
Ws = [(torch.randn(o, i, generator=g) / i ** 0.5).to(dev) for i, o in zip(dims[:-1], dims[1:])]
bs = [(torch.randn(o, generator=g) * 0.1).to(dev) for o in dims[1:]]
a1, a2, y = (torch.empty(N, d, device=dev) for d in dims[1:])
acts = (L.ACT_GELU, L.ACT_GELU, L.ACT_SIGMOID)
def fused(x): ops.mlp3_fwd(x, Ws[0], bs[0], acts[0], Ws[1], bs[1], acts[1], Ws[2], bs[2], acts[2], out=y)
def three(x):
    ops.linear_fwd(x, L.ACT_NONE, Ws[0], bs[0], acts[0], out=a1)
    ops.linear_fwd(a1, L.ACT_NONE, Ws[1], bs[1], acts[1], out=a2)
    ops.linear_fwd(a2, L.ACT_NONE, Ws[2], bs[2], acts[2], out=y)
for name, fn in (("fused", fused), ("three", three)):
    for x in xs: fn(x)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for x in xs: fn(x)
    torch.cuda.current_stream().wait_stream(s)
    with torch.cuda.graph(gr):
        for x in xs: fn(x)
    gr.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): gr.replay()
    b.record(); torch.cuda.synchronize()
    print(f"{name}: {a.elapsed_time(b) * 1e3 / 80:.1f} us per encoder forward at N={N}")
