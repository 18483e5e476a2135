"""One MS-TCN layer (forward, input gradient, weight gradient) at the shipped width (E = 512): tensor-core contractions
(conv_tc.cu) vs the FFMA tiles."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
import lipvq as L
from lipvq import ops
B, T, E = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (4096, 16, 512)
g = torch.Generator().manual_seed(0)
x = torch.randn(B * T, E, generator=g).cuda()
W = (torch.randn(E, E, 3, generator=g) / (3 * E) ** 0.5).cuda()
b = (torch.randn(E, generator=g) * 0.1).cuda()
Wf, Wb = ops.conv_pack_weights(W, True, True)
go = torch.randn(B * T, E, generator=g).cuda()
_, mask = ops.dilated_conv_fwd(x, T, Wf, b, 2)
dWf, db = torch.zeros(E, 3, E).cuda(), torch.zeros(E).cuda()
for algo in ("tc", "tc-bf16", "simt"):
    os.environ.pop("LIPVQ_CONV_PRECISION", None)
    if algo == "tc-bf16":
        os.environ["LIPVQ_CONV_PRECISION"] = "bf16"
    if algo == "simt":
        os.environ["LIPVQ_CONV_ALGO"] = "simt"
    for name, fn in (("forward", lambda: ops.dilated_conv_fwd(x, T, Wf, b, 2)),
                     ("bwd_x", lambda: ops.dilated_conv_bwd_x(go, mask, T, Wb, 2)),
                     ("bwd_w", lambda: ops.dilated_conv_bwd_w(go, mask, x, T, 2, dWf, db))):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            fn()
        c.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(c) / 5
        print(f"{algo}: {ms:.3f} ms per layer {name} at N={B*T}, E={E} ({2.0 * B * T * E * 3 * E / ms / 1e9:.1f} TFLOP/s)")
