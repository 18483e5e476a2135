"""Per-role timeline of the tcgen05 linear kernel (developer tool; needs `make -C ...csrc trace`).
usage: python scripts/trace_linear.py N IN OUT ACT_IN ACT_OUT     (forward;  acts: none|relu|gelu|sigmoid)"""
import sys

import trace_search as ts  # sets LIPVQ_LIB to the trace build before importing lipvq
import torch

import lipvq
from lipvq import ops

ACTS = {"none": lipvq.ACT_NONE, "relu": lipvq.ACT_RELU, "gelu": lipvq.ACT_GELU, "sigmoid": lipvq.ACT_SIGMOID}


def main():
    N, IN, OUT = (int(v) for v in sys.argv[1:4])
    a_in, a_out = ACTS[sys.argv[4]], ACTS[sys.argv[5]]
    g = torch.Generator().manual_seed(0)
    x = torch.randn(N, IN, generator=g).cuda()
    W = (torch.randn(OUT, IN, generator=g) / IN ** 0.5).cuda()
    b = torch.randn(OUT, generator=g).cuda()
    y = torch.empty(N, OUT, device="cuda")
    if "bwdw" in sys.argv:  # N IN OUT ACT_X none bwdw: dW[OUT, IN] += dY^T act(X)
        dY = torch.randn(N, OUT, generator=g).cuda()
        dW, db = torch.zeros(OUT, IN, device="cuda"), torch.zeros(OUT, device="cuda")
        for _ in range(3):
            ops.linear_bwd_w(dY, x, a_in, dW, db)
    elif "bwdx" in sys.argv:  # N IN OUT none none bwdx: dX[N, IN] = (dY[N, OUT] @ W) * given derivative
        dY = torch.randn(N, OUT, generator=g).cuda()
        dX = torch.empty(N, IN, device="cuda")
        for _ in range(3):
            ops.linear_bwd_x(dY, W, x, lipvq.ACT_GIVEN, out=dX)
    elif "dual" in sys.argv:  # training-time forward: value + derivative
        y2 = torch.empty(N, OUT, device="cuda")
        for _ in range(3):
            ops.linear_fwd_dual(x, a_in, W, b, a_out, y, y2)
    else:
        for _ in range(3):
            ops.linear_fwd(x, a_in, W, b, a_out, out=y)
    torch.cuda.synchronize()
    ts.NAMES.update({10: "epi:wait_acc", 11: "epi:acc_full", 12: "epi:group_math", 13: "epi:group_stored", 14: "epi:stored", 20: "cvt:wait_a_empty", 21: "cvt:a_empty",
                     23: "cvt:a_filled", 30: "mma:a_full", 31: "mma:acc_empty", 32: "mma:issued"})
    ts.report("lipvq_debug_trace_linear", 32, 1024, {0: "epi p0", 12: "epi p3", 16: "cvt", 23: "cvt last", 24: "mma"})


if __name__ == "__main__":
    main()
