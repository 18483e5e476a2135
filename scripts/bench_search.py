"""Nearest-code search sweep (BASELINE config 3): K in {512,4096,65536} x D in {64,256}, N = 65536 tokens, fp32 in,
int64 indices out.  Times the whole lipvq_nn_search_f32 call (prep + fp16 pass + 3-term pass + exact scan) with CUDA
events, L2 flushed between iterations, and reports algorithmic 2*N*K*D flop / time against the measured bf16 peak.
usage: python scripts/bench_search.py [--quick] [--algo tc|simt]"""
import json
import math
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
import lipvq  # noqa: E402
from lipvq import ops  # noqa: E402


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), float(p["bf16_tflops"])
    except Exception:
        return 6541.5, 1657.6


def case(N, K, D, mode, seed=0):
    g = torch.Generator().manual_seed(seed)
    if mode == "uniform":  # SURVEY 8(d) raw-latent mode
        z = torch.rand(N, D, generator=g)
        cb = (torch.rand(K, D, generator=g) * 2 - 1) * math.sqrt(6.0 / D)
    else:  # "t1": sigmoid latents of a fresh LipschitzMLP: tight cluster around 0.5
        z = torch.sigmoid(0.05 * torch.randn(N, D, generator=g))
        cb = (torch.rand(K, D, generator=g) * 2 - 1) * math.sqrt(6.0 / D)
    return z.cuda(), cb.cuda()


def main():
    quick = "--quick" in sys.argv
    algo = lipvq.ALGO_SIMT_F32 if "simt" in sys.argv else lipvq.ALGO_TCGEN05
    hbm, tf = peaks()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    shapes = [(65536, 512, 64), (65536, 4096, 64), (65536, 512, 256), (65536, 4096, 256)]
    if not quick:
        shapes += [(65536, 65536, 64), (65536, 65536, 256), (1 << 20, 512, 64)]
    for N, K, D in shapes:
        for mode in ("uniform", "t1"):
            z, cb = case(N, K, D, mode)
            ws = torch.empty(ops.search_workspace_bytes(N, D, K, lipvq.METRIC_L2, algo), dtype=torch.uint8, device="cuda")
            idx = torch.empty(N, dtype=torch.int64, device="cuda")
            durs = []
            iters = 4 if K * D >= 65536 * 64 else 14
            if "--profile" in sys.argv:
                iters = 3
            for i in range(iters):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                ops.nn_search(z, cb, lipvq.METRIC_L2, algo, out=idx, workspace=ws)
                b.record()
                torch.cuda.synchronize()
                if i >= 2:
                    durs.append(a.elapsed_time(b))
            _, st = ops.nn_search(z, cb, lipvq.METRIC_L2, algo, want_stats=True, workspace=ws)
            ms = statistics.median(durs)
            flops = 2.0 * N * K * D
            byt = N * (4 * D + 8) + K * D * 4
            print(json.dumps({"N": N, "K": K, "D": D, "data": mode, "us": round(ms * 1e3, 2), "min_us": round(min(durs) * 1e3, 2),
                              "tflops": round(flops / ms / 1e9, 1), "tensor_frac": round(flops / ms / 1e9 / tf, 4),
                              "hbm_frac": round(byt / ms / 1e6 / hbm, 4), "rows_exact_scan": int(st[0]),
                              "rows_level1_uncertified": int(st[1])}), flush=True)
            del z, cb, ws


if __name__ == "__main__":
    main()
