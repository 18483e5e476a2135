"""Developer check of the single-launch first pass (csrc/search_p1.cu) on a GPU: exactness against a direct fp32/fp64
evaluation, uncertified-row counts and time per variant (fold 0/1, slab descriptor order, p1 off = round-1 path).
usage: python scripts/p1_check.py [--quick]"""
import math
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
import lipvq  # noqa: E402
from lipvq import ops  # noqa: E402


def ref_argmin(z, cb):
    out = []
    for s in range(0, z.shape[0], 2048):
        d = (z[s:s + 2048, None, :] - cb[None]).norm(dim=-1)
        out.append(d.argmin(dim=1))
    return torch.cat(out)


def classify(z, cb, got, ref):
    bad = (got != ref).nonzero().flatten()
    if bad.numel() == 0:
        return 0, 0.0
    zz = z[bad].double()
    vg = (zz - cb[got[bad]].double()).norm(dim=-1)
    vr = (zz - cb[ref[bad]].double()).norm(dim=-1)
    return int(bad.numel()), float(((vg - vr).abs() / vr.clamp_min(1e-30)).max())


def case(N, K, D, mode, seed=0):
    g = torch.Generator().manual_seed(seed)
    cb = (torch.rand(K, D, generator=g) * 2 - 1) * math.sqrt(6.0 / D)
    z = torch.rand(N, D, generator=g) if mode == "uniform" else torch.sigmoid(0.05 * torch.randn(N, D, generator=g))
    return z.cuda(), cb.cuda()


def timed(z, cb, ws, idx, flush, iters=12):
    durs = []
    for i in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ops.nn_search(z, cb, lipvq.METRIC_L2, lipvq.ALGO_TCGEN05, out=idx, workspace=ws)
        b.record()
        torch.cuda.synchronize()
        if i >= 3:
            durs.append(a.elapsed_time(b) * 1e3)
    return statistics.mean(durs), min(durs)


def main():
    quick = "--quick" in sys.argv
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    shapes = [(65536, 512, 64, "t1"), (65536, 512, 64, "uniform"), (4096, 512, 64, "uniform"), (1000, 77, 12, "uniform"),
              (300, 130, 256, "uniform"), (65536, 256, 128, "t1"), (3000, 300, 36, "uniform"), (3000, 300, 30, "uniform")]
    if quick:
        shapes = shapes[:3]
    variants = [("p1 fold1", {"LIPVQ_SEARCH_P1": "1", "LIPVQ_P1_FOLD": "1", "LIPVQ_SLAB_DESC_SWAP": "0"}),
                ("p1 fold0", {"LIPVQ_SEARCH_P1": "1", "LIPVQ_P1_FOLD": "0", "LIPVQ_SLAB_DESC_SWAP": "0"}),
                ("p1 fold1 swap", {"LIPVQ_SEARCH_P1": "1", "LIPVQ_P1_FOLD": "1", "LIPVQ_SLAB_DESC_SWAP": "1"}),
                ("round-1 path", {"LIPVQ_SEARCH_P1": "0", "LIPVQ_P1_FOLD": "1", "LIPVQ_SLAB_DESC_SWAP": "0"})]
    for N, K, D, mode in shapes:
        z, cb = case(N, K, D, mode)
        ref = ref_argmin(z, cb)
        ws = torch.empty(ops.search_workspace_bytes(N, D, K, lipvq.METRIC_L2, lipvq.ALGO_TCGEN05), dtype=torch.uint8, device="cuda")
        idx = torch.empty(N, dtype=torch.int64, device="cuda")
        for name, env in variants:
            os.environ.update(env)
            try:
                idx.fill_(-1)
                got, stats = ops.nn_search(z, cb, lipvq.METRIC_L2, lipvq.ALGO_TCGEN05, want_stats=True, out=idx, workspace=ws)
                torch.cuda.synchronize()
                nbad, gap = classify(z, cb, got, ref)
                mean_us, min_us = timed(z, cb, ws, idx, flush)
                print(f"N={N} K={K} D={D} {mode:8s} {name:14s} mismatches={nbad} worst_rel_gap={gap:.2e} "
                      f"uncertified_p1={int(stats[1])} exact_scan={int(stats[0])} us={mean_us:.1f} min_us={min_us:.1f}", flush=True)
            except Exception as e:  # noqa: BLE001
                print(f"N={N} K={K} D={D} {mode} {name}: FAILED {e}", flush=True)
                return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
