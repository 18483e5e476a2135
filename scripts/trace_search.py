"""Per-role timeline of the fused search kernel's fp16 pass (developer tool).
Needs the trace build: `make -C robot-manipulation-vq-vae_b200/csrc trace` (liblipvq_trace.so, -DLIPVQ_TRACE): lane 0 of
every warp of CTA 0 logs clock64 at the role's hand-over points.  Prints, per role, where the cycles go.
usage: python scripts/trace_search.py N K D [t1|uniform]"""
import collections
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["LIPVQ_LIB"] = os.path.join(ROOT, "robot-manipulation-vq-vae_b200", "lipvq", "liblipvq_trace.so")
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import torch  # noqa: E402

import lipvq  # noqa: E402
from lipvq import _cabi, ops  # noqa: E402
from bench_search import case  # noqa: E402

NAMES = {0: "start", 1: "setup_done", 2: "role_done", 3: "exit", 4: "mu_done", 5: "codebook_done", 6: "dep_wait_done", 40: "mu_loads_issued", 41: "slice_loads_issued",
         42: "mu_reduced", 43: "mu_synced", 44: "slice_synced", 45: "all_arrived", 46: "bulk_issued", 47: "image_landed",
         10: "epi:wait_acc", 11: "epi:acc_full", 12: "epi:last_ld", 13: "epi:en_stored", 14: "epi:folded",
         24: "cvt:converted", 25: "cvt:stats", 26: "cvt:fenced", 20: "cvt:wait_a_empty", 21: "cvt:a_empty", 22: "cvt:z_full", 23: "cvt:a_filled",
         30: "mma:a_full", 31: "mma:acc_empty", 32: "mma:issued"}


def report(fn_name, nwarps, cap, roles):
    fn = getattr(_cabi.lib, fn_name)
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p]
    buf = (C.c_ulonglong * (nwarps * cap))()
    cnt = (C.c_int * nwarps)()
    assert fn(buf, cnt) == cap
    t0 = min(buf[w * cap] >> 8 for w in range(nwarps) if cnt[w] > 0)
    for w, role in roles.items():
        ev = [(buf[w * cap + i] & 0xFF, (buf[w * cap + i] >> 8) - t0) for i in range(cnt[w])]
        print(f"--- warp {w} ({role}): {len(ev)} events, last at {ev[-1][1]} cycles")
        # time attributed to the interval between consecutive hand-over points
        acc = collections.defaultdict(lambda: [0, 0])
        for (e0, c0), (e1, c1) in zip(ev, ev[1:]):
            k = f"{NAMES.get(e0, e0)} -> {NAMES.get(e1, e1)}"
            acc[k][0] += c1 - c0
            acc[k][1] += 1
        for k, (tot, n) in sorted(acc.items(), key=lambda kv: -kv[1][0]):
            print(f"   {k:42s} total {tot:9d} cyc  n={n:5d}  mean {tot / n:9.1f}")
        if "-v" in sys.argv:
            print("   first events:", [(NAMES.get(e, e), c) for e, c in ev[:40]])


def main():
    N, K, D = (int(v) for v in sys.argv[1:4])
    mode = sys.argv[4] if len(sys.argv) > 4 else "t1"
    z, cb = case(N, K, D, mode)
    nt = next((int(a.split("=")[1]) for a in sys.argv if a.startswith("--ties=")), 0)
    if nt:  # nt rows sit (almost) midway between two codes: the fp16 pass cannot certify them
        g = torch.Generator().manual_seed(1)
        rows = torch.randperm(N, generator=g)[:nt]
        a_, b_ = torch.randint(0, K, (nt,), generator=g), torch.randint(0, K, (nt,), generator=g)
        mid = 0.5 * (cb.cpu()[a_] + cb.cpu()[b_]) + 1e-7 * torch.randn(nt, D, generator=g)
        z[rows.cuda()] = mid.cuda()
    algo = lipvq.ALGO_TCGEN05
    ws = torch.empty(ops.search_workspace_bytes(N, D, K, lipvq.METRIC_L2, algo), dtype=torch.uint8, device="cuda")
    idx = torch.empty(N, dtype=torch.int64, device="cuda")
    prepared = "--prepared" in sys.argv  # lipvq_nn_search_prepare_f32 once, then the one-launch search
    if prepared:
        ops.nn_search_prepare(cb, ws, lipvq.METRIC_L2, center=z)
    run = (lambda: ops.nn_search_prepared(z, cb, ws, lipvq.METRIC_L2, out=idx)) if prepared else \
          (lambda: ops.nn_search(z, cb, lipvq.METRIC_L2, algo, out=idx, workspace=ws))
    for _ in range(3):
        run()
    if "--cold" in sys.argv:  # L2 (data AND the kernel's code) evicted, as between the kernels of a training step
        torch.empty(256 << 20, dtype=torch.uint8, device="cuda").zero_()
        run()
    torch.cuda.synchronize()
    if os.environ.get("LIPVQ_SEARCH_P1", "1") != "0" and hasattr(_cabi.lib, "lipvq_debug_trace_p1"):
        report("lipvq_debug_trace_p1", 16, 4096, {0: "epi g0", 4: "epi g1", 8: "cvt", 12: "mma"})
    else:
        report("lipvq_debug_trace", 16, 4096, {0: "epi g0", 4: "epi g1", 8: "cvt", 12: "mma", 13: "tma"})


if __name__ == "__main__":
    main()
