"""Per-role timeline of mlp3_tc_kernel (developer tool; needs `make -C robot-manipulation-vq-vae_b200/csrc trace`)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["LIPVQ_LIB"] = os.path.join(ROOT, "robot-manipulation-vq-vae_b200", "lipvq", "liblipvq_trace.so")
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
import torch
import lipvq as L
from lipvq import _cabi, ops
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
g = torch.Generator().manual_seed(0)
dims = (12, 64, 128, 64)
x = (torch.rand(N, dims[0], generator=g) * 2 - 1).cuda()
Ws = [(torch.randn(o, i, generator=g) / i ** 0.5).cuda() for i, o in zip(dims[:-1], dims[1:])]
bs = [(torch.randn(o, generator=g) * 0.1).cuda() for o in dims[1:]]
for _ in range(3):
    y = ops.mlp3_fwd(x, Ws[0], bs[0], L.ACT_GELU, Ws[1], bs[1], L.ACT_GELU, Ws[2], bs[2], L.ACT_SIGMOID)
torch.cuda.synchronize()
NAMES = {0: "start", 1: "setup", 2: "role_done", 3: "exit", 4: "weights_staged", 10: "w:step_begin", 11: "w:arrived", 12: "w:done_seen",
         13: "w:output_stored", 30: "m:wait_full", 31: "m:full_seen", 32: "m:issued"}
fn = _cabi.lib.lipvq_debug_trace_mlp3
fn.restype = C.c_int; fn.argtypes = [C.c_void_p, C.c_void_p]
cap = 512
buf = (C.c_ulonglong * (32 * cap))(); cnt = (C.c_int * 32)()
assert fn(buf, cnt) == cap
t0 = min(buf[w * cap] >> 8 for w in range(32) if cnt[w] > 0)
for w, role in ((0, "worker P0"), (12, "worker P1"), (24, "mma")):
    ev = [(buf[w * cap + i] & 0xFF, (buf[w * cap + i] >> 8) - t0) for i in range(cnt[w])]
    print(f"--- warp {w} ({role}): {len(ev)} events")
    print("   ", " ".join(f"{NAMES.get(e, e)}@{c}" for e, c in ev[:80]))
