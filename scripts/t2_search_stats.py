"""The nearest-code searches of one HierarchicalVQVAE forward at the shipped config: shapes, metric, rows the tensor pass
could not certify (stats), time per call."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
import lipvq
from lipvq import ops
CFG = dict(embed_dim=512, num_subclusters=1024, num_clusters=512, num_stages=2, num_layers_per_stage=4, dropout=0.0, kmeans_init=False)
B, T = (int(a) for a in sys.argv[1:3]) if len(sys.argv) > 2 else (4096, 16)
torch.manual_seed(0)
m = lipvq.HierarchicalVQVAE(12, **CFG).cuda().eval()
x = torch.rand(B, T, 12, device="cuda") * 2 - 1
orig = ops.nn_search
def spy(z, cb, metric=lipvq.METRIC_L2, algo=lipvq.ALGO_AUTO, **kw):
    for rep in range(2):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        idx, st = orig(z, cb, metric, algo, want_stats=True)
        b.record(); torch.cuda.synchronize()
    zc = z - z.mean(0, keepdim=True)
    print(f"search N={z.shape[0]} D={z.shape[1]} K={cb.shape[0]} metric={metric} algo={ops.pick_algo(z.shape[0], z.shape[1], cb.shape[0], metric)}: "
          f"{a.elapsed_time(b):.3f} ms stats={st.tolist()} |z|={z.norm(dim=1).mean().item():.3f} |z-mean|={zc.norm(dim=1).mean().item():.4f} "
          f"|e|={cb.norm(dim=1).mean().item():.3f}", flush=True)
    if "--save" in sys.argv and z.shape[0] > 20000:
        torch.save({"z": z[:4096].cpu(), "cb": cb.cpu(), "metric": metric, "idx": idx[:4096].cpu()}, os.path.join(ROOT, "gpurun_out", f"t2_search_{cb.shape[0]}.pt"))
    return orig(z, cb, metric, algo, **kw)
nsteps = int(sys.argv[3]) if len(sys.argv) > 3 else 0
if nsteps:
    m.train()
    opt = torch.optim.AdamW(m.parameters(), lr=1e-3, weight_decay=1e-4)
    for i in range(nsteps):
        xi = torch.rand(B, T, 12, device="cuda") * 2 - 1
        opt.zero_grad(set_to_none=True)
        out = m(xi, training=True)
        m.compute_vqvae_loss(out, xi, lambda_rec=0.002)["vqvae_loss"].backward()
        opt.step()
    m.eval()
ops.nn_search = spy
import lipvq.hvq as H
H.ops.nn_search = spy
with torch.no_grad():
    m(x, training=False)
