"""Timing of the fork's live tokenizer, HierarchicalVQVAE (T2: MS-TCN encoder / cosine two-level VQ with EMA / MS-TCN
decoder; reference robomimic/algo/heirarchical_vqvae.py:8-706), at the shipped config (config/icl_config.py:130-135:
embed_dim 512, 2 stages x 4 layers, 1024 / 512 codes) -- forward + compute_vqvae_loss + backward + AdamW, CUDA events,
synthetic action chunks -- next to the oracle port of the same step on the host cores (bounded sample).
usage: python scripts/bench_t2.py [--steps K] [--no-cpu] [--big-only]      (one JSON line per shape)"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import lipvq  # noqa: E402

CFG = dict(embed_dim=512, num_subclusters=1024, num_clusters=512, num_stages=2, num_layers_per_stage=4, dropout=0.0,
           kmeans_init=False)
A_DIM = 12


def ours(B, T, steps):
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    m = lipvq.HierarchicalVQVAE(A_DIM, **CFG).to(dev).train()
    opt = torch.optim.AdamW(m.parameters(), lr=1e-3, weight_decay=1e-4)
    xs = [torch.rand(B, T, A_DIM, device=dev) * 2 - 1 for _ in range(4)]

    def step(x):
        opt.zero_grad(set_to_none=True)
        out = m(x, training=True)
        loss = m.compute_vqvae_loss(out, x, lambda_rec=0.002)["vqvae_loss"]
        loss.backward()
        opt.step()
        return loss

    for i in range(3):
        step(xs[i % 4])
    torch.cuda.synchronize()
    c0 = lipvq.launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        loss = step(xs[i % 4])
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    launches = (lipvq.launch_count() - c0) / steps
    # forward only (tokenisation: encode + both quantisers)
    with torch.no_grad():
        m.eval()
        for i in range(2):
            m(xs[i % 4], training=False)
        torch.cuda.synchronize()
        a.record()
        for i in range(steps):
            m(xs[i % 4], training=False)
        b.record()
        torch.cuda.synchronize()
    ms_fwd = a.elapsed_time(b) / steps
    return ms, ms_fwd, launches, float(loss)


def ours_graph(B, T, steps):
    """the same step through lipvq.HvqTrainer (one CUDA graph per step; dead codes replaced with static shapes)"""
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    m = lipvq.HierarchicalVQVAE(A_DIM, **CFG).to(dev).train()
    tr = lipvq.HvqTrainer(m, lr=1e-3, weight_decay=1e-4, lambda_rec=0.002)
    xs = [torch.rand(B, T, A_DIM, device=dev) * 2 - 1 for _ in range(4)]
    for i in range(3):
        tr.step(xs[i % 4])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        losses = tr.step(xs[i % 4])
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps, float(losses["vqvae_loss"])


def cpu_port(B, T, max_seconds=15.0):
    import lipvq_oracle as O
    torch.manual_seed(0)
    p = O.t2_init(A_DIM, CFG["embed_dim"], CFG["num_subclusters"], CFG["num_clusters"], CFG["num_stages"], CFG["num_layers_per_stage"])
    x = torch.rand(B, T, A_DIM) * 2 - 1
    n, t0 = 0, time.perf_counter()
    while True:
        with torch.no_grad():
            O.t2_forward(p, x, training=False, num_stages=CFG["num_stages"], num_layers=CFG["num_layers_per_stage"])
        n += 1
        if time.perf_counter() - t0 > max_seconds or n >= 20:
            break
    return (time.perf_counter() - t0) / n * 1e3


def main():
    steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 20
    for B, T in (((4096, 16),) if "--big-only" in sys.argv else ((16, 10), (4096, 16))):
        ms, ms_fwd, launches, loss = ours(B, T, steps)
        line = {"workload": f"HierarchicalVQVAE training step (fwd + loss + bwd + AdamW), B={B}, T={T}, A={A_DIM}, E=512, "
                            "2x4 MS-TCN layers per side, 1024/512 codes, fp32, eager autograd over liblipvq kernels",
                "ms_per_step": ms, "chunks_per_s": B / (ms * 1e-3), "ms_forward_eval": ms_fwd,
                "chunks_per_s_forward_eval": B / (ms_fwd * 1e-3), "own_kernel_launches_per_step": launches, "final_loss": loss}
        if "--no-graph" not in sys.argv:
            ms_g, loss_g = ours_graph(B, T, steps)
            line["graph_step"] = {"ms_per_step": ms_g, "chunks_per_s": B / (ms_g * 1e-3), "final_loss": loss_g,
                                  "how": "lipvq.HvqTrainer: the whole step as one CUDA graph"}
        if "--no-cpu" not in sys.argv:
            Bc = min(B, 256)
            cpu_ms = cpu_port(Bc, T)
            line["cpu_port_forward_eval"] = {"ms": cpu_ms, "chunks_per_s": Bc / (cpu_ms * 1e-3), "sample": f"B={Bc}",
                                             "cores": torch.get_num_threads(), "kind": "port"}
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
