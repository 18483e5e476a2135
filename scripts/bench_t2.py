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


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "measured (MEASURED_PEAKS.json, sustained dense bf16)"
    except Exception:
        return 1400.0, "fallback (B200_PROFILING.md)"


def conv_roofline(B, T):
    """the dominant kernel of the step: one MS-TCN layer forward (conv_tc_kernel<fwd>, or the FFMA tiles below its gate) at this
    shape's token count, timed live with CUDA events, against the dense bf16 tensor peak.  Algorithmic flops = 2 N E 3E (the
    fp32 contraction the reference performs); the kernel issues six bf16 MMA chains for it, so the MMA-level rate is 6 x."""
    from lipvq import ops
    E, N = CFG["embed_dim"], B * T
    g = torch.Generator().manual_seed(0)
    x = torch.randn(N, E, generator=g).cuda()
    W = (torch.randn(E, E, 3, generator=g) / (3 * E) ** 0.5).cuda()
    b = torch.zeros(E).cuda()
    Wf, _ = ops.conv_pack_weights(W, True, False)
    xs = [x.clone() for _ in range(4)]           # 4 x 134 MB at the large shape: every launch reads its rows from HBM
    for i in range(3):
        ops.dilated_conv_fwd(xs[i % 4], T, Wf, b, 2)
    torch.cuda.synchronize()
    a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(8):
        ops.dilated_conv_fwd(xs[i % 4], T, Wf, b, 2)
    c.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(c) / 8 * 1e3
    flops = 2.0 * N * E * 3 * E
    peak, src = peaks()
    return {"bound": "tensor", "kernel": "lipvq_dilated_conv_fwd_f32 (conv_tc_kernel<fwd>: E = 512 layer forward; 48 such launches per "
            "step: 16 layers x forward / input gradient / weight gradient)", "achieved": flops / us / 1e6, "peak": peak,
            "unit": "TFLOP/s", "frac": flops / us / 1e6 / peak, "frac_mma_level": 6 * flops / us / 1e6 / peak, "traffic": None,
            "us_per_launch": us, "algorithmic_flops": flops, "peak_source": src,
            "note": "ncu: profiles/r2f_conv_tc_full.md (tensor pipe active 48.5 %, shared-memory data pipe 87 %)"}


def e2e_step(B, T, steps):
    """the step through the public API with HOST inputs: pinned [B, T, A] batch -> device copy -> HvqTrainer.step -> loss on host"""
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    m = lipvq.HierarchicalVQVAE(A_DIM, **CFG).to(dev).train()
    tr = lipvq.HvqTrainer(m, lr=1e-3, weight_decay=1e-4, lambda_rec=0.002)
    xh = [(torch.rand(B, T, A_DIM) * 2 - 1).pin_memory() for _ in range(4)]
    xd = torch.empty(B, T, A_DIM, device=dev)
    for i in range(3):
        xd.copy_(xh[i % 4], non_blocking=True)
        float(tr.step(xd)["vqvae_loss"])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(steps):
        xd.copy_(xh[i % 4], non_blocking=True)
        loss = float(tr.step(xd)["vqvae_loss"])     # device -> host read of the step's loss
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / steps * 1e3
    return {"value": B / (ms * 1e-3), "unit": "chunks/s", "ms_per_step": ms, "h2d_bytes_per_step": B * T * A_DIM * 4,
            "d2h_bytes_per_step": 4, "final_loss": loss}


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
        if "--no-graph" not in sys.argv:
            line["e2e"] = e2e_step(B, T, steps)
        line["roofline"] = conv_roofline(B, T)
        if "--no-cpu" not in sys.argv:
            Bc = min(B, 256)
            cpu_ms = cpu_port(Bc, T)
            line["cpu_port_forward_eval"] = {"ms": cpu_ms, "chunks_per_s": Bc / (cpu_ms * 1e-3), "sample": f"B={Bc}",
                                             "cores": torch.get_num_threads(), "kind": "port"}
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
