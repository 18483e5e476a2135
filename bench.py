#!/usr/bin/env python
"""bench.py -- LipVQ-VAE tokenizer hot path on B200 (see DESIGN.md "Measurement").

Workload (BASELINE.json configs[1]): tokenizer TRAINING step with EMA codebook update on synthetic
RoboCasa-shaped action chunks -- 4096 chunks x 16 steps = 65536 tokens per GPU, 12-DoF actions, codebook 512x64,
fp32.  One "step" = encode -> nearest-code search -> gather/loss -> EMA statistics -> decode -> backward -> (one
all-reduce when N>1) -> EMA update -> AdamW.  `value` = chunks/s over all GPUs with inputs resident in HBM; `e2e`
= the same through TokenizerTrainer.step() with pinned-host inputs (H2D inside) and a D2H read of the loss.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
N>1 is launched by torchrun (one rank per GPU, NCCL).  --impl reference times the CPU oracle port of the reference's
PyTorch path (oracle/lipvq_oracle.py; /root/reference itself does not exist on the GPU box) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))

import torch  # noqa: E402

CHUNKS, T, F_, D, K, H = 4096, 16, 12, 64, 512, 128
N_TOK = CHUNKS * T
CPU_SAMPLE_CHUNKS = 512  # bounded sample of the same workload for the CPU legs (8192 tokens per step)
WORKLOAD = (f"LipVQ-VAE tokenizer training step + EMA codebook update, {CHUNKS} chunks x T={T} per GPU, "
            f"F={F_}, K={K}, D={D}, AdamW(1e-3,1e-4)")


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), float(p["bf16_tflops"]), float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "measured"
    except Exception:
        return 6650.0, 1590.0, 1400.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        # under load = upper half of the samples (idle samples before/after drag the median down)
        sm.sort()
        load = sm[len(sm) // 2:] if sm else []
        return {"sm_mhz": statistics.median(load) if load else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def _synthetic_actions(n_tokens: int, seed: int) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return torch.rand(n_tokens, F_, generator=g) * 2 - 1  # tanh-range actions


# ----------------------------------------------------------------------------------------------
# CPU legs: the oracle port of the reference's PyTorch path
# ----------------------------------------------------------------------------------------------
def _cpu_step_fn(threads: int):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import lipvq_oracle as O
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    p = O.t1_init(F_, D, K)
    ema = {"ema_cluster_size": torch.zeros(K), "ema_codebook": torch.randn(K, D)}
    x = _synthetic_actions(CPU_SAMPLE_CHUNKS * T, 123)
    st = {}

    def step():
        return O.t1_train_step(p, x, st, ema=ema)["loss"].item()
    return step


def cpu_baseline(max_seconds: float = 25.0, threads: int = 0, sample_chunks: int = 0):
    global CPU_SAMPLE_CHUNKS
    threads = threads or (os.cpu_count() or 1)
    if sample_chunks:
        CPU_SAMPLE_CHUNKS = sample_chunks
    step = _cpu_step_fn(threads)
    step()
    t0 = time.perf_counter()
    n = 0
    while n < 2 or (time.perf_counter() - t0 < max_seconds / 3 and n < 8):
        step()
        n += 1
    dt = (time.perf_counter() - t0) / n
    out = {"value": CPU_SAMPLE_CHUNKS / dt, "unit": "chunks/s", "cores": threads, "kind": "port",
           "sample": f"{n} training steps of {CPU_SAMPLE_CHUNKS} chunks ({CPU_SAMPLE_CHUNKS * T} tokens), same F/K/D, "
                     f"oracle/lipvq_oracle.py (torch CPU restatement of the reference modules), {dt * 1e3:.0f} ms/step"}
    torch.set_num_threads(os.cpu_count() or 1)
    return out


def cpu_cdist_argmin_us(N, K, D, rows=512, threads=0):
    """The comparator BASELINE config 3 names: torch.cdist + argmin (heirarchical_vqvae.py:384,402,432) on the host cores,
    on `rows` rows (row-chunked), scaled to N rows."""
    torch.set_num_threads(threads or (os.cpu_count() or 1))
    g = torch.Generator().manual_seed(0)
    z, cb = torch.rand(rows, D, generator=g), (torch.rand(K, D, generator=g) * 2 - 1) * (6.0 / D) ** 0.5
    torch.cdist(z[:64], cb).argmin(dim=1)
    t0 = time.perf_counter()
    for s0 in range(0, rows, 128):
        torch.cdist(z[s0:s0 + 128], cb).argmin(dim=1)
    return (time.perf_counter() - t0) * 1e6 * (N / rows)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    step = _cpu_step_fn(threads)
    for _ in range(max(1, min(args.warmup, 2))):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    val = CPU_SAMPLE_CHUNKS / dt
    line = {"impl": "reference", "metric": "action_chunks_per_sec", "value": val, "unit": "chunks/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample_chunks_per_step": CPU_SAMPLE_CHUNKS},
            "cpu_baseline": {"value": val, "unit": "chunks/s", "cores": threads, "kind": "port",
                             "sample": f"each step = {CPU_SAMPLE_CHUNKS} chunks of the workload on {threads} host threads"},
            "e2e": {"value": val, "unit": "chunks/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
def _sync_all(world):
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def _time_events(fn, iters, flush=None):
    """mean / min duration in us of fn() on the current stream, L2 flushed (written) between iterations"""
    durs = []
    for i in range(iters + 3):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        if i >= 3:
            durs.append(a.elapsed_time(b) * 1e3)
    return statistics.mean(durs), min(durs)


def _time_rotating(fn, bufs, iters, streamed=False):
    """mean / min duration in us of fn(buf) over calls that walk through `bufs` (together larger than the L2, so every
    call finds its input in HBM; nothing is flushed, small per-launch state such as a prepared codebook stays in L2 as it
    does in a real run).
    streamed=False: an event pair and a synchronize around every single call; a short spin kernel is queued first so the
    host has issued the launch before the GPU reaches the first event (otherwise the event pair times the Python call).
    streamed=True: the len(bufs) calls are captured once as a CUDA graph (no host time between launches) and replayed
    back to back between ONE event pair: mean = total / launches."""
    n = len(bufs)
    for i in range(n):
        fn(bufs[i])
    torch.cuda.synchronize()
    if streamed:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for i in range(n):
                fn(bufs[i])
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(n):
                fn(bufs[i])
        reps = max(1, iters // n)
        g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            g.replay()
        b.record()
        torch.cuda.synchronize()
        t = a.elapsed_time(b) * 1e3 / (reps * n)
        del g
        return t, t
    durs = []
    for i in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(200000)
        a.record()
        fn(bufs[i % n])
        b.record()
        torch.cuda.synchronize()
        durs.append(a.elapsed_time(b) * 1e3)
    return statistics.mean(durs), min(durs)


def t2_leg(dev, B=4096, T=16, steps=5):
    """The fork's live tokenizer (HierarchicalVQVAE, SURVEY section 8 'next' row) at its shipped config: training step through
    lipvq.HvqTrainer (one CUDA graph) and eval forward, synthetic chunks resident in HBM.  Not the headline; an extra key."""
    import lipvq
    cfg = dict(embed_dim=512, num_subclusters=1024, num_clusters=512, num_stages=2, num_layers_per_stage=4, dropout=0.0,
               kmeans_init=False)
    torch.manual_seed(0)
    m = lipvq.HierarchicalVQVAE(12, **cfg).to(dev).train()
    tr = lipvq.HvqTrainer(m, lr=1e-3, weight_decay=1e-4, lambda_rec=0.002)
    xs = [torch.rand(B, T, 12, device=dev) * 2 - 1 for _ in range(4)]
    for i in range(3):
        tr.step(xs[i % 4])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        losses = tr.step(xs[i % 4])
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
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
    ms_f = a.elapsed_time(b) / steps
    out = {"workload": f"HierarchicalVQVAE (heirarchical_vqvae.py:8-706) E=512, 2x4 MS-TCN layers per side, 1024/512 codes, "
                       f"B={B}, T={T}, fp32; 4 input batches rotate (activations >> L2)",
           "train_step": {"ms_per_step": ms, "chunks_per_s": B / (ms * 1e-3), "how": "lipvq.HvqTrainer, one CUDA graph per step",
                          "final_loss": float(losses["vqvae_loss"])},
           "forward_eval": {"ms": ms_f, "chunks_per_s": B / (ms_f * 1e-3)},
           "see": "scripts/bench_t2.py (eager step, B=16 shape, oracle port on the host cores), DESIGN.md section 7"}
    del tr, m, xs
    torch.cuda.empty_cache()
    return out


def search_sweep(dev, hbm, tf_burst, flush, cpu_rows=256):
    """BASELINE config 3: K in {512, 4096, 65536} x D in {64, 256}, N = 65536, raw latents z ~ U(0,1), with the
    two baselines the config names: torch.cdist + argmin on the host cores and eager on the same B200."""
    import lipvq
    from lipvq import ops
    out = []
    N = N_TOK
    for K, D in ((512, 64), (4096, 64), (65536, 64), (512, 256), (4096, 256), (65536, 256)):
        g = torch.Generator().manual_seed(5 + K + D)
        z = torch.rand(N, D, generator=g).to(dev)
        cb = ((torch.rand(K, D, generator=g) * 2 - 1) * (6.0 / D) ** 0.5).to(dev)
        algo = lipvq.ALGO_TCGEN05
        ws = torch.empty(ops.search_workspace_bytes(N, D, K, lipvq.METRIC_L2, algo), dtype=torch.uint8, device=dev)
        idx = torch.empty(N, dtype=torch.int64, device=dev)
        iters = 3 if K * D >= 65536 * 64 else 8
        us, mn = _time_events(lambda: ops.nn_search(z, cb, lipvq.METRIC_L2, algo, out=idx, workspace=ws), iters, flush)
        _, st = ops.nn_search(z, cb, lipvq.METRIC_L2, algo, want_stats=True, workspace=ws)
        us_prep = None
        if ops.search_prepared_supported(N, D, K):  # resident codebook: the prepared pair (one launch per search)
            ops.nn_search_prepare(cb, ws, lipvq.METRIC_L2, center=z)
            us_prep, _ = _time_events(lambda: ops.nn_search_prepared(z, cb, ws, lipvq.METRIC_L2, out=idx), iters, flush)
        eager_us, _ = _time_events(lambda: torch.cdist(z, cb).argmin(dim=1), 2 if K >= 65536 else 4, flush)
        ref_idx = torch.cdist(z[:2048], cb, compute_mode="donot_use_mm_for_euclid_dist").argmin(dim=1)
        flops = 2.0 * N * K * D
        out.append({"N": N, "K": K, "D": D, "latents": "uniform", "us": round(us, 1), "min_us": round(mn, 1),
                    "tflops": round(flops / us / 1e6, 1), "frac_tensor": round(flops / us / 1e6 / tf_burst, 4),
                    "frac_hbm": round((N * (4 * D + 8) + K * D * 4) / us / 1e3 / hbm, 4),
                    "rows_uncertified_by_fp16_pass": int(st[1]), "rows_exact_scan": int(st[0]),
                    "index_mismatch_vs_eager_direct_2048": int((idx[:2048] != ref_idx).sum()),
                    "us_prepared_codebook": None if us_prep is None else round(us_prep, 1),
                    "baseline_gpu_eager_us": round(eager_us, 1),
                    "baseline_cpu_us": round(cpu_cdist_argmin_us(N, K, D, rows=cpu_rows if K < 65536 else 64), 0)})
        del z, cb, ws, idx
    return out


def run_ours(args):
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU leg)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    extra = {}
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import lipvq
    from lipvq import ops
    from lipvq.trainer import TokenizerTrainer

    torch.manual_seed(0)  # identical replicas on every rank
    model = lipvq.LLFQVAE_V4(F_, D, K)
    model.quantizer = lipvq.LFQQuantizerEMA(K, D)
    model.to(dev).train()
    tr = TokenizerTrainer(model, N_TOK, use_graph=True)
    n_in = 4
    xs_host = [_synthetic_actions(N_TOK, 1000 + 17 * rank + i).pin_memory() for i in range(n_in)]
    xs_dev = [x.to(dev) for x in xs_host]

    # launches per step, counted on one eager (un-captured) step: graph replays do not pass through the C ABI
    snap = tr._snapshot()
    tr.use_graph = False
    c0 = lipvq.launch_count()
    tr.step(xs_dev[0])
    torch.cuda.synchronize()
    launches_per_step = lipvq.launch_count() - c0
    tr._restore(snap)
    tr.use_graph = True

    # data-parallel self-check (the driver's test box has one GPU): the peer all-reduce against NCCL on the same slab,
    # and bit-identical replicas after a few steps
    collective_check = None
    if world > 1:
        for i in range(2):
            tr.step(xs_dev[i % n_in])
        torch.cuda.synchronize()
        ok_sum = True
        if tr.peer is not None:  # our kernel's sum vs NCCL's on the same slab (with NCCL itself there is nothing to compare)
            ref = tr.red.detach().clone()
            dist.all_reduce(ref, op=dist.ReduceOp.SUM)
            ok_sum = bool(torch.allclose(tr.red_out.detach(), ref, rtol=1e-5, atol=1e-6))
        flat = tr.flat_p.detach()
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        ok_rep = all(torch.equal(gathered[0], t) for t in gathered)
        flag = torch.tensor([int(ok_sum and ok_rep)], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        collective_check = "ok" if int(flag) == 1 else f"FAILED (sum {ok_sum}, replicas {ok_rep})"

    for i in range(max(args.warmup, 3)):
        tr.step(xs_dev[i % n_in])
    _sync_all(world)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _sync_all(world)
    e0.record()
    for i in range(args.steps):
        tr.step(xs_dev[i % n_in])
    e1.record()
    _sync_all(world)
    ms = e0.elapsed_time(e1)
    loss_final = float(tr.losses_red[0].item()) / world

    # e2e: public API, pinned-host inputs, loss read back every step
    for i in range(3):
        tr.step(xs_host[i % n_in], prefetch=xs_host[(i + 1) % n_in]); tr.losses_red.cpu()
    _sync_all(world)
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for i in range(args.steps):
        # the next batch's H2D copy is issued inside this call (prefetch=) and runs under this step's kernels; every
        # step still copies its own 3 MB from pinned host memory and reads its loss back
        tr.step(xs_host[i % n_in], prefetch=xs_host[(i + 1) % n_in])
        _ = tr.losses_red.cpu()
    e3.record()
    _sync_all(world)
    ms_e2e = e2.elapsed_time(e3)

    # bulk tokenisation (encode + VQ), the metric's literal reading, on EVERY rank (the dataset shards over ranks, no
    # collective): resident inputs, and end to end = pinned-host actions in, int64 indices back to pinned host memory
    idx_host = torch.empty(N_TOK, dtype=torch.int64).pin_memory()
    # (the weights are fixed while a dataset is tokenised: the first call derives the clamped to_latent weight and
    #  prepares the codebook for the search, the following calls -- prepared=True -- are three linear layers + one search)
    tr.tokenize(xs_dev[0])
    for i in range(3):
        tr.tokenize(xs_dev[i % n_in], prepared=True)
    _sync_all(world)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(args.steps):
        tr.tokenize(xs_dev[i % n_in], prepared=True)
    t1.record()
    _sync_all(world)
    ms_tok = t0.elapsed_time(t1)
    # the same with the encoder in its bf16 mode (bf16 operands, fp32 accumulation): reported beside the fp32 number
    tr.tokenize(xs_dev[0], bf16=True)
    for i in range(3):
        tr.tokenize(xs_dev[i % n_in], prepared=True, bf16=True)
    _sync_all(world)
    tb0, tb1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tb0.record()
    for i in range(args.steps):
        tr.tokenize(xs_dev[i % n_in], prepared=True, bf16=True)
    tb1.record()
    _sync_all(world)
    ms_tok_bf16 = tb0.elapsed_time(tb1)
    tr.tokenize(xs_dev[0])  # (back to the fp32 state of the prepared buffers)
    for i in range(2):
        idx_host.copy_(tr.tokenize(xs_host[i % n_in], prepared=True), non_blocking=True)
    _sync_all(world)
    t2, t3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t2.record()
    for i in range(args.steps):
        idx_host.copy_(tr.tokenize(xs_host[i % n_in], prepared=True), non_blocking=True)
    t3.record()
    _sync_all(world)
    ms_tok_e2e = t2.elapsed_time(t3)
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([ms, ms_e2e, ms_tok, ms_tok_e2e, ms_tok_bf16], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_tok, ms_tok_e2e, ms_tok_bf16 = (float(v) for v in t)

    if rank == 0:
        hbm, tf_burst, tf_sus, src = _peaks()
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
        # --- dominant kernel of the headline target: the fused distance+argmin search, timed alone on the step's own
        #     z_e / codebook, L2 flushed (written) before every call ---
        z_e, cb = tr.bufs.t["z_e"], tr.p["quantizer.codebook"]
        algo = ops.pick_algo(N_TOK, D, K)
        ws = tr.bufs.t["ws"]
        idx_buf = tr.bufs.t["idx"]
        un, un_min = _time_events(lambda: ops.nn_search(z_e, cb, lipvq.METRIC_L2, algo, out=idx_buf, workspace=ws), 20, flush)
        # ten copies of the step's latents (10 x 16.8 MB > the 126 MB L2), used in turn: every call reads its rows from HBM
        zrot = [z_e.clone() for _ in range(10)]
        if tr.search_prepared:
            # the product path: the codebook is prepared once per codebook version (after each training step's update,
            # once for a tokenisation run); a search is then ONE launch
            pu, _ = _time_events(lambda: ops.nn_search_prepare(cb, ws, lipvq.METRIC_L2, center=z_e), 20, flush)
            call = lambda zz: ops.nn_search_prepared(zz, cb, ws, lipvq.METRIC_L2, out=idx_buf)
            wf, wf_min = _time_events(lambda: call(z_e), 20, flush)
            _, st = ops.nn_search_prepared(z_e, cb, ws, lipvq.METRIC_L2, want_stats=True)
            kname = ("lipvq_nn_search_prepared_f32 = search_p1_kernel, ONE launch (bulk copy of the prepared code image, fp16 "
                     "tcgen05 pass with a best-only fold + exact near-best mask, certification, in-kernel settling of "
                     "uncertified rows); the codebook-only preparation (search_prep_kernel, once per codebook version) is "
                     "timed separately as us_prepare; us_unprepared = lipvq_nn_search_f32 (one call doing both)")
        else:
            pu = None
            call = lambda zz: ops.nn_search(zz, cb, lipvq.METRIC_L2, algo, out=idx_buf, workspace=ws)
            wf, wf_min = un, un_min
            _, st = ops.nn_search(z_e, cb, lipvq.METRIC_L2, algo, want_stats=True, workspace=ws)
            kname = "lipvq_nn_search_f32; algo=" + {1: "simt_f32", 2: "tcgen05"}[algo]
        iso, iso_min = _time_rotating(call, zrot, 40)
        sd, sd_min = _time_rotating(call, zrot, 200, streamed=True)
        del zrot
        flops = 2.0 * N_TOK * K * D
        bytes_alg = N_TOK * (4 * D + 8) + K * D * 4
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "search_traffic.json")) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        except Exception:
            pass
        step_us = ms / args.steps * 1e3
        extra["roofline"] = {"bound": "tensor", "kernel": kname,
                             "achieved": flops / sd / 1e6, "peak": tf_burst, "unit": "TFLOP/s",
                             "frac": flops / sd / 1e6 / tf_burst, "traffic": traffic,
                             "peak_source": src + " (bf16 dense burst)", "us_per_launch": sd, "us_per_launch_min": sd_min,
                             "timing": "us_per_launch: 200 back-to-back launches (a captured graph of ten, replayed twenty times) "
                                       "between ONE CUDA-event pair, total / 200, the inputs rotating through ten copies of the "
                                       "step's latents (168 MB > L2: every launch reads its rows from HBM; no flush); us_per_launch_isolated: the same rotation with an "
                                       "event pair and a synchronize around every single launch (adds the launch latency of a "
                                       "lone kernel); us_per_launch_writeflush: isolated, with 256 MB written through the L2 "
                                       "before every launch (the L2 is then full of dirty lines that every miss must evict "
                                       "first; the r1 / r2a way of measuring)",
                             "us_per_launch_isolated": iso, "us_per_launch_isolated_min": iso_min,
                             "us_per_launch_writeflush": wf, "us_per_launch_writeflush_min": wf_min,
                             "frac_isolated": flops / iso / 1e6 / tf_burst, "frac_writeflush": flops / wf / 1e6 / tf_burst,
                             "us_prepare": pu, "us_unprepared": un, "us_unprepared_min": un_min,
                             "algorithmic_flops": flops, "algorithmic_bytes": bytes_alg,
                             "hbm_frac": bytes_alg / sd / 1e3 / hbm, "share_of_step": sd / step_us,
                             "rows_rescored_exactly": int(st[0]), "rows_uncertified_by_fp16_pass": int(st[1]),
                             "codes_evaluated_exactly_for_those_rows": int(st[2]),
                             "l2": "inputs larger than L2 (ten 16.8 MB latent buffers in turn); see timing"}
        # --- the step-dominant kernel: the tensor-core linear layer (11 forward / dX launches + 6 dW launches per step).
        #     Representative launch: encoder.2 (64 -> 128, GELU + derivative), HBM-bound by bytes ---
        a1, a2, da2 = tr.bufs.t["a1"], tr.bufs.t["a2"], tr.bufs.t["da2_act"]
        W, b = tr.p["encoder.2.weight"], tr.p["encoder.2.bias"]
        lu, lu_min = _time_events(lambda: ops.linear_fwd_dual(a1, lipvq.ACT_NONE, W, b, lipvq.ACT_GELU, a2, da2), 20, flush)
        lbytes = N_TOK * (64 + 128 + 128) * 4 + W.numel() * 4
        extra["roofline_step"] = {"bound": "hbm", "kernel": "linear_tc_kernel (lipvq_linear_fwd_dual_f32, 64 -> 128, GELU and its "
                                  "derivative in the epilogue)", "achieved": lbytes / lu / 1e3, "peak": hbm, "unit": "GB/s",
                                  "frac": lbytes / lu / 1e3 / hbm, "traffic": None, "us_per_launch": lu, "us_per_launch_min": lu_min,
                                  "algorithmic_bytes": lbytes, "peak_source": src + " (copy bandwidth)",
                                  "launches_of_this_family_per_step": 17}
        if not args.no_sweep:
            extra["sweep"] = search_sweep(dev, hbm, tf_burst, flush)
        del flush
        if not args.no_t2 and world == 1:
            extra["t2"] = t2_leg(dev)

    if world > 1:
        dist.barrier()
    if rank == 0:
        chunks = world * CHUNKS * args.steps
        line = {"metric": "action_chunks_per_sec", "value": chunks / (ms * 1e-3), "unit": "chunks/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "tokens_per_gpu": N_TOK, "parallelism": f"dp{world}",
                           "l2": "per-step working set (activations+gradients ~0.3 GB) exceeds the 126 MB L2; "
                                 "4 input batches rotate", "cuda_graph": True, "final_loss": loss_final,
                           "collective": tr.collective, "collective_check": collective_check},
                "e2e": {"value": chunks / (ms_e2e * 1e-3), "unit": "chunks/s",
                        "h2d_bytes_per_step": N_TOK * F_ * 4, "d2h_bytes_per_step": 16},
                "tokenize": {"value": chunks / (ms_tok * 1e-3), "unit": "chunks/s", "ms_per_step": ms_tok / args.steps,
                             "note": "encode+VQ only (the metric's literal reading), every rank on its own shard, inputs resident; "
                                     "fixed weights: clamped to_latent weight and prepared codebook derived once before the loop"},
                "tokenize_bf16": {"value": chunks / (ms_tok_bf16 * 1e-3), "unit": "chunks/s", "ms_per_step": ms_tok_bf16 / args.steps,
                                  "note": "as tokenize, the encoder's matrix products with bf16 operands (fp32 accumulation, "
                                          "bias, activation and I/O; latents within 1e-2 of the fp32 path); NOT the headline dtype"},
                "tokenize_e2e": {"value": chunks / (ms_tok_e2e * 1e-3), "unit": "chunks/s", "ms_per_step": ms_tok_e2e / args.steps,
                                 "h2d_bytes_per_step": N_TOK * F_ * 4, "d2h_bytes_per_step": N_TOK * 8,
                                 "note": "TokenizerTrainer.tokenize(pinned host actions) -> int64 indices copied back to pinned "
                                         "host memory, every step, every rank"},
                "gpu_launches": int(launches_per_step) * args.steps, "clocks": clocks}
        line.update(extra)
    _teardown(tr, world)
    if rank == 0:
        # The timed CPU legs run on rank 0 AFTER the process group is gone and the other ranks have left: no GPU sits in a
        # device-side barrier while the host cores work (round 1 kept seven B200s spinning for half a minute here).
        if not args.no_cpu_baseline:
            print("[bench] GPU section done; timing the CPU legs on rank 0", file=sys.stderr, flush=True)
            line["cpu_baseline"] = cpu_baseline()
            line["cpu_baseline_1thread"] = cpu_baseline(max_seconds=12.0, threads=1, sample_chunks=128)
        print(json.dumps(line), flush=True)


def _teardown(tr, world):
    """Release the captured graphs BEFORE the communicator (NCCL's destructor can block while a live graph still holds
    its kernels), then destroy the group.  The destroy runs in a helper thread with a deadline: ranks > 0 leave if it does
    not finish, rank 0 carries on (it still has the CPU legs to time and the line to print)."""
    import gc
    import torch.distributed as dist
    if world <= 1:
        return
    sys.stdout.flush()
    torch.cuda.synchronize()
    dist.barrier()
    for name in ("_graph_full", "_graph_a", "_graph_b", "_graph_tok", "_graph_tok_fast", "_graph_tok_bf16", "_graph_tok_fast_bf16"):
        setattr(tr, name, None)
    gc.collect()
    torch.cuda.synchronize()
    th = threading.Thread(target=dist.destroy_process_group, daemon=True)
    th.start()
    th.join(30.0)
    if th.is_alive() and int(os.environ.get("RANK", "0")) != 0:
        os._exit(0)


# ----------------------------------------------------------------------------------------------
# BASELINE config 5: ICRT + LipVQ-VAE in-context policy training step (bf16 transformer, random-init weights, synthetic
# multi-trajectory prompts).  Shipped shapes: B = 16 prompt + 16 query trajectories, T = 10, observation width 208,
# A = 12, embed 512, 6 layers, 8 heads, 5-mode GMM (config/icl_config.py:118-135, train_config.json).
# ----------------------------------------------------------------------------------------------
C5 = dict(B=16, T=10, A=12, embed=512, layers=6, heads=8)
C5_OBS = (("robot0_eef_pos", (3,)), ("robot0_eef_quat", (4,)), ("robot0_gripper_qpos", (2,)), ("lang_emb", (199,)))
C5_WORKLOAD = ("ICRT in-context policy training step with the LipVQ-VAE tokenizer (T1 protocol: tokenizer AdamW step, then "
               "policy Adam step, clip 100), 16 prompt + 16 query trajectories x T=10 per GPU, obs width 208, A=12, "
               "GPT 6 x 512 x 8 heads, 5-mode GMM")


def _c5_batch(seed, device=None, pin=False):
    from collections import OrderedDict
    g = torch.Generator().manual_seed(seed)
    B2, T_, A_ = 2 * C5["B"], C5["T"], C5["A"]
    obs = OrderedDict((k, torch.randn(B2, T_, *s, generator=g)) for k, s in C5_OBS)
    act = torch.rand(B2, T_, A_, generator=g) * 2 - 1
    if pin:
        obs = OrderedDict((k, v.pin_memory()) for k, v in obs.items())
        act = act.pin_memory()
    if device is not None:
        obs = OrderedDict((k, v.to(device)) for k, v in obs.items())
        act = act.to(device)
    return {"obs": obs, "actions": act}


def _c5_policy(factory=None, dropout=0.1):
    from collections import OrderedDict
    from lipvq import icrt
    return icrt.ICLTransformerGMMActorNetwork(
        obs_shapes=OrderedDict(C5_OBS), ac_dim=C5["A"], transformer_embed_dim=C5["embed"], transformer_num_layers=C5["layers"],
        transformer_num_heads=C5["heads"], transformer_context_length=C5["T"], transformer_vq_vae_enabled=True,
        transformer_nn_parameter_for_timesteps=True, num_modes=5, min_std=0.0001, transformer_emb_dropout=dropout,
        transformer_attn_dropout=dropout, transformer_block_output_dropout=dropout, action_network_factory=factory)


def run_c5(args):
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (use --impl reference for the CPU leg)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import lipvq
    from lipvq import icl_algo
    torch.manual_seed(0)
    policy = _c5_policy().to(dev).train()
    algo = icl_algo.ICLTransformerGMM(policy, bf16=True)
    hosts = [_c5_batch(100 + 7 * rank + i, pin=True) for i in range(4)]
    devs = [{"obs": {k: v.to(dev) for k, v in b["obs"].items()}, "actions": b["actions"].to(dev)} for b in hosts]
    c0 = lipvq.launch_count()
    for i in range(max(args.warmup, 3)):
        algo.train_on_batch(devs[i % 4])
    torch.cuda.synchronize()
    launches_per_step = (lipvq.launch_count() - c0) // max(args.warmup, 3)
    _sync_all(world)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        info = algo.train_on_batch(devs[i % 4])
    e1.record()
    _sync_all(world)
    ms = e0.elapsed_time(e1)
    # e2e: pinned-host prompts in (H2D inside), the two losses read back, every step
    to_dev = lambda b: {"obs": {k: v.to(dev, non_blocking=True) for k, v in b["obs"].items()},
                        "actions": b["actions"].to(dev, non_blocking=True)}
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for i in range(args.steps):
        info = algo.train_on_batch(to_dev(hosts[i % 4]))
        _ = (float(info["action_loss"]), float(info["vq_vae_loss"]))
    e3.record()
    _sync_all(world)
    ms_e2e = e2.elapsed_time(e3)
    # share of the step inside the tokenizer: its forward + backward + AdamW alone on the same prompt actions
    tok = policy.vq_vae_model
    opt = torch.optim.AdamW(tok.parameters(), lr=1e-3, weight_decay=1e-4)
    pa = devs[0]["actions"][:C5["B"]].reshape(-1, C5["A"]).contiguous()
    for _ in range(3):
        opt.zero_grad(); tok(pa)[1].backward(); opt.step()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(args.steps):
        opt.zero_grad(); tok(pa)[1].backward(); opt.step()
    t1.record()
    torch.cuda.synchronize()
    ms_tok = t0.elapsed_time(t1)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
    if rank == 0:
        bytes_in = sum(v.numel() * 4 for v in hosts[0]["obs"].values()) + hosts[0]["actions"].numel() * 4
        n_par = sum(p.numel() for p in policy.parameters())
        line = {"metric": "policy_training_steps_per_sec", "value": world * args.steps / (ms * 1e-3) / world, "unit": "steps/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": C5_WORKLOAD, "parallelism": f"dp{world}", "parameters": n_par,
                           "trajectories_per_sec": world * 2 * C5["B"] * args.steps / (ms * 1e-3),
                           "tokenizer": "lipvq.LLFQVAE_V4 (fp32 CUDA kernels; K=1024, D=208)", "transformer": "torch bf16 autocast + SDPA",
                           "l2": "4 prompt batches rotate; launch-bound at this size", "final_action_loss": float(info["action_loss"])},
                "e2e": {"value": args.steps / (ms_e2e * 1e-3), "unit": "steps/s", "h2d_bytes_per_step": bytes_in, "d2h_bytes_per_step": 8},
                "tokenizer_share_of_step": ms_tok / ms, "tokenizer_ms_per_step": ms_tok / args.steps,
                "gpu_launches": int(launches_per_step) * args.steps, "clocks": clocks}
        print(json.dumps(line), flush=True)
    if world > 1:
        sys.stdout.flush()
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()


def run_reference_c5(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import lipvq_oracle as O
    from lipvq import icl_algo  # pure torch (no kernel is touched on this arm: the tokenizer slot holds the oracle)
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    policy = _c5_policy(factory=lambda a, d: O.T1Module(a, d)).train()
    algo = icl_algo.ICLTransformerGMM(policy, bf16=False)
    batches = [_c5_batch(100 + i) for i in range(2)]
    for i in range(max(1, min(args.warmup, 2))):
        algo.train_on_batch(batches[i % 2])
    t0 = time.perf_counter()
    for i in range(args.steps):
        algo.train_on_batch(batches[i % 2])
    dt = (time.perf_counter() - t0) / args.steps
    line = {"impl": "reference", "metric": "policy_training_steps_per_sec", "value": 1.0 / dt, "unit": "steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": {"workload": C5_WORKLOAD},
            "cpu_baseline": {"value": 1.0 / dt, "unit": "steps/s", "cores": threads, "kind": "port",
                             "sample": "the same step on the host cores: restated ICRT policy (torch fp32) with the oracle "
                                       "tokenizer in the action_network slot"},
            "e2e": {"value": 1.0 / dt, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the timed CPU leg (profiling runs only)")
    ap.add_argument("--no-sweep", action="store_true", help="skip the codebook-size sweep (BASELINE config 3) in the line")
    ap.add_argument("--no-t2", action="store_true", help="skip the HierarchicalVQVAE leg (extra key t2) in the line")
    ap.add_argument("--config", default="c2", choices=["c2", "c5"],
                    help="c2: tokenizer training step (default, the headline); c5: ICRT+LipVQ-VAE policy training step")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_c5(args) if args.config == "c5" else run_reference(args)
    else:
        run_c5(args) if args.config == "c5" else run_ours(args)


if __name__ == "__main__":
    main()
