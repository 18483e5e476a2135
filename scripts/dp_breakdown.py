"""Where a data-parallel step spends its time (run under torchrun, N>=2): phase A graph, the slab all-reduce, phase B
graph, the whole step.  CUDA events on the current stream, max over ranks."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))


def timed(fn, n=30, pre=None):
    ts = []
    for i in range(n + 5):
        if pre:
            pre()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        if i >= 5:
            ts.append(a.elapsed_time(b))
    t = torch.tensor([sum(ts) / len(ts), min(ts)], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [round(float(x) * 1e3, 1) for x in t]


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import lipvq
    from lipvq.parallel import all_reduce_slab
    from lipvq.trainer import TokenizerTrainer
    torch.manual_seed(0)
    model = lipvq.LLFQVAE_V4(12, 64, 512)
    model.quantizer = lipvq.LFQQuantizerEMA(512, 64)
    model.to(dev).train()
    os.environ["LIPVQ_GRAPH_COLLECTIVE"] = "0"
    tr = TokenizerTrainer(model, 65536, use_graph=True)
    x = (torch.rand(65536, 12, device=dev) * 2 - 1)
    for _ in range(5):
        tr.step(x)
    torch.cuda.synchronize()
    res = {
        "slab_bytes": tr.red.numel() * 4,
        "phase_a_us": timed(tr._graph_a.replay),
        "all_reduce_us(after barrier)": timed(lambda: all_reduce_slab(tr.red, tr.pg), pre=dist.barrier),
        "all_reduce_small_us": timed(lambda: dist.all_reduce(tr.losses_red), pre=dist.barrier),
        "phase_b_us": timed(tr._graph_b.replay),
        "step_us(2 graphs + eager collective)": timed(lambda: tr.step(x)),
    }
    if dist.get_rank() == 0:
        print(res, flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    os._exit(0)


if __name__ == "__main__":
    main()
