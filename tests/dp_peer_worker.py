"""Worker of tests/test_gpu_dp_peer.py (one process per GPU under torchrun): the NVLink peer all-reduce must give the
same training trajectory as the NCCL all-reduce, with bit-identical replicas on every rank, eagerly and in a graph."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))


def run(collective: str, use_graph: bool, dev, rank):
    import lipvq
    from lipvq.trainer import TokenizerTrainer
    os.environ["LIPVQ_DP_COLLECTIVE"] = collective
    torch.manual_seed(0)
    model = lipvq.LLFQVAE_V4(12, 64, 512)
    model.quantizer = lipvq.LFQQuantizerEMA(512, 64)
    model.to(dev).train()
    N = 8192
    tr = TokenizerTrainer(model, N, use_graph=use_graph)
    assert (tr.peer is not None) == (collective == "peer"), "peer path was not taken"
    losses = []
    for s in range(4):
        g = torch.Generator().manual_seed(100 + 7 * rank + s)
        x = torch.rand(N, 12, generator=g).to(dev)
        losses.append(tr.step(x).clone())
    torch.cuda.synchronize()
    return tr.flat_p.clone(), torch.stack(losses), model.quantizer.ema_codebook.clone()


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    for use_graph in (False, True):
        p_n, l_n, e_n = run("nccl", use_graph, dev, rank)
        p_p, l_p, e_p = run("peer", use_graph, dev, rank)
        # same sums up to the order of the W additions
        assert torch.allclose(p_p, p_n, rtol=1e-5, atol=1e-6), (p_p - p_n).abs().max().item()
        assert torch.allclose(l_p, l_n, rtol=1e-5, atol=1e-6), (l_p - l_n).abs().max().item()
        assert torch.allclose(e_p, e_n, rtol=1e-5, atol=1e-6)
        # replicas stay bit-identical: every rank added the copies in rank order
        gathered = [torch.empty_like(p_p) for _ in range(world)]
        dist.all_gather(gathered, p_p)
        assert all(torch.equal(gathered[0], t) for t in gathered), "replicas diverged"
    dist.barrier()
    torch.cuda.synchronize()
    if rank == 0:
        print("PEER_OK", flush=True)
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
