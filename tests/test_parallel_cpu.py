"""Host logic of the data-parallel step on CPU (gloo, world_size 2): shard ranges, the single reduction slab, and
that sum-all-reduce + 1/world scaling of per-shard results reproduces the full-batch gradients, EMA counts/sums and
losses of the oracle (which plays the role of the per-rank compute here)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def test_shard_range_partitions_exactly():
    from lipvq.parallel import shard_range
    for n in (0, 1, 7, 4096, 65537):
        for w in (1, 2, 3, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def test_slab_layout_offsets():
    from lipvq.parallel import SlabLayout
    L = SlabLayout(["a", "b"], [(3, 5), (7,)], K=8, D=4, ema=True)
    assert L.P == 22 and L.P4 == 24 and L.off_counts == 24 and L.off_sums == 32 and L.off_losses == 64 and L.length == 68
    slab = torch.arange(L.length, dtype=torch.float32)
    assert L.grad_view(slab, "b").tolist() == list(range(15, 22))
    assert L.sums(slab).shape == (8, 4) and L.losses(slab).tolist() == [64., 65., 66., 67.]
    L2 = SlabLayout(["a"], [(4,)], K=8, D=4, ema=False)
    assert L2.length == 8


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    import sys
    for p in (os.path.join(ROOT, "robot-manipulation-vq-vae_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        sys.path.insert(0, p)
    import lipvq_oracle as O
    from lipvq.parallel import SlabLayout, all_reduce_slab, shard_range
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    F_, D, K, N = 7, 16, 32, 64
    torch.manual_seed(0)
    p = O.t1_init(F_, D, K)
    g = torch.Generator().manual_seed(123)
    x = torch.rand(N, F_, generator=g) * 2 - 1
    names = list(p.keys())
    L = SlabLayout(names, [tuple(p[k].shape) for k in names], K, D, ema=True)
    b, e = shard_range(N, rank, world)
    leaves = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    out = O.t1_forward(leaves, x[b:e])
    out["loss"].backward()
    slab = torch.zeros(L.length)
    for k in names:
        L.grad_view(slab, k).copy_(leaves[k].grad)
    idx = out["indices"]
    L.counts(slab).copy_(torch.bincount(idx, minlength=K).float())
    L.sums(slab).index_add_(0, idx, out["z_e"].detach())
    L.losses(slab).copy_(torch.stack([out["loss"], out["recon_loss"], out["commitment_loss"], out["codebook_loss"]]).detach())
    w = all_reduce_slab(slab)
    assert w == world
    if rank == 0:
        full = {k: v.clone().requires_grad_(True) for k, v in p.items()}
        fo = O.t1_forward(full, x)
        fo["loss"].backward()
        ok = True
        for k in names:
            ok &= torch.allclose(L.grad_view(slab, k) / world, full[k].grad, rtol=1e-5, atol=1e-8)
        ok &= torch.equal(L.counts(slab), torch.bincount(fo["indices"], minlength=K).float())
        ref_sums = torch.zeros(K, D).index_add_(0, fo["indices"], fo["z_e"].detach())
        ok &= torch.allclose(L.sums(slab), ref_sums, rtol=1e-6, atol=1e-6)
        ok &= torch.allclose(L.losses(slab)[0] / world, fo["loss"].detach(), rtol=1e-6)
        ret.put(bool(ok))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_rank_reduction_matches_full_batch():
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p_ in procs:
        p_.start()
    ok = ret.get(timeout=150)
    for p_ in procs:
        p_.join(60)
        assert p_.exitcode == 0
    assert ok
