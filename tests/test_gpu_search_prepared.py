"""-m gpu: the search split in two -- lipvq_nn_search_prepare_f32 (once per codebook version) and
lipvq_nn_search_prepared_f32 (one launch per search: bulk copy of the prepared image, fp16 tcgen05 pass, in-kernel exact
scan of the rows it cannot certify) -- against the CPU oracle and against lipvq_nn_search_f32, through the C-ABI.
Reference op: LFQQuantizer.forward's [N,K,D] broadcast + norm + argmin (backbone_lfqvae_v5.py:39-46)."""
import math

import pytest
import torch

import lipvq_oracle as O
from test_gpu_kernels import DEV, _rand_case, assert_indices

pytestmark = pytest.mark.gpu


def _ws(L, N, D, K):
    return torch.empty(L.ops.search_workspace_bytes(N, D, K, L.METRIC_L2, L.ALGO_TCGEN05), dtype=torch.uint8, device=DEV)


@pytest.mark.parametrize("N,K,D", [
    (65536, 512, 64),    # the bench shape
    (4096, 512, 64),     # BASELINE C1
    (9000, 77, 12),      # ragged: D padded 12 -> 64, K padded to one tile, N not a multiple of 128
    (5000, 130, 256),    # four latent chunks
    (6000, 300, 30),     # D % 4 != 0: rows loaded from global, scalar staging of the codebook for the exact scan
    (8192, 1024, 64),    # resident codebook too large to stage in fp32: the exact scan reads it from global memory
])
@pytest.mark.parametrize("latents", ["uniform", "t1"])
def test_prepared_search_vs_oracle_and_unprepared(N, K, D, latents):
    import lipvq as L
    assert L.ops.search_prepared_supported(1 << 22, D, K)
    g = torch.Generator().manual_seed(7 + N + K + D)
    z, cb = _rand_case(N, K, D, 300 + N + K + D)
    if latents == "t1":  # sigmoid latents of a trained encoder: a tight cloud around 0.5, codes drawn from it
        z = torch.sigmoid(0.3 * torch.randn(N, D, generator=g))
        cb = z[torch.randperm(N, generator=g)[:K]] + 0.02 * torch.randn(K, D, generator=g)
    zd, cd = z.to(DEV), cb.to(DEV)
    ws = _ws(L, N, D, K)
    ref_rows = torch.randperm(N, generator=torch.Generator().manual_seed(0))[:2048]
    ref = O.nearest_code_l2(z[ref_rows], cb)
    base = L.ops.nn_search(zd, cd, L.METRIC_L2, L.ALGO_TCGEN05)
    for center in (None, zd):
        L.ops.nn_search_prepare(cd, ws, L.METRIC_L2, center=center)
        idx, st = L.ops.nn_search_prepared(zd, cd, ws, L.METRIC_L2, want_stats=True)
        idx2, st2 = L.ops.nn_search_prepared(zd, cd, ws, L.METRIC_L2, want_stats=True)  # counters reset themselves
        torch.cuda.synchronize()
        assert torch.equal(idx, idx2) and torch.equal(st, st2)
        assert 0 <= int(st[0]) == int(st[1]) <= N
        assert_indices(z[ref_rows], cb, idx.cpu()[ref_rows], ref)
        assert_indices(z, cb, idx, base.cpu())
    if latents == "t1":
        assert int(st[0]) < 0.10 * N, f"{int(st[0])} of {N} rows left to the exact scan with centred latents"


def test_prepared_search_other_metrics_and_sizes():
    """L2SQ and DOT through the prepared pair; one workspace prepared once serves several N."""
    import lipvq as L
    z, cb = _rand_case(20000, 512, 64, 42)
    zd, cd = z.to(DEV), (cb * 0.01).to(DEV)
    ws = _ws(L, 20000, 64, 512)
    L.ops.nn_search_prepare(cd, ws, L.METRIC_L2SQ)
    for n in (20000, 8192, 5001):
        idx = L.ops.nn_search_prepared(zd[:n], cd, ws, L.METRIC_L2SQ)
        assert_indices(z[:n], cb * 0.01, idx, O.nearest_code_l2(z[:n], cb * 0.01, squared=True), max_near_tie_frac=0.05)
    zn = torch.nn.functional.normalize(torch.randn(9000, 64), dim=-1)
    cn = torch.nn.functional.normalize(torch.randn(257, 64), dim=-1)
    ws = _ws(L, 9000, 64, 257)
    L.ops.nn_search_prepare(cn.to(DEV), ws, L.METRIC_DOT)
    idx = L.ops.nn_search_prepared(zn.to(DEV), cn.to(DEV), ws, L.METRIC_DOT)
    assert_indices(zn, cn, idx, torch.argmax(zn @ cn.t(), dim=1), metric="dot")


def test_unprepared_workspace_is_slow_not_wrong():
    """A workspace that was never prepared (or prepared for another shape) is detected on the device: nothing is
    certified and every row takes the exact scan."""
    import lipvq as L
    N, K, D = 8192, 512, 64
    z, cb = _rand_case(N, K, D, 5)
    zd, cd = z.to(DEV), cb.to(DEV)
    ws = _ws(L, N, D, K)
    ws.zero_()
    idx, st = L.ops.nn_search_prepared(zd, cd, ws, L.METRIC_L2, want_stats=True)
    assert int(st[0]) == N
    assert_indices(z, cb, idx, O.nearest_code_l2(z, cb))
    L.ops.nn_search_prepare(cd, ws, L.METRIC_L2SQ)  # another metric
    idx, st = L.ops.nn_search_prepared(zd, cd, ws, L.METRIC_L2, want_stats=True)
    assert int(st[0]) == N
    assert_indices(z, cb, idx, O.nearest_code_l2(z, cb))
    # the unprepared entry point on the same workspace leaves a prepared state intact
    L.ops.nn_search_prepare(cd, ws, L.METRIC_L2)
    a = L.ops.nn_search_prepared(zd, cd, ws, L.METRIC_L2).clone()
    b = L.ops.nn_search(zd, cd, L.METRIC_L2, L.ALGO_TCGEN05, workspace=ws)
    c = L.ops.nn_search_prepared(zd, cd, ws, L.METRIC_L2)
    assert torch.equal(a, c)
    assert_indices(z, cb, a, b.cpu())


def test_prepared_search_degenerate_codebooks():
    """Duplicates resolve to the lowest index, huge-norm codes (EMA rows divided by ~eps, hq_vae.py:277-279) are
    excluded from the contraction without disturbing the result, NaN rows get index 0 like the exact-scan kernels."""
    import lipvq as L
    N, K, D = 8192, 512, 64
    z, cb = _rand_case(N, K, D, 77)
    cb[100:200] = torch.randn(100, D) * 1e5
    cb[300] = cb[7]
    z[:64] = cb[7] + 1e-3
    z[64] = float("nan")
    zd, cd = z.to(DEV), cb.to(DEV)
    ws = _ws(L, N, D, K)
    L.ops.nn_search_prepare(cd, ws, L.METRIC_L2, center=zd[128:])
    idx, st = L.ops.nn_search_prepared(zd, cd, ws, L.METRIC_L2, want_stats=True)
    ok = torch.ones(N, dtype=torch.bool)
    ok[64] = False
    assert_indices(z[ok], cb, idx.cpu()[ok], O.nearest_code_l2(z[ok], cb))
    assert (idx[:64] == 7).all() and int(idx[64]) == 0
    assert int(st[0]) < 0.3 * N
    same = torch.zeros(K, D) + 0.25   # a collapsed codebook: every row ties over all codes -> index 0 everywhere
    L.ops.nn_search_prepare(same.to(DEV), ws, L.METRIC_L2)
    idx = L.ops.nn_search_prepared(zd, same.to(DEV), ws, L.METRIC_L2)
    assert int(idx.max()) == 0


def test_prepared_search_mostly_dead_ema_codebook():
    """An EMA codebook early in training: a dozen live codes, every other row divided by ~eps (hq_vae.py:277-279).  The
    huge rows are left out of the fp16 contraction; rows the pass cannot certify must not pay an exact evaluation of all
    of them (the per-code triangle-inequality test of the in-kernel scan)."""
    import lipvq as L
    N, K, D = 65536, 512, 64
    g = torch.Generator().manual_seed(3)
    z = torch.sigmoid(0.2 * torch.randn(N, D, generator=g))
    cb = torch.randn(K, D, generator=g) * 3e4
    live = torch.randperm(K, generator=g)[:12]
    cb[live] = z[torch.randperm(N, generator=g)[:12]] + 0.01 * torch.randn(12, D, generator=g)
    cb[live[1]] = cb[live[0]] + 1e-6     # a near-duplicate pair: the rows nearest to it cannot be certified in fp16
    zd, cd = z.to(DEV), cb.to(DEV)
    ws = _ws(L, N, D, K)
    L.ops.nn_search_prepare(cd, ws, L.METRIC_L2, center=zd)
    idx, st = L.ops.nn_search_prepared(zd, cd, ws, L.METRIC_L2, want_stats=True)
    rows = torch.randperm(N, generator=torch.Generator().manual_seed(0))[:4096]
    assert_indices(z[rows], cb, idx.cpu()[rows], O.nearest_code_l2(z[rows], cb), max_near_tie_frac=0.5)
    assert torch.isin(idx.cpu(), live).all()
    assert int(st[0]) > 0 and int(st[3]) == 0
    assert int(st[2]) <= 16 * int(st[0]), f"{int(st[2])} exact evaluations for {int(st[0])} queued rows"


def test_trainer_steps_and_tokenizes_on_the_prepared_search():
    """TokenizerTrainer prepares the codebook at the end of every step; the indices of a step and of the two tokenize
    modes equal those of the unprepared entry point on the same latents."""
    import lipvq as L
    from lipvq.trainer import TokenizerTrainer
    torch.manual_seed(0)
    N = 8192
    m = L.LLFQVAE_V4(12, 64, 512)
    m.quantizer = L.LFQQuantizerEMA(512, 64)
    m.to(DEV).train()
    tr = TokenizerTrainer(m, N)
    assert tr.search_prepared
    x = torch.rand(N, 12, device=DEV)
    for _ in range(3):
        tr.step(x)
        torch.cuda.synchronize()
    # the step's indices were computed with the codebook BEFORE the step's update: recompute with tokenize instead
    a = tr.tokenize(x).clone()
    ref = L.ops.nn_search(tr.bufs.t["z_e"], tr.p["quantizer.codebook"], L.METRIC_L2, L.ALGO_AUTO)
    assert_indices(tr.bufs.t["z_e"], tr.p["quantizer.codebook"], a, ref.cpu())
    b = tr.tokenize(x, prepared=True).clone()
    c = tr.tokenize(x, prepared=True)
    assert torch.equal(a, b) and torch.equal(a, c)
    with torch.no_grad():   # parameters changed behind the trainer's back -> refresh()
        tr.p["quantizer.codebook"].mul_(-1.0)
    tr.refresh()
    d = tr.tokenize(x, prepared=True)
    ref = L.ops.nn_search(tr.bufs.t["z_e"], tr.p["quantizer.codebook"], L.METRIC_L2, L.ALGO_AUTO)
    assert_indices(tr.bufs.t["z_e"], tr.p["quantizer.codebook"], d, ref.cpu())


@pytest.mark.parametrize("K,D,metric", [(1024, 512, "dot"), (2048, 256, "l2"), (1024, 384, "dot")])
def test_long_queue_exact_scan_wide_batches_equal_one_row_batches(K, D, metric):
    """A collapsed codebook (the cosine quantisers of a HierarchicalVQVAE a few steps into training: codes with mutual cosine
    0.9999, exact duplicates) leaves EVERY row to the exact scan; its long-queue form (sixteen rows per batch) must give the
    one-row form's indices bit for bit, and the fp64 winner wherever the fp64 gap is above fp32 rounding of the scores."""
    import os
    import lipvq as L
    g = torch.Generator().manual_seed(5)
    N = 8192
    base = torch.nn.functional.normalize(torch.randn(1, D, generator=g), dim=1)
    cb = torch.nn.functional.normalize(base + 0.01 * torch.randn(K, D, generator=g) / D ** 0.5, dim=1)
    cb[K // 2:] = cb[:K // 2]      # every code twice: each row's winner is an exact tie (never certifiable); lower index wins
    z = torch.nn.functional.normalize(base + 0.3 * torch.randn(N, D, generator=g) / D ** 0.5, dim=1)
    zd, cd = z.to(DEV), cb.to(DEV)
    M = L.METRIC_DOT if metric == "dot" else L.METRIC_L2
    idx, st = L.ops.nn_search(zd, cd, M, L.ALGO_TCGEN05, want_stats=True)
    os.environ["LIPVQ_RESCORE_WIDE"] = "0"
    try:
        idx1 = L.ops.nn_search(zd, cd, M, L.ALGO_TCGEN05)
    finally:
        del os.environ["LIPVQ_RESCORE_WIDE"]
    torch.cuda.synchronize()
    assert st[0].item() > N // 2, st.tolist()              # the regime this test is about
    assert torch.equal(idx, idx1)
    assert (idx >= K // 2).sum().item() == 0
    h = cb[:K // 2].double()
    s = z.double() @ h.T if metric == "dot" else -torch.cdist(z.double(), h)
    top = s.topk(2, dim=1)
    clear = (top.values[:, 0] - top.values[:, 1]) > 1e-5
    assert torch.equal(idx.cpu()[clear], top.indices[:, 0][clear])
