"""-m gpu: each kernel of liblipvq.so (through the C ABI) against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): indices bit-exact except near-ties (relative distance gap < 1e-6, counted);
gathered codewords bit-exact; fp32 values within 1e-4 relative (tighter where the arithmetic allows)."""
import math

import pytest
import torch

import lipvq_oracle as O
from conftest import load_golden, sub

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module")
def L():
    import lipvq
    return lipvq


def assert_indices(z, cb, got, ref, metric="l2", max_near_tie_frac=0.01):
    """exact match, or a documented near-tie: the two winners' distances differ by < 1e-6 relative (fp64 check)."""
    got, ref = got.cpu(), ref.cpu()
    bad = (got != ref).nonzero().flatten()
    if bad.numel() == 0:
        return 0
    zz = z.cpu()[bad].double()
    eg, er = cb.cpu()[got[bad]].double(), cb.cpu()[ref[bad]].double()
    if metric == "dot":
        vg, vr = (zz * eg).sum(-1), (zz * er).sum(-1)
        rel = (vg - vr).abs()
    else:
        vg, vr = (zz - eg).norm(dim=-1), (zz - er).norm(dim=-1)
        rel = (vg - vr).abs() / vr.clamp_min(1e-30)
    assert rel.max().item() < 1e-6, f"{bad.numel()} index mismatches, worst relative gap {rel.max().item():.3e}"
    assert bad.numel() <= max(1, int(max_near_tie_frac * got.numel())), f"too many near-ties: {bad.numel()}"
    return bad.numel()


def _rand_case(N, K, D, seed):
    g = torch.Generator().manual_seed(seed)
    z = torch.rand(N, D, generator=g)
    b = math.sqrt(6.0 / D)
    cb = (torch.rand(K, D, generator=g) * 2 - 1) * b
    return z, cb


# ------------------------------------------------------------------------------------------------ K3
@pytest.mark.parametrize("N,K,D", [(256, 64, 32), (4096, 512, 64), (80, 1024, 208), (1, 5, 7), (1000, 77, 12),
                                   (300, 130, 256), (5000, 4096, 64)])
def test_search_l2_vs_oracle(L, N, K, D):
    z, cb = _rand_case(N, K, D, 100 + N + K + D)
    ref = O.nearest_code_l2(z, cb)
    idx, best, stats = L.ops.nn_search(z.to(DEV), cb.to(DEV), L.METRIC_L2, L.ALGO_SIMT_F32, want_best=True, want_stats=True)
    assert idx.dtype == torch.int64 and idx.shape == (N,)
    assert_indices(z, cb, idx, ref)
    dref = torch.norm(z - cb[ref], dim=-1)
    assert torch.allclose(best.cpu(), dref, rtol=1e-5, atol=1e-7)
    assert 0 <= int(stats[0]) <= N


def test_search_golden_quantizer_fixture(L):
    g = load_golden("quantizers")
    cb = sub(g, "pre")["codebook"]
    idx = L.ops.nn_search(g["z"].to(DEV), cb.to(DEV))
    assert_indices(g["z"], cb, idx, g["indices"])


def test_search_sqdist_and_dot_metrics(L):
    z, cb = _rand_case(2048, 300, 48, 7)
    cb = cb * 0.01  # classic VQ-VAE init scale U(-1/K,1/K): distances nearly equal -> exact path decides
    ref = O.nearest_code_l2(z, cb, squared=True)
    idx = L.ops.nn_search(z.to(DEV), cb.to(DEV), L.METRIC_L2SQ, L.ALGO_SIMT_F32)
    assert_indices(z, cb, idx, ref, max_near_tie_frac=0.05)
    zn = torch.nn.functional.normalize(torch.randn(1500, 64), dim=-1)
    cn = torch.nn.functional.normalize(torch.randn(257, 64), dim=-1)
    ref = torch.argmax(zn @ cn.t(), dim=1)
    idx, best = L.ops.nn_search(zn.to(DEV), cn.to(DEV), L.METRIC_DOT, L.ALGO_SIMT_F32, want_best=True)
    assert_indices(zn, cn, idx, ref, metric="dot")
    assert torch.allclose(best.cpu(), (zn @ cn.t()).max(dim=1).values, atol=1e-5)


def test_search_ties_take_lowest_index_and_edges(L):
    z, cb = _rand_case(512, 64, 32, 3)
    cb[40] = cb[3]
    cb[17] = cb[3]  # exact duplicates: torch.argmin returns the first
    z[:100] = cb[3] + 1e-3
    ref = O.nearest_code_l2(z, cb)
    idx, stats = L.ops.nn_search(z.to(DEV), cb.to(DEV), want_stats=True)
    assert_indices(z, cb, idx, ref)
    assert (idx[:100] == 3).all() and int(stats[0]) >= 100  # ties must have gone through the exact path
    # K = 1, and empty input
    one = L.ops.nn_search(z.to(DEV), cb[:1].contiguous().to(DEV))
    assert (one == 0).all()
    empty = L.ops.nn_search(torch.empty(0, 32, device=DEV), cb.to(DEV))
    assert empty.shape == (0,)
    with pytest.raises(RuntimeError):
        L.ops.nn_search(torch.zeros(4, 31, device=DEV), cb.to(DEV))


def test_search_full_size_properties(L):
    """BASELINE C3 size (N=65536, K=512, D=64): oracle on a row subsample + optimality of every row against random
    codes + determinism."""
    z, cb = _rand_case(65536, 512, 64, 11)
    zd, cd = z.to(DEV), cb.to(DEV)
    idx, best = L.ops.nn_search(zd, cd, want_best=True)
    idx2 = L.ops.nn_search(zd, cd)
    assert torch.equal(idx, idx2)
    rows = torch.randperm(65536, generator=torch.Generator().manual_seed(0))[:4096]
    ref = O.nearest_code_l2(z[rows], cb)
    assert_indices(z[rows], cb, idx.cpu()[rows], ref)
    # no random code may beat the reported winner by more than fp32 noise
    rk = torch.randint(0, 512, (65536, 8), generator=torch.Generator().manual_seed(1))
    d_rand = (z[:, None, :] - cb[rk]).norm(dim=-1)
    assert (best.cpu()[:, None] <= d_rand * (1 + 1e-6)).all()
    # histogram sanity: every index in range
    assert int(idx.min()) >= 0 and int(idx.max()) < 512


# ------------------------------------------------------------------------------------------------ K4
def test_gather_fwd_bwd(L):
    z, cb = _rand_case(1000, 50, 36, 5)
    idx = O.nearest_code_l2(z, cb)
    zd, cd, id_ = z.to(DEV), cb.to(DEV), idx.to(DEV)
    sq = torch.zeros(1, dtype=torch.float64, device=DEV)
    zq, zst = L.ops.gather_fwd(zd, id_, cd, sq, want_ste=True)
    assert torch.equal(zq.cpu(), cb[idx])  # bit-exact codewords
    assert torch.equal(zst.cpu(), z + (cb[idx] - z))  # the reference's STE expression, same roundings
    ref_sq = ((cb[idx] - z).double() ** 2).sum()
    assert abs(sq.item() - ref_sq.item()) <= 1e-6 * ref_sq.item()
    # backward vs autograd of 0.25*mse(zq.detach(), z) + 0.5*mse(zq, z.detach()) + <up, zq>
    up = torch.randn(1000, 36, generator=torch.Generator().manual_seed(9))
    zl, cl = z.clone().requires_grad_(True), cb.clone().requires_grad_(True)
    q = cl[idx]
    loss = 0.25 * torch.nn.functional.mse_loss(q.detach(), zl) + 0.5 * torch.nn.functional.mse_loss(q, zl.detach()) + (up * q).sum()
    (3.0 * loss).backward()
    dz = torch.empty_like(zd)
    dcb = torch.zeros_like(cd)
    gs = torch.tensor([3.0], device=DEV)
    L.ops.gather_bwd(zd, zq, id_, (3.0 * up).to(DEV), gs, 0.25, 0.5, False, None, 50, dz, dcb)
    assert torch.allclose(dz.cpu(), zl.grad, rtol=1e-5, atol=1e-9)
    assert torch.allclose(dcb.cpu(), cl.grad, rtol=1e-4, atol=1e-6)
    # STE routing + fused sigmoid derivative
    dz2 = torch.empty_like(zd)
    y = torch.sigmoid(torch.randn(1000, 36)).to(DEV)
    L.ops.gather_bwd(zd, zq, id_, up.to(DEV), None, 0.25, 0.5, True, y, 50, dz2, None)
    coef = 2.0 / (1000 * 36)
    ref = (0.25 * coef * (z - cb[idx]) + up) * (y.cpu() * (1 - y.cpu()))
    assert torch.allclose(dz2.cpu(), ref, rtol=1e-5, atol=1e-9)
    # scalar (D % 4 != 0) path
    z3, cb3 = _rand_case(100, 9, 7, 6)
    i3 = O.nearest_code_l2(z3, cb3)
    sq.zero_()
    zq3, _ = L.ops.gather_fwd(z3.to(DEV), i3.to(DEV), cb3.to(DEV), sq)
    assert torch.equal(zq3.cpu(), cb3[i3])


def test_gather_bwd_and_stats_shared_memory_path(L):
    """N >= 4096 with a table that fits shared memory: the privatised scatter kernels (quantize.cu), on a collapsed
    assignment (all tokens on 5 codes, the state of a fresh EMA quantizer) and on a spread one."""
    for spread in (False, True):
        N, K, D = 8192 + 37, 512, 64
        g = torch.Generator().manual_seed(31 + spread)
        z = torch.rand(N, D, generator=g)
        cb = torch.randn(K, D, generator=g)
        idx = torch.randint(0, K if spread else 5, (N,), generator=g)
        up = torch.randn(N, D, generator=g)
        zl, cl = z.clone().requires_grad_(True), cb.clone().requires_grad_(True)
        q = cl[idx]
        loss = 0.25 * torch.nn.functional.mse_loss(q.detach(), zl) + 0.25 * torch.nn.functional.mse_loss(q, zl.detach()) + (up * q).sum()
        loss.backward()
        zd, id_ = z.to(DEV), idx.to(DEV)
        zq = cb[idx].to(DEV)
        dz, dcb = torch.empty_like(zd), torch.zeros(K, D, device=DEV)
        L.ops.gather_bwd(zd, zq, id_, up.to(DEV), None, 0.25, 0.25, False, None, K, dz, dcb)
        assert torch.allclose(dz.cpu(), zl.grad, rtol=1e-5, atol=1e-9)
        assert torch.allclose(dcb.cpu(), cl.grad, rtol=1e-4, atol=1e-2 if not spread else 1e-4)
        counts, sums = torch.zeros(K, device=DEV), torch.zeros(K, D, device=DEV)
        L.ops.ema_stats(zd, id_, counts, sums)
        assert torch.equal(counts.cpu(), torch.bincount(idx, minlength=K).float())
        ref = torch.zeros(K, D, dtype=torch.float64).index_add_(0, idx, z.double()).float()
        assert torch.allclose(sums.cpu(), ref, rtol=1e-4, atol=1e-3)


def test_scatter_deterministic_sorted_path(L):
    """D % 4 == 0, D <= 128, K <= 8192: the stable counting sort + ordered segment sums (scatter_det.cu).  Sums must be
    bit-identical run to run (torch.equal), match index_add, and handle a collapsed assignment (one code owns everything:
    its segment spans every piece), empty codes, out-of-range indices (clamped) and N not a multiple of the piece size."""
    for case in ("spread", "collapsed", "one", "ragged"):
        N, K, D = {"spread": (65536, 512, 64), "collapsed": (8192 + 37, 512, 64), "one": (5000, 64, 32),
                   "ragged": (4097, 1000, 128)}[case]
        g = torch.Generator().manual_seed(77)
        z = torch.rand(N, D, generator=g)
        cb = torch.randn(K, D, generator=g)
        if case == "spread":
            idx = torch.randint(0, K, (N,), generator=g)
        elif case == "collapsed":
            idx = torch.randint(0, 5, (N,), generator=g) * 100
        elif case == "one":
            idx = torch.full((N,), 7, dtype=torch.int64)
        else:
            idx = torch.randint(-3, K + 3, (N,), generator=g)
        idc = idx.clamp(0, K - 1)
        up = torch.randn(N, D, generator=g)
        zd, id_, upd = z.to(DEV), idx.to(DEV), up.to(DEV)
        zq = cb[idc].to(DEV)
        outs = []
        ws = torch.empty(L.ops.code_sort_workspace_bytes(N, D, K), dtype=torch.uint8, device=DEV)
        assert ws.numel() > 0
        for _ in range(2):
            ws.fill_(0xAB)  # nothing may depend on the scratch's previous contents
            counts, sums = torch.zeros(K, device=DEV), torch.zeros(K, D, device=DEV)
            L.ops.code_sort(id_, D, K, ws)
            L.ops.ema_stats_sorted(zd, ws, counts, sums)
            dz, dcb = torch.empty_like(zd), torch.zeros(K, D, device=DEV)
            L.ops.gather_bwd_sorted(zd, zq, upd, None, 0.25, 0.25, False, None, K, ws, dz, dcb)
            outs.append((counts.cpu(), sums.cpu(), dz.cpu(), dcb.cpu()))
        for a, b in zip(*outs):
            assert torch.equal(a, b), case
        counts, sums, dz, dcb = outs[0]
        assert torch.equal(counts, torch.bincount(idc, minlength=K).float()), case
        ref = torch.zeros(K, D, dtype=torch.float64).index_add_(0, idc, z.double())
        assert torch.allclose(sums.double(), ref, rtol=1e-5, atol=1e-4 * max(1.0, float(counts.max()) / 1000)), case
        coef = 2.0 / (N * D)
        q = cb[idc]
        assert torch.allclose(dz, 0.25 * coef * (z - q), rtol=1e-5, atol=1e-9), case
        refg = torch.zeros(K, D, dtype=torch.float64).index_add_(0, idc, (0.25 * coef * (q - z) + up).double())
        assert torch.allclose(dcb.double(), refg, rtol=1e-4, atol=2e-3 if case in ("collapsed", "one") else 1e-4), case
        # accumulate semantics: a second call adds on top
        counts2, sums2 = counts.clone().to(DEV), sums.clone().to(DEV)
        L.ops.ema_stats_sorted(zd, ws, counts2, sums2)
        assert torch.equal(counts2.cpu(), 2 * counts), case
        assert torch.allclose(sums2.cpu(), 2 * sums, rtol=1e-6, atol=1e-6), case


# ------------------------------------------------------------------------------------------------ K5 / K6
@pytest.mark.parametrize("N,K,D", [(256, 64, 32), (4096, 512, 64), (333, 20, 208), (100, 9, 7)])
def test_ema_stats_and_apply(L, N, K, D):
    z, cb = _rand_case(N, K, D, 21)
    idx = O.nearest_code_l2(z, cb)
    if N >= 256:
        idx[:64] = idx[0]  # a run of equal codes exercises the run-length aggregation
    counts = torch.zeros(K, device=DEV)
    sums = torch.zeros(K, D, device=DEV)
    L.ops.ema_stats(z.to(DEV), idx.to(DEV), counts, sums)
    assert torch.equal(counts.cpu(), torch.bincount(idx, minlength=K).float())
    ref_sums = torch.zeros(K, D).index_add_(0, idx, z)
    assert torch.allclose(sums.cpu(), ref_sums, rtol=1e-5, atol=1e-5)
    g = torch.Generator().manual_seed(4)
    size0, sum0 = torch.rand(K, generator=g) * 5, torch.randn(K, D, generator=g)
    scratch = torch.zeros(1, dtype=torch.float64, device=DEV)
    for variant, fn in ((L.EMA_SIMPLE, O.ema_update_simple), (L.EMA_LAPLACE, O.ema_update_laplace),
                        (L.EMA_NORMALIZE, O.ema_update_cosine)):
        rs, re, rc = fn(z, idx, size0, sum0)
        s, e, c = size0.clone().to(DEV), sum0.clone().to(DEV), torch.empty(K, D, device=DEV)
        L.ops.ema_apply(counts, sums, s, e, c, 0.99, 1e-5, variant, scratch)
        assert torch.allclose(s.cpu(), rs, rtol=1e-6, atol=1e-7)
        assert torch.allclose(e.cpu(), re, rtol=1e-5, atol=1e-6)
        assert torch.allclose(c.cpu(), rc, rtol=1e-4, atol=1e-6), variant


def test_ema_golden_fixture_and_usage(L):
    g = load_golden("quantizers")
    pre, post = sub(g, "pre"), sub(g, "post")
    q = L.LFQQuantizerEMA(64, 32)
    q.load_state_dict(pre)
    q.to(DEV).train()
    z_q, idx = q(g["z"].to(DEV))
    assert torch.equal(idx.cpu(), g["indices"]) and torch.equal(z_q.cpu(), g["z_q"])
    sd = q.state_dict()
    for k in ("ema_cluster_size", "ema_codebook", "codebook"):
        assert torch.allclose(sd[k].cpu(), post[k], rtol=1e-5, atol=1e-6), k
    lp, lq = sub(g, "lap_pre"), sub(g, "lap_post")
    counts = torch.bincount(g["lap_indices"], minlength=64).float().to(DEV)
    uc, uma, ent = lp["usage_counts"].to(DEV), lp["usage_ma"].to(DEV), lp["entropy_ma"].reshape(1).to(DEV)
    L.ops.usage_stats(counts, uc, uma, ent)
    assert torch.allclose(uc.cpu(), lq["usage_counts"]) and torch.allclose(uma.cpu(), lq["usage_ma"])
    assert torch.allclose(ent.cpu().reshape(()), lq["entropy_ma"], rtol=1e-5)
    size = torch.tensor([0.0, 0.5, 2.0, 1.0, 0.0, 3.0], device=DEV)
    assert L.ops.codebook_usage(size, 1.0).tolist() == [4, 3]


# ------------------------------------------------------------------------------------------------ K1/K2/K7
ACTS = {0: lambda t: t, 1: torch.relu, 2: torch.nn.functional.gelu, 3: torch.sigmoid}


@pytest.mark.parametrize("N,i,o,act_in,act_out", [(300, 7, 64, 0, 0), (1000, 64, 128, 2, 0), (257, 128, 64, 2, 3),
                                                   (129, 128, 12, 2, 0), (64, 208, 64, 0, 1), (500, 64, 208, 1, 0),
                                                   (200, 512, 512, 1, 0), (1, 12, 7, 0, 0),
                                                   # N >= 4096: the tcgen05 three-way-bf16 path (linear_tc.cu)
                                                   (4096, 12, 64, 0, 0), (5000, 64, 128, 2, 0), (4500, 128, 64, 2, 3),
                                                   (4100, 128, 12, 2, 0), (6000, 64, 64, 0, 1), (4097, 7, 30, 1, 0),
                                                   (8192, 128, 128, 2, 2), (4200, 100, 208, 1, 0)])
def test_linear_fwd_bwd(L, N, i, o, act_in, act_out):
    g = torch.Generator().manual_seed(N + i + o)
    X = torch.randn(N, i, generator=g)
    W = torch.randn(o, i, generator=g) / math.sqrt(i)
    b = torch.randn(o, generator=g)
    ref = ACTS[act_out](torch.nn.functional.linear(ACTS[act_in](X), W, b))
    Y = L.ops.linear_fwd(X.to(DEV), act_in, W.to(DEV), b.to(DEV), act_out)
    assert torch.allclose(Y.cpu(), ref, rtol=1e-5, atol=2e-6)
    dY = torch.randn(N, o, generator=g)
    Xl, Wl, bl = X.clone().requires_grad_(True), W.clone().requires_grad_(True), b.clone().requires_grad_(True)
    (torch.nn.functional.linear(ACTS[act_in](Xl), Wl, bl) * dY).sum().backward()
    dX = L.ops.linear_bwd_x(dY.to(DEV), W.to(DEV), X.to(DEV), act_in)
    assert torch.allclose(dX.cpu(), Xl.grad, rtol=1e-4, atol=1e-5)
    dW, db = torch.zeros(o, i, device=DEV), torch.zeros(o, device=DEV)
    L.ops.linear_bwd_w(dY.to(DEV), X.to(DEV), act_in, dW, db)
    scale = math.sqrt(N)
    assert torch.allclose(dW.cpu(), Wl.grad, rtol=1e-4, atol=2e-5 * scale)
    assert torch.allclose(db.cpu(), bl.grad, rtol=1e-4, atol=2e-5 * scale)


@pytest.mark.parametrize("N,i,o,act", [(300, 12, 64, 2), (4096, 12, 64, 2), (5001, 64, 128, 2), (4500, 64, 64, 3),
                                        (4100, 128, 12, 1), (4099, 64, 200, 2), (777, 64, 128, 0)])
def test_linear_fwd_dual_and_given_derivative(L, N, i, o, act):
    """Training forward: activation value and derivative from one evaluation in the producing layer (tcgen05 epilogue
    for N >= 4096, FFMA tiles + elementwise pass below), consumed by linear_bwd_x with ACT_GIVEN."""
    g = torch.Generator().manual_seed(N + o)
    X = torch.randn(N, i, generator=g)
    W = torch.randn(o, i, generator=g) / math.sqrt(i)
    b = torch.randn(o, generator=g)
    pre = torch.nn.functional.linear(X.double(), W.double(), b.double()).requires_grad_(True)  # fp64 reference
    ref = ACTS[act](pre)
    ref.sum().backward()
    Y, dYd = torch.empty(N, o, device=DEV), torch.empty(N, o, device=DEV)
    L.ops.linear_fwd_dual(X.to(DEV), 0, W.to(DEV), b.to(DEV), act, Y, dYd)
    # fp32-accurate pre-activation (2e-6 of sum |x||w|, as in the wide-range test) through activations whose value and
    # derivative are at most 1.13-Lipschitz; relu'/given derivatives may flip only where |pre| is inside that bound
    bound = 2.5e-6 * (X.abs() @ W.abs().t() + b.abs()) + 1e-6
    assert ((Y.cpu().double() - ref.detach()).abs() <= bound).all()
    derr = (dYd.cpu().double() - pre.grad).abs()
    if act == 1:
        derr = derr * (pre.detach().abs() > bound)
    assert (derr <= bound).all()
    # next layer's input gradient through the stored derivative == through the activation itself
    o2 = 48
    W2 = torch.randn(o2, o, generator=g) / math.sqrt(o)
    dY2 = torch.randn(N, o2, generator=g)
    viaX = L.ops.linear_bwd_x(dY2.to(DEV), W2.to(DEV), pre.detach().float().to(DEV), act)
    given = L.ops.linear_bwd_x(dY2.to(DEV), W2.to(DEV), dYd, L.ACT_GIVEN)
    off = ((given - viaX).abs() > 2e-5 + 1e-4 * viaX.abs()).sum().item()
    assert off <= (o2 if act == 1 else 0)  # relu: one pre-activation within rounding of 0 may flip one row of dX


def test_linear_tc_bwd_w_plain_and_simt_agree(L, monkeypatch):
    """dW/db of a layer without input activation, ragged token count: the tcgen05 kernel (MN-major operands, default for
    N >= 4096) against fp64, and the FFMA tiles (LIPVQ_LINEAR_ALGO=simt) on the same data so both stay covered."""
    g = torch.Generator().manual_seed(5)
    N, i, o = 10007, 64, 64
    X, dY = torch.randn(N, i, generator=g), torch.randn(N, o, generator=g)
    dW, db = torch.zeros(o, i, device=DEV), torch.zeros(o, device=DEV)
    L.ops.linear_bwd_w(dY.to(DEV), X.to(DEV), 0, dW, db)
    assert torch.allclose(dW.cpu(), (dY.double().t() @ X.double()).float(), rtol=1e-4, atol=2e-3)
    assert torch.allclose(db.cpu(), dY.double().sum(0).float(), rtol=1e-4, atol=2e-3)
    monkeypatch.setenv("LIPVQ_LINEAR_ALGO", "simt")
    dW2, db2 = torch.zeros(o, i, device=DEV), torch.zeros(o, device=DEV)
    L.ops.linear_bwd_w(dY.to(DEV), X.to(DEV), 0, dW2, db2)
    assert torch.allclose(dW2, dW, rtol=1e-4, atol=2e-3) and torch.allclose(db2, db, rtol=1e-4, atol=2e-3)


def test_linear_tc_tiny_gradients_and_wide_range(L):
    """The tensor path must keep fp32 range: gradients of 1e-8 next to activations of 1e3 (a bf16 split keeps the
    8-bit exponent; an fp16 split would flush them)."""
    g = torch.Generator().manual_seed(11)
    N, i, o = 8192, 128, 64
    W = torch.randn(o, i, generator=g) / math.sqrt(i)
    dY = torch.randn(N, o, generator=g) * 1e-8
    X = torch.randn(N, i, generator=g) * torch.logspace(-3, 3, N).unsqueeze(1)
    dX = L.ops.linear_bwd_x(dY.to(DEV), W.to(DEV), X.to(DEV), 0)
    ref = (dY.double() @ W.double()).float()
    assert torch.allclose(dX.cpu(), ref, rtol=2e-5, atol=1e-12)
    Y = L.ops.linear_fwd(X.to(DEV), 0, W.to(DEV), None, 0)
    refY = (X.double() @ W.double().t()).float()
    err = (Y.cpu() - refY).abs() / (X.abs() @ W.abs().t())
    assert err.max().item() < 2e-6, f"forward error {err.max().item():.2e} relative to sum |x||w|"


def test_linear_bwd_w_many_token_slices(L):
    g = torch.Generator().manual_seed(2)
    N, i, o = 70000, 64, 128
    X, dY = torch.randn(N, i, generator=g), torch.randn(N, o, generator=g) * 0.01
    dW, db = torch.zeros(o, i, device=DEV), torch.zeros(o, device=DEV)
    L.ops.linear_bwd_w(dY.to(DEV), X.to(DEV), 0, dW, db)
    assert torch.allclose(dW.cpu(), (dY.double().t() @ X.double()).float(), rtol=1e-3, atol=1e-3)
    assert torch.allclose(db.cpu(), dY.double().sum(0).float(), rtol=1e-3, atol=1e-4)


def test_lipschitz_layer(L):
    g = torch.Generator().manual_seed(8)
    W = torch.randn(64, 128, generator=g)
    W[5] *= 0.001  # row with ||W||_1 < softplus(ci): scale clamps at 1 and passes no gradient through the norm
    ci = torch.ones(64) + 0.1 * torch.randn(64, generator=g)
    b = torch.randn(64, generator=g) * 0.1
    x = torch.randn(200, 128, generator=g)
    Wn = L.ops.lipschitz_scale_fwd(W.to(DEV), ci.to(DEV))
    assert torch.allclose(Wn.cpu(), W * O.linf_row_scale(W, ci), rtol=1e-6, atol=1e-8)
    m = L.LipschitzMLP(128, 64)
    m.load_state_dict({"W": W, "b": b, "ci": ci})
    m.to(DEV)
    xd = x.to(DEV).requires_grad_(True)
    y = m(xd)
    Wl, bl, cl, xl = (t.clone().requires_grad_(True) for t in (W, b, ci, x))
    yr = O.lipschitz_sigmoid(xl, Wl, bl, cl)
    assert torch.allclose(y.detach().cpu(), yr.detach(), rtol=1e-5, atol=1e-6)
    up = torch.randn(200, 64, generator=g)
    (y * up.to(DEV)).sum().backward()
    (yr * up).sum().backward()
    assert torch.allclose(xd.grad.cpu(), xl.grad, rtol=1e-4, atol=1e-6)
    assert torch.allclose(m.W.grad.cpu(), Wl.grad, rtol=1e-4, atol=1e-5)
    assert torch.allclose(m.b.grad.cpu(), bl.grad, rtol=1e-4, atol=1e-5)
    assert torch.allclose(m.ci.grad.cpu(), cl.grad, rtol=1e-4, atol=1e-5)
    assert m.ci.grad[5].item() == 0.0


def test_mse_and_adamw(L):
    g = torch.Generator().manual_seed(12)
    y, t = torch.randn(999, 7, generator=g), torch.randn(999, 7, generator=g)
    sq = torch.zeros(1, dtype=torch.float64, device=DEV)
    dy = torch.empty(999, 7, device=DEV)
    L.ops.mse_fwd_bwd(y.to(DEV), t.to(DEV), sq, gscale=torch.tensor([2.0], device=DEV), dY=dy)
    assert abs(sq.item() / y.numel() - torch.nn.functional.mse_loss(y, t).item()) < 1e-6
    assert torch.allclose(dy.cpu(), 2.0 * 2.0 / y.numel() * (y - t), rtol=1e-5, atol=1e-9)
    # AdamW: 3 steps against torch.optim.AdamW
    p0, grads = torch.randn(5001, generator=g), [torch.randn(5001, generator=g) for _ in range(3)]
    pr = p0.clone().requires_grad_(True)
    opt = torch.optim.AdamW([pr], lr=1e-3, weight_decay=1e-4)
    p, m, v = p0.clone().to(DEV), torch.zeros(5001, device=DEV), torch.zeros(5001, device=DEV)
    ctr = torch.zeros(1, dtype=torch.int32, device=DEV)
    for s, gr in enumerate(grads, 1):
        pr.grad = gr.clone()
        opt.step()
        if s < 3:
            L.ops.adamw_step(p, gr.to(DEV), m, v, s)
        else:  # device-resident step counter variant
            L.ops.counter_add(ctr, 3)
            L.ops.adamw_step(p, (gr * 2).to(DEV), m, v, 0, step_dev=ctr, grad_scale=0.5)
    # (the reference runs on whatever host CPU the box has: its vector width changes the last bits of sqrt/div chains)
    assert torch.allclose(p.cpu(), pr.detach(), rtol=2e-5, atol=1e-6)
