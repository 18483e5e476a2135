"""GPU parity of the other tokenizer families behind the reference's module API, through liblipvq.so:
T4 classic VQVAE (backbone.py:6-76), T2 HierarchicalVQVAE (heirarchical_vqvae.py), T3 LFQQuantizerEMA_KMeans /
HierarchicalLFQHVQVAE (hq_vae.py) -- against fixtures generated from the reference modules (oracle/gen_golden.py) and
against the CPU oracle on seeded inputs.  Tolerances: indices bit-exact, codewords bit-exact, losses/grads 1e-4 rel."""
import pytest
import torch

from conftest import load_golden, sub

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _close(a, b, rtol=1e-4, atol=1e-6, what=""):
    a, b = a.detach().cpu().float(), b.detach().cpu().float()
    assert a.shape == b.shape, f"{what}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    assert torch.allclose(a, b, rtol=rtol, atol=atol), f"{what}: max abs err {(a - b).abs().max():.3e}"


@pytest.fixture(scope="module")
def L():
    import lipvq
    return lipvq


# ------------------------------------------------------------------------------------------------ T4
def test_t4_classic_vs_reference_fixture(L):
    g = load_golden("t4_classic")
    m = L.VQVAE(7, 32, num_embeddings=128)
    m.load_state_dict(sub(g, "sd"), strict=True)
    m.to(DEV)
    c0 = L.launch_count()
    z_latent, loss = m(g["x"].to(DEV))
    assert z_latent.grad_fn is None and loss.dim() == 0
    assert torch.equal(m.last_indices.cpu(), g["indices"])
    _close(z_latent, g["z_latent"], rtol=1e-6, atol=1e-7, what="z_latent")
    _close(loss, g["loss"], what="loss")
    loss.backward()
    assert L.launch_count() - c0 >= 20
    for k, gr in sub(g, "grad").items():
        _close(dict(m.named_parameters())[k].grad, gr, rtol=1e-3, atol=1e-7, what="grad " + k)


def test_t4_larger_batch_vs_oracle(L):
    import lipvq_oracle as O
    torch.manual_seed(3)
    p = O.t4_init(12, 64, 256)
    with torch.no_grad():
        p["embedding.weight"].mul_(40.0)  # spread the codes so that many of them are selected
    x = torch.rand(2048, 12) * 2 - 1
    m = L.VQVAE(12, 64, num_embeddings=256)
    m.load_state_dict(p, strict=True)
    m.to(DEV)
    z_latent, loss = m(x.to(DEV))
    leaves = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    ref = O.t4_forward(leaves, x)
    assert int((m.last_indices.cpu() != ref["indices"]).sum()) == 0
    _close(z_latent, ref["z_latent"], rtol=1e-6, atol=1e-7)
    _close(loss, ref["loss"])
    loss.backward(), ref["loss"].backward()
    for k, v in m.named_parameters():
        _close(v.grad, leaves[k].grad, rtol=2e-3, atol=1e-7, what="grad " + k)


# ------------------------------------------------------------------------------------------------ K8 kernels
@pytest.mark.parametrize("B,T,E,dil", [(4, 16, 64, 1), (3, 10, 64, 2), (5, 16, 128, 4), (2, 16, 64, 16), (7, 12, 32, 8)])
def test_dilated_conv_layer_vs_torch(L, B, T, E, dil):
    torch.manual_seed(B * 100 + dil)
    conv = torch.nn.Conv1d(E, E, 3, padding=dil, dilation=dil)
    x = torch.randn(B, T, E)
    xr = x.clone().requires_grad_(True)
    ref = torch.relu(conv(xr.transpose(1, 2))) + xr.transpose(1, 2)
    ref = ref.transpose(1, 2)
    gout = torch.randn(B, T, E)
    ref.backward(gout)
    layer = L.DilatedTemporalConvLayer(E, E, 3, dilation=dil, dropout=0.0)
    layer.conv.load_state_dict(conv.state_dict())
    layer.to(DEV)
    xd = x.to(DEV).reshape(B * T, E).requires_grad_(True)
    out = layer.forward_tokens(xd, T)
    _close(out.reshape(B, T, E), ref, rtol=1e-4, atol=1e-5, what="conv fwd")
    out.backward(gout.to(DEV).reshape(B * T, E))
    _close(xd.grad.reshape(B, T, E), xr.grad, rtol=1e-4, atol=1e-5, what="conv dx")
    _close(layer.conv.weight.grad, conv.weight.grad, rtol=1e-3, atol=1e-5, what="conv dW")
    _close(layer.conv.bias.grad, conv.bias.grad, rtol=1e-3, atol=1e-5, what="conv db")
    # the reference's own [B,C,T] entry point
    out2 = layer(x.to(DEV).transpose(1, 2))
    _close(out2.transpose(1, 2), ref, rtol=1e-4, atol=1e-5)


def test_dropout_keep_rate_and_scaling(L):
    torch.manual_seed(0)
    layer = L.DilatedTemporalConvLayer(64, 64, 3, dilation=1, dropout=0.25).to(DEV).train()
    x = torch.randn(64 * 16, 64, device=DEV)
    y_eval = layer.eval().forward_tokens(x, 16) - x          # relu(conv)
    y_train = layer.train().forward_tokens(x, 16) - x        # dropped / scaled
    pos = y_eval > 1e-3
    kept = (y_train.abs() > 0) & pos
    rate = kept.sum().item() / pos.sum().item()
    assert abs(rate - 0.75) < 0.01, rate
    ratio = (y_train[kept] / y_eval[kept])
    assert torch.allclose(ratio, torch.full_like(ratio, 1 / 0.75), rtol=1e-3)


def test_layernorm_and_normalize_backward(L):
    torch.manual_seed(1)
    x = torch.randn(300, 96)
    ln = torch.nn.LayerNorm(96)
    with torch.no_grad():
        ln.weight.uniform_(0.5, 1.5), ln.bias.uniform_(-0.3, 0.3)
    xr = x.clone().requires_grad_(True)
    yr = ln(xr)
    g = torch.randn_like(yr)
    yr.backward(g)
    from lipvq import hvq
    xd = x.to(DEV).requires_grad_(True)
    w, b = ln.weight.detach().to(DEV).requires_grad_(True), ln.bias.detach().to(DEV).requires_grad_(True)
    y = hvq._LayerNormFn.apply(xd, w, b, ln.eps)
    _close(y, yr, atol=1e-5)
    y.backward(g.to(DEV))
    _close(xd.grad, xr.grad, atol=1e-5), _close(w.grad, ln.weight.grad, atol=1e-4), _close(b.grad, ln.bias.grad, atol=1e-4)
    xr2 = x.clone().requires_grad_(True)
    nr = torch.nn.functional.normalize(xr2, dim=-1)
    nr.backward(g)
    xd2 = x.to(DEV).requires_grad_(True)
    n = hvq.normalize(xd2)
    _close(n, nr, atol=1e-6)
    n.backward(g.to(DEV))
    _close(xd2.grad, xr2.grad, atol=1e-6)


# ------------------------------------------------------------------------------------------------ T2
def _t2_model(L, g):
    m = L.HierarchicalVQVAE(7, embed_dim=64, num_subclusters=64, num_clusters=32, num_stages=2, num_layers_per_stage=3,
                            dropout=0.0, replace_threshold_z=0.0, replace_threshold_q=0.0, kmeans_init=False)
    m.load_state_dict(sub(g, "sd"), strict=True)
    return m.to(DEV)


def test_t2_hierarchical_vs_reference_fixture(L):
    g = load_golden("t2_hvq")
    m = _t2_model(L, g)
    actions = g["actions"].to(DEV)
    out = m(actions, training=True)
    ref = sub(g, "out")
    assert set(out.keys()) == set(ref.keys())
    assert torch.equal(out["z_indices"].cpu(), ref["z_indices"]) and torch.equal(out["q_indices"].cpu(), ref["q_indices"])
    for k in ("embeddings", "quantized_z", "quantized_q", "reconstructed_actions"):
        _close(out[k], ref[k], rtol=1e-4, atol=1e-5, what=k)
    for k in ("z_distances", "q_distances"):  # lazily produced, from the pre-update codebooks
        _close(out[k], ref[k], rtol=1e-4, atol=1e-5, what=k)
    losses = m.compute_vqvae_loss(out, actions, lambda_rec=0.002)
    for k, v in sub(g, "loss").items():
        _close(losses[k], v, what=k)
    losses["vqvae_loss"].backward()
    params = dict(m.named_parameters())
    for k, gr in sub(g, "grad").items():
        _close(params[k].grad, gr, rtol=2e-3, atol=1e-7, what="grad " + k)
    sd = m.state_dict()
    for k, v in sub(g, "post").items():
        _close(sd[k], v, rtol=1e-4, atol=1e-6, what="post " + k)
    u = m.get_codebook_usage()
    assert u["z_used"] == int(g["usage_z_used"]) and u["q_used"] == int(g["usage_q_used"])
    assert set(u) == {"z_used", "z_total", "z_usage_pct", "z_utilization", "q_utilization", "q_used", "q_total",
                      "q_usage_pct", "z_dead", "q_dead"}


def test_t2_eval_does_not_touch_buffers_and_wide_embed(L):
    import lipvq_oracle as O
    torch.manual_seed(11)
    p = O.t2_init(12, embed_dim=128, num_subclusters=96, num_clusters=40, num_stages=1, num_layers_per_stage=5)
    m = L.HierarchicalVQVAE(12, embed_dim=128, num_subclusters=96, num_clusters=40, num_stages=1, num_layers_per_stage=5,
                            dropout=0.1, kmeans_init=False)
    sd = dict(p)
    sd["codebooks_initialized"] = torch.tensor(True)
    m.load_state_dict(sd, strict=True)
    m.to(DEV).eval()
    actions = torch.rand(32, 10, 12) * 2 - 1
    before = {k: v.clone() for k, v in m.state_dict().items()}
    out = m(actions.to(DEV), training=False)
    for k, v in m.state_dict().items():
        assert torch.equal(v, before[k]), k
    ref = O.t2_forward({k: v.clone() for k, v in p.items()}, actions, training=False, num_stages=1, num_layers=5)
    mism = int((out["z_indices"].cpu() != ref["z_indices"]).sum())
    assert mism <= 1  # cosine near-ties within the reference matmul's own rounding
    if mism == 0:
        _close(out["reconstructed_actions"], ref["reconstructed_actions"], rtol=1e-3, atol=1e-5)
    _close(out["embeddings"], ref["embeddings"], rtol=1e-4, atol=1e-5)


def test_t2_dead_code_replacement_masks(L):
    g = load_golden("t2_hvq")
    m = L.HierarchicalVQVAE(7, embed_dim=64, num_subclusters=64, num_clusters=32, num_stages=2, num_layers_per_stage=3,
                            dropout=0.0, kmeans_init=False)  # thresholds 1.0: codes with EMA size < 1 are replaced
    m.load_state_dict(sub(g, "sd"), strict=True)
    m.to(DEV)
    torch.manual_seed(0)
    m(g["actions"].to(DEV), training=True)
    # after one step from zero sizes every code is below 1.0 -> all replaced by unit-norm batch rows, size := 1
    assert torch.all(m.cluster_size_z == 1.0) and torch.all(m.cluster_size_q == 1.0)
    assert torch.allclose(m.codebook_z.norm(dim=-1), torch.ones(64, device=DEV), atol=1e-5)
    assert torch.equal(m.codebook_z, m.embed_avg_z)


# ------------------------------------------------------------------------------------------------ T3
def test_t3_kmeans_quantizer_vs_reference_fixture(L):
    g = load_golden("quantizers")
    pre, post = sub(g, "lap_pre"), sub(g, "lap_post")
    q = L.LFQQuantizerEMA_KMeans(64, 32, dead_threshold=0)
    q.load_state_dict(pre, strict=True)
    q.initialized = True
    q.to(DEV)
    z_q, idx = q(g["z"].to(DEV))
    assert torch.equal(idx.cpu(), g["lap_indices"]) and torch.equal(z_q.cpu(), g["lap_z_q"])
    sd = q.state_dict()
    for k in ("ema_cluster_size", "ema_codebook", "usage_counts", "usage_ma", "entropy_ma"):
        _close(sd[k], post[k], rtol=1e-5, atol=1e-6, what=k)
    _close(sd["codebook"], post["codebook"], rtol=1e-4, atol=1e-4, what="codebook")


def test_t3_dead_code_nearest_replacement(L):
    torch.manual_seed(2)
    q = L.LFQQuantizerEMA_KMeans(16, 8, dead_threshold=3, replace_strategy="nearest")
    q.initialized = True
    q.to(DEV)
    with torch.no_grad():
        q.usage_counts.copy_(torch.tensor([9.0] * 8 + [0.0] * 8, device=DEV))
    z = torch.rand(64, 8, device=DEV)
    z_q, idx = q(z)
    cb = q.codebook.data.cpu()
    dead = (q.usage_counts < 3).cpu()
    alive = ~dead
    # every dead row now equals one of the alive rows, namely its nearest one under the post-EMA codebook
    for i in dead.nonzero().flatten().tolist():
        assert any(torch.equal(cb[i], cb[j]) for j in alive.nonzero().flatten().tolist())


def test_t3_hierarchical_lipschitz_vs_reference_fixture(L):
    """HierarchicalLFQHVQVAE forward/backward against the REFERENCE's own run of hq_vae.py:289-415
    (tests/golden/t3_hier.npz, written by oracle/gen_golden.py: same weights, k-means skipped, dead-code replacement
    disabled): 11-key dict, indices and gathered codewords bit-exact, every parameter gradient, and the EMA / usage
    buffers + overwritten codebooks of both quantizers after the step."""
    g = load_golden("t3_hier")
    m = L.HierarchicalLFQHVQVAE(7, 32, 16, num_z_codes=64, num_q_codes=32)
    for qz in (m.z_quantizer, m.q_quantizer):
        qz.initialized, qz.dead_threshold = True, 0
    m.load_state_dict(sub(g, "sd"), strict=True)
    m.to(DEV)
    out = m(g["x"].to(DEV))
    out["loss"].backward()
    ref = sub(g, "out")
    assert list(out.keys()) == ["loss", "recon_loss", "z_commit", "z_codebook", "q_commit", "q_codebook", "x_recon", "z_q",
                                "q_q", "z_indices", "q_indices"]
    assert torch.equal(out["z_indices"].cpu(), ref["z_indices"]) and torch.equal(out["q_indices"].cpu(), ref["q_indices"])
    assert torch.equal(out["z_q"].detach().cpu(), ref["z_q"]) and torch.equal(out["q_q"].detach().cpu(), ref["q_q"])
    for k in ("loss", "recon_loss", "z_commit", "z_codebook", "q_commit", "q_codebook"):
        _close(out[k], ref[k], rtol=1e-4, atol=1e-7, what=k)
    _close(out["x_recon"], ref["x_recon"], rtol=1e-4, atol=1e-5, what="x_recon")
    grads = sub(g, "grad")
    named = dict(m.named_parameters())
    assert set(grads) == {k for k, v in named.items() if v.grad is not None}
    for k, gr in grads.items():
        _close(named[k].grad, gr, rtol=2e-3, atol=1e-7, what="grad " + k)
    post, sd = sub(g, "post"), m.state_dict()
    for pref in ("z_quantizer", "q_quantizer"):
        for k in ("ema_cluster_size", "ema_codebook", "usage_counts", "usage_ma", "entropy_ma"):
            # north_star gate: EMA buffers within 1e-4 relative after one step (q-level inputs passed through the GPU's
            # own Lipschitz layer, so they carry its ~1e-6 fp32 noise)
            _close(sd[f"{pref}.{k}"], post[f"{pref}.{k}"], rtol=1e-4, atol=1e-5, what=f"{pref}.{k}")
        _close(sd[f"{pref}.codebook"], post[f"{pref}.codebook"], rtol=1e-3, atol=1e-3, what=f"{pref}.codebook")


def _match_rows(a, b):
    """permutation p with a[p] ~ b (nearest-row matching; centres are well separated)"""
    return torch.cdist(b, a).argmin(dim=1)


def test_kmeans_init_vs_reference_fixture(L):
    """a13: HierarchicalVQVAE.initialize_codebooks_with_current_batch with kmeans_init=True against the reference's own
    run (heirarchical_vqvae.py:321-419; sklearn random_state=42) -- centres, cluster sizes and embedding sums; and
    LFQQuantizerEMA_KMeans.kmeans_init (hq_vae.py:65-78) on the blob latents of the same fixture."""
    import numpy as np
    g = load_golden("kmeans_init")
    t2 = sub(g, "t2")
    torch.manual_seed(0)
    m = L.HierarchicalVQVAE(7, embed_dim=16, num_subclusters=16, num_clusters=8, num_stages=1, num_layers_per_stage=1,
                            dropout=0.0, kmeans_init=True).to(DEV)
    assert not bool(m.codebooks_initialized)
    np.random.seed(0)
    m.initialize_codebooks_with_current_batch(g["emb"].to(DEV))
    assert bool(m.codebooks_initialized)
    sd = m.state_dict()
    for lvl in ("z", "q"):
        cb, ref_cb = sd["codebook_" + lvl].cpu(), t2["codebook_" + lvl]
        perm = _match_rows(cb, ref_cb)  # same sklearn seed -> identity in practice; matching makes the test order-free
        assert sorted(perm.tolist()) == list(range(cb.shape[0]))
        _close(cb[perm], ref_cb, rtol=1e-4, atol=1e-5, what="codebook_" + lvl)
        assert torch.equal(sd["cluster_size_" + lvl].cpu()[perm], t2["cluster_size_" + lvl])
        _close(sd["embed_avg_" + lvl].cpu()[perm], t2["embed_avg_" + lvl], rtol=1e-4, atol=1e-4, what="embed_avg_" + lvl)
    # second call is a no-op (heirarchical_vqvae.py:327-328)
    before = sd["codebook_z"].clone()
    m.initialize_codebooks_with_current_batch(torch.randn(2, 4, 16, device=DEV))
    assert torch.equal(m.codebook_z, before)
    # hq_vae.py kmeans_init: first forward of an un-initialised quantizer
    torch.manual_seed(1)
    q = L.LFQQuantizerEMA_KMeans(8, 8, dead_threshold=0).to(DEV)
    torch.manual_seed(5)
    np.random.seed(5)
    assert not q.initialized
    q.kmeans_init(g["z"].to(DEV))
    assert q.initialized and torch.equal(q.codebook.data, q.ema_codebook.data)
    cb = q.codebook.data.cpu()
    perm = _match_rows(cb, g["hq_codebook"])
    assert sorted(perm.tolist()) == list(range(8))
    _close(cb[perm], g["hq_codebook"], rtol=1e-4, atol=1e-5, what="hq kmeans centres")
    z_q, idx = q(g["z"].to(DEV))  # forward after init: every blob owns exactly one code
    assert torch.equal(torch.bincount(idx.cpu(), minlength=8), torch.full((8,), 32))


def test_hvq_trainer_graph_step_equals_eager_step():
    """lipvq.HvqTrainer: the captured step (one CUDA graph) against the same step run eagerly, from the same initial state on
    the same batches, with the dead-code thresholds at zero (no random replacement: the two runs are then deterministic up to
    the order of the atomic adds of the per-code sums)."""
    import copy
    import lipvq as L
    torch.manual_seed(0)
    cfg = dict(embed_dim=128, num_subclusters=64, num_clusters=32, num_stages=2, num_layers_per_stage=2, dropout=0.0,
               kmeans_init=False, replace_threshold_z=0.0, replace_threshold_q=0.0)
    m1 = L.HierarchicalVQVAE(7, **cfg).to(DEV).train()
    xs = [torch.rand(8, 10, 7, device=DEV) * 2 - 1 for _ in range(6)]
    with torch.no_grad():
        m1.initialize_codebooks_with_current_batch(m1.encoder(xs[0]))
    m2 = copy.deepcopy(m1)
    t1 = L.HvqTrainer(m1, lr=1e-3, weight_decay=1e-4, lambda_rec=0.5, graph=True, warmup=3)
    t2 = L.HvqTrainer(m2, lr=1e-3, weight_decay=1e-4, lambda_rec=0.5, graph=False)
    # the capture's warm-up steps are real steps on the first batch: give the eager trainer the same three
    l1 = t1.step(xs[0])
    for _ in range(3):
        t2.step(xs[0])
    l2 = t2.step(xs[0])
    assert torch.allclose(l1["vqvae_loss"], l2["vqvae_loss"], rtol=1e-4, atol=1e-7), (l1["vqvae_loss"], l2["vqvae_loss"])
    for x in xs[1:]:
        l1, l2 = t1.step(x), t2.step(x)
        assert torch.allclose(l1["vqvae_loss"], l2["vqvae_loss"], rtol=2e-3, atol=1e-6), (l1["vqvae_loss"], l2["vqvae_loss"])
    for (n, a), (_, b) in zip(m1.named_parameters(), m2.named_parameters()):
        assert torch.allclose(a, b, rtol=1e-2, atol=2e-4), n
    assert torch.allclose(m1.cluster_size_z, m2.cluster_size_z, rtol=1e-3, atol=1e-3)


def test_replace_dead_codes_static_replaces_exactly_the_dead_codes():
    import lipvq as L
    torch.manual_seed(1)
    m = L.HierarchicalVQVAE(7, embed_dim=64, num_subclusters=32, num_clusters=16, num_stages=1, num_layers_per_stage=1,
                            kmeans_init=False).to(DEV)
    emb = torch.randn(4, 10, 64, device=DEV)
    m.cluster_size_z.copy_(torch.arange(32, device=DEV).float() % 2)       # every second code dead
    m.cluster_size_q.fill_(5.0)
    cz, cq = m.codebook_z.clone(), m.codebook_q.clone()
    m.replace_dead_codes_static(emb, emb)
    dead = torch.arange(32, device=DEV) % 2 == 0
    assert torch.equal(m.codebook_z[~dead], cz[~dead]) and torch.equal(m.codebook_q, cq)
    rows = torch.nn.functional.normalize(emb.reshape(-1, 64), dim=-1)
    hit = (m.codebook_z[dead] @ rows.T).max(dim=1).values                    # each replacement is a (normalised) batch row
    assert torch.all(hit > 1 - 1e-5)
    assert torch.all(m.cluster_size_z[dead] == 1.0) and torch.equal(m.embed_avg_z[dead], m.codebook_z[dead])


def test_t2_shipped_width_vs_oracle(L):
    """HierarchicalVQVAE at the fork's shipped config (config/icl_config.py:130-135: embed_dim 512, 2 x 4 MS-TCN layers per
    side, 1024 / 512 codes) against the oracle port, forward in eval mode: the E = 512 convolutions run on the tensor cores
    (conv_tc.cu, 160 tokens >= its gate), the searches on the streamed tcgen05 pass + exact scan.  Embeddings within 2e-4
    (eight layers of three-term bf16 contractions against fp32 torch); indices equal except cosine near-ties inside the
    reference matmul's own rounding, which at random initialisation (gaps of 1e-3 between codes) means equal."""
    import lipvq_oracle as O
    torch.manual_seed(5)
    cfg = dict(embed_dim=512, num_subclusters=1024, num_clusters=512, num_stages=2, num_layers_per_stage=4)
    p = O.t2_init(12, **cfg)
    m = L.HierarchicalVQVAE(12, dropout=0.0, kmeans_init=False, **cfg)
    sd = dict(p)
    sd["codebooks_initialized"] = torch.tensor(True)
    m.load_state_dict(sd, strict=True)
    m.to(DEV).eval()
    actions = torch.rand(16, 10, 12) * 2 - 1
    out = m(actions.to(DEV), training=False)
    ref = O.t2_forward({k: v.clone() for k, v in p.items()}, actions, training=False, num_stages=2, num_layers=4)
    _close(out["embeddings"], ref["embeddings"], rtol=2e-4, atol=2e-5)
    mz = int((out["z_indices"].cpu() != ref["z_indices"]).sum())
    mq = int((out["q_indices"].cpu() != ref["q_indices"]).sum())
    assert mz <= 2 and mq <= 2, (mz, mq)
    if mz == 0 and mq == 0:
        _close(out["reconstructed_actions"], ref["reconstructed_actions"], rtol=1e-3, atol=2e-5)
