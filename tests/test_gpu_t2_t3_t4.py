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


def test_t3_hierarchical_lipschitz_vs_torch_autograd(L):
    """HierarchicalLFQHVQVAE forward/backward against a torch-CPU restatement of hq_vae.py:358-415 built from the
    oracle's functions (same weights, k-means skipped, dead-code replacement disabled)."""
    import lipvq_oracle as O
    F = torch.nn.functional
    torch.manual_seed(4)
    m = L.HierarchicalLFQHVQVAE(7, 32, 16, num_z_codes=64, num_q_codes=32)
    for qz in (m.z_quantizer, m.q_quantizer):
        qz.initialized, qz.dead_threshold = True, 0
    with torch.no_grad():
        m.z_quantizer.codebook.copy_(torch.rand(64, 32))
        m.q_quantizer.codebook.copy_(torch.rand(32, 16))
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    x = torch.rand(200, 7) * 2 - 1
    m.to(DEV)
    out = m(x.to(DEV))
    out["loss"].backward()
    p = {k: v.clone().requires_grad_(v.dtype.is_floating_point and "ema" not in k and "usage" not in k and "entropy" not in k)
         for k, v in sd.items()}
    h = F.gelu(F.linear(F.gelu(F.linear(x, p["encoder.0.weight"], p["encoder.0.bias"])), p["encoder.2.weight"],
                        p["encoder.2.bias"]))
    z_e = O.lipschitz_sigmoid(h, p["to_z_latent.W"], p["to_z_latent.b"], p["to_z_latent.ci"])
    zi = O.nearest_code_l2(z_e.detach(), p["z_quantizer.codebook"].detach())
    z_q = p["z_quantizer.codebook"][zi]
    q_e = O.lipschitz_sigmoid(z_q.detach(), p["q_encoder.W"], p["q_encoder.b"], p["q_encoder.ci"])
    qi = O.nearest_code_l2(q_e.detach(), p["q_quantizer.codebook"].detach())
    q_q = p["q_quantizer.codebook"][qi]
    d = F.gelu(F.linear(F.gelu(F.linear(q_q, p["decoder.0.weight"], p["decoder.0.bias"])), p["decoder.2.weight"],
                        p["decoder.2.bias"]))
    xr = F.linear(d, p["to_output.weight"], p["to_output.bias"])
    loss = (F.mse_loss(xr, x) + 0.25 * (F.mse_loss(z_e, z_q.detach()) + F.mse_loss(z_q, z_e.detach()))
            + 0.25 * (F.mse_loss(q_e, q_q.detach()) + F.mse_loss(q_q, q_e.detach())))
    loss.backward()
    assert torch.equal(out["z_indices"].cpu(), zi) and torch.equal(out["q_indices"].cpu(), qi)
    assert torch.equal(out["z_q"].cpu(), z_q.detach()) and torch.equal(out["q_q"].cpu(), q_q.detach())
    _close(out["loss"], loss), _close(out["x_recon"], xr, atol=1e-5)
    assert set(out) == {"loss", "recon_loss", "z_commit", "z_codebook", "q_commit", "q_codebook", "x_recon", "z_q", "q_q",
                        "z_indices", "q_indices"}
    for k, v in m.named_parameters():
        if p[k].grad is not None:
            _close(v.grad, p[k].grad, rtol=2e-3, atol=1e-7, what="grad " + k)
    # EMA side effects of both quantizers (Laplace variant) against the oracle
    s, e, c = O.ema_update_laplace(z_e.detach(), zi, sd["z_quantizer.ema_cluster_size"], sd["z_quantizer.ema_codebook"])
    _close(m.z_quantizer.ema_cluster_size, s, rtol=1e-5), _close(m.z_quantizer.ema_codebook, e, rtol=1e-4, atol=1e-6)
    _close(m.z_quantizer.codebook, c, rtol=1e-3, atol=1e-3)
