"""Pins oracle/lipvq_oracle.py to fixtures produced by the reference's own modules
(oracle/gen_golden.py).  CPU only."""
import torch

import lipvq_oracle as O
from conftest import load_golden, sub

RT = dict(rtol=1e-6, atol=1e-7)  # same-host runs are bit-equal; slack is for other BLAS builds


def _close(a, b, **kw):
    kw = kw or RT
    assert torch.allclose(a, b, **kw), f"max abs diff {(a - b).abs().max().item():.3e}"


def test_init_draw_order_matches_reference_ctor():
    g = load_golden("t1_c1")
    torch.manual_seed(0)
    p = O.t1_init(7, 64, 512)
    sd = sub(g, "sd")
    assert set(p) == set(sd)
    for k in sd:
        assert torch.equal(p[k], sd[k]), k
    g4 = load_golden("t4_classic")
    torch.manual_seed(0)
    p4 = O.t4_init(7, 32, 128)
    for k, v in sub(g4, "sd").items():
        assert torch.equal(p4[k], v), k
    g2 = load_golden("t2_hvq")
    torch.manual_seed(0)
    p2 = O.t2_init(7, 64, 64, 32, 2, 3)
    sd2 = sub(g2, "sd")
    for k, v in sd2.items():
        if k == "codebooks_initialized":
            continue
        assert torch.equal(p2[k], v), k


def _t1_case(name):
    g = load_golden(name)
    p = {k: v.clone().requires_grad_(True) for k, v in sub(g, "sd").items()}
    out = O.t1_forward(p, g["x"])
    assert torch.equal(out["indices"], g["indices"])
    assert torch.equal(out["z_latent"], g["z_latent"])
    _close(out["z_e"], g["z_e"])
    _close(out["loss"], g["loss"])
    out["loss"].backward()
    for k, gr in sub(g, "grad").items():
        _close(p[k].grad, gr, rtol=1e-5, atol=1e-8)
    return g


def test_t1_forward_backward_c1():
    _t1_case("t1_c1")


def test_t1_forward_backward_reference_shape():
    _t1_case("t1_real")


def test_t1_adamw_trajectory():
    for name in ("t1_c1", "t1_real"):
        g = load_golden(name)
        p = {k: v.clone() for k, v in sub(g, "sd").items()}
        st = {}
        losses = [O.t1_train_step(p, g["x"], st)["loss"] for _ in range(g["traj_losses"].numel())]
        _close(torch.stack(losses), g["traj_losses"], rtol=1e-5, atol=1e-7)
        for k, v in sub(g, "final").items():
            _close(p[k], v, rtol=1e-4, atol=1e-6)


def test_t1_ema_trajectory():
    g = load_golden("t1_ema")
    sd = sub(g, "sd")
    ema = {"ema_cluster_size": sd.pop("quantizer.ema_cluster_size").clone(),
           "ema_codebook": sd.pop("quantizer.ema_codebook").clone()}
    p = {k: v.clone() for k, v in sd.items()}
    st = {}
    losses = [O.t1_train_step(p, g["x"], st, ema=ema)["loss"] for _ in range(g["traj_losses"].numel())]
    _close(torch.stack(losses), g["traj_losses"], rtol=1e-5, atol=1e-7)
    fin = sub(g, "final")
    _close(ema["ema_cluster_size"], fin["quantizer.ema_cluster_size"], rtol=1e-5, atol=1e-7)
    _close(ema["ema_codebook"], fin["quantizer.ema_codebook"], rtol=1e-4, atol=1e-6)
    _close(p["quantizer.codebook"], fin["quantizer.codebook"], rtol=1e-4, atol=1e-6)


def test_quantizers_ema_simple_and_laplace():
    g = load_golden("quantizers")
    pre, post = sub(g, "pre"), sub(g, "post")
    idx = O.nearest_code_l2(g["z"], pre["codebook"])
    assert torch.equal(idx, g["indices"])
    assert torch.equal(pre["codebook"][idx], g["z_q"])
    size, esum, cb = O.ema_update_simple(g["z"], idx, pre["ema_cluster_size"], pre["ema_codebook"])
    _close(size, post["ema_cluster_size"]), _close(esum, post["ema_codebook"]), _close(cb, post["codebook"])
    pre, post = sub(g, "lap_pre"), sub(g, "lap_post")
    idx = O.nearest_code_l2(g["z"], pre["codebook"])
    assert torch.equal(idx, g["lap_indices"])
    size, esum, cb = O.ema_update_laplace(g["z"], idx, pre["ema_cluster_size"], pre["ema_codebook"])
    _close(size, post["ema_cluster_size"]), _close(esum, post["ema_codebook"])
    _close(cb, post["codebook"], rtol=1e-5, atol=1e-6)
    counts = torch.bincount(idx, minlength=size.numel()).float()
    uc, uma, ent = O.usage_stats(counts, pre["usage_counts"], pre["usage_ma"], pre["entropy_ma"])
    _close(uc, post["usage_counts"]), _close(uma, post["usage_ma"]), _close(ent, post["entropy_ma"])


def test_t4_classic_ste():
    g = load_golden("t4_classic")
    p = {k: v.clone().requires_grad_(True) for k, v in sub(g, "sd").items()}
    out = O.t4_forward(p, g["x"])
    assert torch.equal(out["indices"], g["indices"])
    _close(out["z_latent"], g["z_latent"])
    _close(out["loss"], g["loss"])
    out["loss"].backward()
    for k, gr in sub(g, "grad").items():
        _close(p[k].grad, gr, rtol=1e-5, atol=1e-8)


def test_t2_hierarchical():
    g = load_golden("t2_hvq")
    sd = sub(g, "sd")
    sd.pop("codebooks_initialized")
    p = {k: v.clone() for k, v in sd.items()}
    for k in p:
        if k not in O.T2_BUFFERS:
            p[k].requires_grad_(True)
    out = O.t2_forward(p, g["actions"], training=True, num_stages=2, num_layers=3)
    ref = sub(g, "out")
    assert torch.equal(out["z_indices"], ref["z_indices"])
    assert torch.equal(out["q_indices"], ref["q_indices"])
    for k in ("embeddings", "quantized_z", "quantized_q", "reconstructed_actions", "z_distances", "q_distances"):
        _close(out[k], ref[k])
    losses = O.t2_loss(out, g["actions"], lambda_rec=0.002)
    for k, v in sub(g, "loss").items():
        _close(losses[k], v)
    losses["vqvae_loss"].backward()
    for k, gr in sub(g, "grad").items():
        _close(p[k].grad, gr, rtol=1e-5, atol=1e-8)
    for k, v in sub(g, "post").items():
        _close(p[k], v)
    u = O.codebook_usage(p["cluster_size_z"], p["cluster_size_q"], 0.0, 0.0)
    assert u["z_used"] == int(g["usage_z_used"]) and u["q_used"] == int(g["usage_q_used"])


def test_row_chunked_search_is_bit_identical():
    torch.manual_seed(5)
    z, cb = torch.rand(300, 48), torch.rand(257, 48) - 0.5
    full = O.nearest_code_l2(z, cb, row_chunk=300)
    for c in (1, 7, 64):
        assert torch.equal(O.nearest_code_l2(z, cb, row_chunk=c), full)


def test_t3_hierarchical_lipschitz():
    """oracle t3_forward vs the reference's HierarchicalLFQHVQVAE run (tests/golden/t3_hier.npz)."""
    g = load_golden("t3_hier")
    sd = sub(g, "sd")
    frozen = ("ema_", "usage_", "entropy_")
    p = {k: v.clone().requires_grad_(v.dtype.is_floating_point and not any(f in k for f in frozen)) for k, v in sd.items()}
    out = O.t3_forward(p, g["x"])
    ref = sub(g, "out")
    assert torch.equal(out["z_indices"], ref["z_indices"]) and torch.equal(out["q_indices"], ref["q_indices"])
    assert torch.equal(out["z_q"].detach(), ref["z_q"]) and torch.equal(out["q_q"].detach(), ref["q_q"])
    for k in ("loss", "recon_loss", "z_commit", "z_codebook", "q_commit", "q_codebook", "x_recon"):
        _close(out[k], ref[k])
    out["loss"].backward()
    for k, gr in sub(g, "grad").items():
        _close(p[k].grad, gr, rtol=1e-5, atol=1e-8)
    post = sub(g, "post")
    for k, v in out["post"].items():
        _close(v, post[k], rtol=1e-5, atol=1e-6)


def test_kmeans_initialisers():
    """oracle hvq_kmeans_init vs HierarchicalVQVAE.initialize_codebooks_with_current_batch run by the reference
    (sklearn random_state=42: deterministic), and the blob centres LFQQuantizerEMA_KMeans.kmeans_init must find."""
    g = load_golden("kmeans_init")
    t2 = sub(g, "t2")
    out = O.hvq_kmeans_init(g["emb"], 16, 8)
    for k, v in out.items():
        _close(v, t2[k], rtol=1e-5, atol=1e-5)
    assert bool(t2["codebooks_initialized"])
    assert t2["cluster_size_z"].sum().item() == g["emb"].shape[0] * g["emb"].shape[1]
    # hq: 8 blobs of 32 rows -> the 8 centres are the blob means, in some order
    cb = g["hq_codebook"]
    assert torch.equal(cb, g["hq_ema_codebook"])
    d = torch.cdist(g["z"], cb)
    assert torch.equal(torch.bincount(d.argmin(1), minlength=8), torch.full((8,), 32))
