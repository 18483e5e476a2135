"""CPU oracle for the LipVQ-VAE action-tokenizer hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``robot-manipulation-vq-vae_b200/`` may import this
file; it is used by ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` as the *checker* and the timed CPU baseline, never
as the product path.

What it is: a functional (stateless, dict-of-tensors) restatement, in plain PyTorch CPU ops,
of the arithmetic of the reference tokenizers.  The reference itself is PyTorch, so using the
same ATen ops in the same order reproduces it bit-for-bit on CPU.  Parity is PINNED: the
committed fixtures under ``tests/golden/`` were produced by importing the reference's own
modules from ``/root/reference`` (``oracle/gen_golden.py``) and ``tests/test_oracle_golden.py``
checks every function here against them (indices/codewords bit-exact, floats to 0 ulp on the
same host, <=1e-6 rel across hosts).

Reference anchors (paths relative to the reference checkout):
  * L-inf weight clamp ............ robomimic/models/vq_vae/backbone_lfqvae_v5.py:6-12
  * Lipschitz sigmoid layer ....... backbone_lfqvae_v5.py:15-24
  * nearest-code search (L2) ...... backbone_lfqvae_v5.py:37-48
  * T1 tokenizer fwd + losses ..... backbone_lfqvae_v5.py:51-84
  * EMA quantizer ................. robomimic/models/vq_vae/hq_vae.py:224-281
  * Laplace/k-means EMA quantizer . hq_vae.py:84-130; k-means init hq_vae.py:65-78 (sklearn, host side)
  * hierarchical Lipschitz VQ-VAE . hq_vae.py:289-415
  * HVQ k-means codebook init ..... robomimic/algo/heirarchical_vqvae.py:321-419 (sklearn + assignment statistics)
  * classic VQ-VAE (STE) .......... robomimic/models/vq_vae/backbone.py:38-76
  * MS-TCN encoder/decoder ........ robomimic/algo/heirarchical_vqvae.py:8-214
  * cosine quantize / EMA / loss .. heirarchical_vqvae.py:421-541, 590-685
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import torch
import torch.nn.functional as F

Params = Dict[str, torch.Tensor]


# --------------------------------------------------------------------------------------
# parameter construction (same RNG draw order as the reference constructors, so that
# torch.manual_seed(s) + these builders == torch.manual_seed(s) + reference ctor)
# --------------------------------------------------------------------------------------
def _linear_init(out_f: int, in_f: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """nn.Linear default init: kaiming_uniform(a=sqrt(5)) weight, U(-1/sqrt(in),1/sqrt(in)) bias."""
    w = torch.empty(out_f, in_f)
    torch.nn.init.kaiming_uniform_(w, a=math.sqrt(5))
    bound = 1.0 / math.sqrt(in_f) if in_f > 0 else 0.0
    b = torch.empty(out_f)
    torch.nn.init.uniform_(b, -bound, bound)
    return w, b


def t1_init(feature_dim: int, latent_dim: int, num_codes: int = 1024, hidden_dim: int = 128) -> Params:
    """Parameters of the T1 tokenizer with the reference's state_dict key names
    (backbone_lfqvae_v5.py:52-68): draw order encoder.0, encoder.2, to_latent.W,
    quantizer.codebook (randn then kaiming_uniform_), decoder.0, decoder.2, to_output."""
    p: Params = {}
    p["encoder.0.weight"], p["encoder.0.bias"] = _linear_init(64, feature_dim)
    p["encoder.2.weight"], p["encoder.2.bias"] = _linear_init(hidden_dim, 64)
    p["to_latent.W"] = torch.randn(latent_dim, hidden_dim)
    p["to_latent.b"] = torch.zeros(latent_dim)
    p["to_latent.ci"] = torch.ones(latent_dim)
    cb = torch.randn(num_codes, latent_dim)
    torch.nn.init.kaiming_uniform_(cb)
    p["quantizer.codebook"] = cb
    p["decoder.0.weight"], p["decoder.0.bias"] = _linear_init(64, latent_dim)
    p["decoder.2.weight"], p["decoder.2.bias"] = _linear_init(hidden_dim, 64)
    p["to_output.weight"], p["to_output.bias"] = _linear_init(feature_dim, hidden_dim)
    return p


# --------------------------------------------------------------------------------------
# building blocks
# --------------------------------------------------------------------------------------
def linf_row_scale(W: torch.Tensor, ci: torch.Tensor) -> torch.Tensor:
    """min(1, softplus(ci)/sum_j|W_ij|) per output row (backbone_lfqvae_v5.py:6-12)."""
    l1 = W.abs().sum(dim=1, keepdim=True)
    return torch.minimum(torch.tensor(1.0), F.softplus(ci).unsqueeze(1) / l1)


def lipschitz_sigmoid(h: torch.Tensor, W: torch.Tensor, b: torch.Tensor, ci: torch.Tensor) -> torch.Tensor:
    """sigmoid(h @ (W*scale)^T + b) (backbone_lfqvae_v5.py:22-24)."""
    Wn = W * linf_row_scale(W, ci)
    return torch.sigmoid(torch.matmul(h, Wn.T) + b)


def nearest_code_l2(z: torch.Tensor, codebook: torch.Tensor, *, squared: bool = False,
                    row_chunk: int | None = None) -> torch.Tensor:
    """argmin_k ||z_n - e_k||_2, ties -> lowest k (torch.argmin).

    ``squared=False`` follows backbone_lfqvae_v5.py:39-46 literally, including the sign mask
    (a numerical no-op, kept so the op sequence is the reference's) and the sqrt inside
    torch.norm.  ``squared=True`` follows backbone.py:57-63 (sum of squares, no sqrt).
    The reference materialises [N,K,D]; the per-(n,k) reduction runs over D only, so
    evaluating rows in chunks is bit-identical and makes large K feasible.
    """
    N, D = z.shape
    K = codebook.shape[0]
    if row_chunk is None:
        row_chunk = max(1, min(N, (1 << 26) // max(1, K * D)))
    out = torch.empty(N, dtype=torch.int64)
    for s in range(0, N, row_chunk):
        zc = z[s:s + row_chunk]
        if squared:
            d = (zc.unsqueeze(1) - codebook).pow(2).sum(-1)
        else:
            sgn = torch.clamp((2 * torch.sign(zc) + 1).unsqueeze(1), max=1)
            d = torch.norm(sgn * (zc.unsqueeze(1) - codebook.unsqueeze(0)), dim=-1)
        out[s:s + row_chunk] = torch.argmin(d, dim=-1)
    return out


def nearest_code_cosine(x: torch.Tensor, codebook: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """argmax_k <normalize(x_n), normalize(c_k)>; also returns the [N,K] similarities
    (heirarchical_vqvae.py:441-449)."""
    xn = F.normalize(x, dim=-1)
    cn = F.normalize(codebook, dim=-1)
    sims = torch.matmul(xn, cn.t())
    return torch.argmax(sims, dim=1), sims


def ema_update_simple(z: torch.Tensor, idx: torch.Tensor, ema_size: torch.Tensor, ema_sum: torch.Tensor,
                      decay: float = 0.99, eps: float = 1e-5) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """hq_vae.py:262-279: size<-g*size+(1-g)*counts; sum<-g*sum+(1-g)*onehot^T z;
    codebook<-sum/(size+eps).  Returns new (size, sum, codebook); inputs untouched."""
    K = ema_size.numel()
    enc = F.one_hot(idx, K).type(z.dtype)
    size = ema_size.clone().mul_(decay).add_(enc.sum(0), alpha=1 - decay)
    esum = ema_sum.clone().mul_(decay).add_(enc.t() @ z, alpha=1 - decay)
    n = size + eps
    return size, esum, esum / n.unsqueeze(1)


def ema_update_laplace(z: torch.Tensor, idx: torch.Tensor, ema_size: torch.Tensor, ema_sum: torch.Tensor,
                       decay: float = 0.99, eps: float = 1e-5) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """hq_vae.py:103-118: same EMA, codebook<-sum / ((size+eps)/(sum(size)+K*eps))."""
    K = ema_size.numel()
    one_hot = F.one_hot(idx, K).float()
    size = ema_size.clone().mul_(decay).add_(one_hot.sum(0), alpha=1 - decay)
    esum = ema_sum.clone().mul_(decay).add_(one_hot.T @ z, alpha=(1 - decay))
    n = size.sum()
    size_norm = (size + eps) / (n + K * eps)
    return size, esum, esum / size_norm.unsqueeze(1)


def ema_update_cosine(x: torch.Tensor, idx: torch.Tensor, size_buf: torch.Tensor, avg_buf: torch.Tensor,
                      decay: float = 0.99, eps: float = 1e-5, normalize: bool = True
                      ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """heirarchical_vqvae.py:498-541: EMA, proto=avg/(size+eps), optional L2 normalise."""
    K = size_buf.numel()
    onehot = F.one_hot(idx, num_classes=K).type_as(x)
    size = size_buf.clone().mul_(decay).add_(onehot.sum(dim=0), alpha=(1.0 - decay))
    avg = avg_buf.clone().mul_(decay).add_(onehot.t() @ x, alpha=(1.0 - decay))
    proto = avg / (size + eps).unsqueeze(1)
    if normalize:
        proto = F.normalize(proto, dim=-1)
    return size, avg, proto


def usage_stats(counts: torch.Tensor, usage_counts: torch.Tensor, usage_ma: torch.Tensor,
                entropy_ma: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """hq_vae.py:123-130 usage bookkeeping."""
    uc = usage_counts + counts
    uma = 0.99 * usage_ma + 0.01 * (counts > 0).float()
    p = counts / (counts.sum() + 1e-8)
    ent = -(p * (p + 1e-8).log()).sum()
    return uc, uma, 0.99 * entropy_ma + 0.01 * ent


# --------------------------------------------------------------------------------------
# T1: LipVQ-VAE tokenizer (backbone_lfqvae_v5.py:51-84), optional EMA quantizer (hq_vae.py:224-281)
# --------------------------------------------------------------------------------------
def t1_encode(p: Params, x: torch.Tensor) -> torch.Tensor:
    """encoder MLP (exact-erf GELU) + Lipschitz sigmoid layer -> z_e in (0,1)^D."""
    h = F.gelu(F.linear(x, p["encoder.0.weight"], p["encoder.0.bias"]))
    h = F.gelu(F.linear(h, p["encoder.2.weight"], p["encoder.2.bias"]))
    return lipschitz_sigmoid(h, p["to_latent.W"], p["to_latent.b"], p["to_latent.ci"])


def t1_decode(p: Params, z_q: torch.Tensor) -> torch.Tensor:
    r = F.gelu(F.linear(z_q, p["decoder.0.weight"], p["decoder.0.bias"]))
    r = F.gelu(F.linear(r, p["decoder.2.weight"], p["decoder.2.bias"]))
    return F.linear(r, p["to_output.weight"], p["to_output.bias"])


def t1_forward(p: Params, x: torch.Tensor, *, row_chunk: int | None = None) -> Dict[str, torch.Tensor]:
    """Full T1 forward.  No straight-through: z_q = codebook[idx] carries grad to the
    codebook only; the encoder learns through the commitment term."""
    z_e = t1_encode(p, x)
    with torch.no_grad():
        idx = nearest_code_l2(z_e.detach(), p["quantizer.codebook"].detach(), row_chunk=row_chunk)
    z_q = p["quantizer.codebook"][idx]
    z_latent = z_q.clone().detach()
    x_recon = t1_decode(p, z_q)
    recon = F.mse_loss(x_recon, x)
    commit = F.mse_loss(z_q.detach(), z_e)
    cbl = F.mse_loss(z_q, z_e.detach())
    loss = recon + 0.25 * commit + 0.25 * cbl
    return {"z_e": z_e, "indices": idx, "z_q": z_q, "z_latent": z_latent, "x_recon": x_recon,
            "recon_loss": recon, "commitment_loss": commit, "codebook_loss": cbl, "loss": loss}


def t1_tokenize(p: Params, x: torch.Tensor, *, row_chunk: int | None = None) -> torch.Tensor:
    """encode + VQ only (the headline metric's path): x -> code indices."""
    with torch.no_grad():
        return nearest_code_l2(t1_encode(p, x), p["quantizer.codebook"], row_chunk=row_chunk)


def adamw_step(params: Params, grads: Params, m: Params, v: Params, step: int, lr: float = 1e-3,
               betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 1e-4) -> None:
    """torch.optim.AdamW single-tensor update rule (decoupled decay), in place.
    Reference protocol: AdamW(lr=1e-3, weight_decay=1e-4) on the tokenizer, robomimic/algo/icl.py:1437-1441."""
    b1, b2 = betas
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    for k, w in params.items():
        g = grads.get(k)
        if g is None:
            continue
        w.mul_(1 - lr * weight_decay)
        m[k].lerp_(g, 1 - b1)
        v[k].mul_(b2).addcmul_(g, g, value=1 - b2)
        denom = (v[k].sqrt() / math.sqrt(bc2)).add_(eps)
        w.addcdiv_(m[k], denom, value=-(lr / bc1))


def t1_train_step(p: Params, x: torch.Tensor, opt_state: Dict, *, ema: Dict | None = None,
                  lr: float = 1e-3, weight_decay: float = 1e-4) -> Dict[str, torch.Tensor]:
    """One tokenizer optimisation step in the reference's order (icl.py:1466,1520-1522):
    zero_grad -> forward -> loss.backward() -> AdamW.step().  With ``ema`` (keys
    ema_cluster_size, ema_codebook, decay, epsilon) the quantizer is hq_vae.py:224-281: the
    codebook is overwritten from the EMA buffers inside forward (after the gather) and still
    receives its gradient/AdamW update afterwards, exactly as the reference does."""
    leaves = {k: v.detach().requires_grad_(True) for k, v in p.items()}
    out = t1_forward(leaves, x)
    if ema is not None:
        with torch.no_grad():
            size, esum, cb = ema_update_simple(out["z_e"].detach(), out["indices"], ema["ema_cluster_size"],
                                               ema["ema_codebook"], ema.get("decay", 0.99), ema.get("epsilon", 1e-5))
            ema["ema_cluster_size"].copy_(size)
            ema["ema_codebook"].copy_(esum)
    out["loss"].backward()
    grads = {k: v.grad for k, v in leaves.items()}
    with torch.no_grad():
        if ema is not None:
            p["quantizer.codebook"].copy_(cb)
        opt_state["step"] = opt_state.get("step", 0) + 1
        if "m" not in opt_state:
            opt_state["m"] = {k: torch.zeros_like(v) for k, v in p.items()}
            opt_state["v"] = {k: torch.zeros_like(v) for k, v in p.items()}
        adamw_step(p, grads, opt_state["m"], opt_state["v"], opt_state["step"], lr=lr, weight_decay=weight_decay)
    res = {k: v.detach() for k, v in out.items()}
    res["grads"] = grads
    return res


# --------------------------------------------------------------------------------------
# T3: HierarchicalLFQHVQVAE (hq_vae.py:289-415): T1 encoder -> Lipschitz -> Laplace-EMA quantizer Z ->
# Lipschitz q_encoder on z_q.detach() -> Laplace-EMA quantizer Q -> T1 decoder
# --------------------------------------------------------------------------------------
def t3_forward(p: Params, x: torch.Tensor) -> Dict[str, torch.Tensor]:
    """Functional restatement of HierarchicalLFQHVQVAE.forward with both quantizers already initialised and dead-code
    replacement off.  Returns the reference's 11 keys plus ``post``: the buffers / codebooks each quantizer holds after
    the call (EMA runs unconditionally, hq_vae.py:103-130)."""
    h = F.gelu(F.linear(F.gelu(F.linear(x, p["encoder.0.weight"], p["encoder.0.bias"])), p["encoder.2.weight"],
                        p["encoder.2.bias"]))
    z_e = lipschitz_sigmoid(h, p["to_z_latent.W"], p["to_z_latent.b"], p["to_z_latent.ci"])
    post: Params = {}

    def quant(pref: str, e: torch.Tensor):
        cb = p[pref + ".codebook"]
        idx = nearest_code_l2(e.detach(), cb.detach())
        q = cb[idx]
        with torch.no_grad():
            size, esum, new_cb = ema_update_laplace(e.detach(), idx, p[pref + ".ema_cluster_size"], p[pref + ".ema_codebook"])
            counts = F.one_hot(idx, cb.shape[0]).float().sum(0)
            uc, uma, ent = usage_stats(counts, p[pref + ".usage_counts"], p[pref + ".usage_ma"], p[pref + ".entropy_ma"])
        post.update({pref + ".ema_cluster_size": size, pref + ".ema_codebook": esum, pref + ".codebook": new_cb,
                     pref + ".usage_counts": uc, pref + ".usage_ma": uma, pref + ".entropy_ma": ent})
        return q, idx

    z_q, z_idx = quant("z_quantizer", z_e)
    commit_z, codebook_z = F.mse_loss(z_e, z_q.detach()), F.mse_loss(z_q, z_e.detach())
    q_e = lipschitz_sigmoid(z_q.detach(), p["q_encoder.W"], p["q_encoder.b"], p["q_encoder.ci"])
    q_q, q_idx = quant("q_quantizer", q_e)
    commit_q, codebook_q = F.mse_loss(q_e, q_q.detach()), F.mse_loss(q_q, q_e.detach())
    d = F.gelu(F.linear(F.gelu(F.linear(q_q, p["decoder.0.weight"], p["decoder.0.bias"])), p["decoder.2.weight"],
                        p["decoder.2.bias"]))
    x_recon = F.linear(d, p["to_output.weight"], p["to_output.bias"])
    recon = F.mse_loss(x_recon, x)
    loss = recon + 0.25 * (commit_z + codebook_z) + 0.25 * (commit_q + codebook_q)
    return {"loss": loss, "recon_loss": recon, "z_commit": commit_z, "z_codebook": codebook_z, "q_commit": commit_q,
            "q_codebook": codebook_q, "x_recon": x_recon, "z_q": z_q, "q_q": q_q, "z_indices": z_idx, "q_indices": q_idx,
            "post": post}


def hvq_assignment_stats(sample: torch.Tensor, codebook: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """heirarchical_vqvae.py:377-394 (cosine branch): counts = bincount(argmax(normalize(x) @ codebook^T)),
    sums[k] = sum of the un-normalised rows assigned to k."""
    assign = torch.argmax(torch.matmul(F.normalize(sample, dim=-1), codebook.t()), dim=1)
    K = codebook.shape[0]
    counts = torch.bincount(assign, minlength=K).float()
    sums = torch.zeros_like(codebook)
    for k in range(K):
        m = (assign == k).nonzero(as_tuple=False).squeeze(-1)
        if m.numel() > 0:
            sums[k] = sample[m].sum(dim=0)
    return counts, sums


def hvq_kmeans_init(embeddings: torch.Tensor, num_subclusters: int, num_clusters: int) -> Params:
    """heirarchical_vqvae.py:321-419 with kmeans_init=True, use_cosine=True (<= sample_limit rows, so no subsampling):
    sklearn KMeans(n_init=10, random_state=42, max_iter=300) for Z and Q, centres L2-normalised, EMA statistics from
    the assignment of the same rows."""
    from sklearn.cluster import KMeans
    flat = embeddings.reshape(-1, embeddings.shape[-1])
    sample = flat.cpu().numpy()
    out: Params = {}
    for tag, k in (("z", num_subclusters), ("q", num_clusters)):
        km = KMeans(n_clusters=k, n_init=10, random_state=42, max_iter=300).fit(sample)
        cb = F.normalize(torch.from_numpy(km.cluster_centers_).float(), dim=-1)
        counts, sums = hvq_assignment_stats(flat, cb)
        out.update({"codebook_" + tag: cb, "cluster_size_" + tag: counts, "embed_avg_" + tag: sums})
    return out


# --------------------------------------------------------------------------------------
# T4: classic VQ-VAE with straight-through estimator (backbone.py:6-76)
# --------------------------------------------------------------------------------------
def t4_init(feature_dim: int, latent_dim: int, num_embeddings: int = 128) -> Params:
    p: Params = {}
    p["encoder.0.weight"], p["encoder.0.bias"] = _linear_init(64, feature_dim)
    p["encoder.2.weight"], p["encoder.2.bias"] = _linear_init(128, 64)
    p["encoder.4.weight"], p["encoder.4.bias"] = _linear_init(latent_dim, 128)
    p["decoder.0.weight"], p["decoder.0.bias"] = _linear_init(128, latent_dim)
    p["decoder.2.weight"], p["decoder.2.bias"] = _linear_init(64, 128)
    p["decoder.4.weight"], p["decoder.4.bias"] = _linear_init(feature_dim, 64)
    emb = torch.empty(num_embeddings, latent_dim)
    torch.nn.init.normal_(emb)  # nn.Embedding default draw (consumes RNG) ...
    emb.uniform_(-1 / num_embeddings, 1 / num_embeddings)  # ... then overwritten (backbone.py:36)
    p["embedding.weight"] = emb
    return p


def t4_forward(p: Params, x: torch.Tensor, commitment_cost: float = 0.25) -> Dict[str, torch.Tensor]:
    h = F.relu(F.linear(x, p["encoder.0.weight"], p["encoder.0.bias"]))
    h = F.relu(F.linear(h, p["encoder.2.weight"], p["encoder.2.bias"]))
    z_e = F.relu(F.linear(h, p["encoder.4.weight"], p["encoder.4.bias"]))
    with torch.no_grad():
        idx = nearest_code_l2(z_e.detach(), p["embedding.weight"].detach(), squared=True)
    z_q = F.embedding(idx, p["embedding.weight"])
    commit = commitment_cost * F.mse_loss(z_q.detach(), z_e)
    emb_loss = F.mse_loss(z_q, z_e.detach())
    qloss = emb_loss + commit
    z_st = z_e + (z_q - z_e).detach()
    z_latent = z_st.clone().detach()
    r = F.relu(F.linear(z_st, p["decoder.0.weight"], p["decoder.0.bias"]))
    r = F.relu(F.linear(r, p["decoder.2.weight"], p["decoder.2.bias"]))
    x_recon = F.relu(F.linear(r, p["decoder.4.weight"], p["decoder.4.bias"]))
    recon = F.mse_loss(x_recon, x)
    return {"z_e": z_e, "indices": idx, "z_q": z_q, "z_latent": z_latent, "x_recon": x_recon,
            "recon_loss": recon, "quantization_loss": qloss, "loss": recon + qloss}


# --------------------------------------------------------------------------------------
# T2: hierarchical VQ-VAE (heirarchical_vqvae.py:8-706) -- MS-TCN encoder/decoder, cosine search,
# STE at both levels, EMA buffers.  k-means init and dead-code replacement are host-side and
# RNG-dependent: excluded (feed pre-initialised codebooks; compare dead masks only).
# --------------------------------------------------------------------------------------
def _conv1d_init(out_c: int, in_c: int, k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    w = torch.empty(out_c, in_c, k)
    torch.nn.init.kaiming_uniform_(w, a=math.sqrt(5))
    bound = 1.0 / math.sqrt(in_c * k)
    b = torch.empty(out_c)
    torch.nn.init.uniform_(b, -bound, bound)
    return w, b


def t2_init(action_dim: int, embed_dim: int = 512, num_subclusters: int = 128, num_clusters: int = 64,
            num_stages: int = 2, num_layers_per_stage: int = 10) -> Params:
    """Same draw order as HierarchicalVQVAE.__init__ (heirarchical_vqvae.py:257-319):
    encoder (input_projection, stages, ln), decoder (stages, output_projection), codebooks."""
    p: Params = {}
    p["encoder.input_projection.weight"], p["encoder.input_projection.bias"] = _linear_init(embed_dim, action_dim)
    for s in range(num_stages):
        for l in range(num_layers_per_stage):
            w, b = _conv1d_init(embed_dim, embed_dim, 3)
            p[f"encoder.stages.{s}.layers.{l}.conv.weight"] = w
            p[f"encoder.stages.{s}.layers.{l}.conv.bias"] = b
    p["encoder.ln.weight"] = torch.ones(embed_dim)
    p["encoder.ln.bias"] = torch.zeros(embed_dim)
    for s in range(num_stages):
        for l in range(num_layers_per_stage):
            w, b = _conv1d_init(embed_dim, embed_dim, 3)
            p[f"decoder.stages.{s}.layers.{l}.conv.weight"] = w
            p[f"decoder.stages.{s}.layers.{l}.conv.bias"] = b
    p["decoder.output_projection.weight"], p["decoder.output_projection.bias"] = _linear_init(action_dim, embed_dim)
    cz = torch.randn(num_subclusters, embed_dim)
    cq = torch.randn(num_clusters, embed_dim)
    torch.nn.init.uniform_(cz, -1.0 / num_subclusters, 1.0 / num_subclusters)
    torch.nn.init.uniform_(cq, -1.0 / num_clusters, 1.0 / num_clusters)
    p["cluster_size_z"] = torch.zeros(num_subclusters)
    p["embed_avg_z"] = cz.clone()
    p["cluster_size_q"] = torch.zeros(num_clusters)
    p["embed_avg_q"] = cq.clone()
    p["codebook_z"] = F.normalize(cz, dim=-1)
    p["codebook_q"] = F.normalize(cq, dim=-1)
    return p


T2_BUFFERS = ("codebook_z", "codebook_q", "cluster_size_z", "cluster_size_q", "embed_avg_z", "embed_avg_q")


def _tcn_stages(p: Params, prefix: str, x_bct: torch.Tensor, num_stages: int, num_layers: int) -> torch.Tensor:
    """[B,C,T] through stages x layers of conv(k=3, dilation 2^l, 'same' pad) -> ReLU -> +x
    (dropout is identity in eval / p=0; heirarchical_vqvae.py:39-57, 75-104)."""
    out = x_bct
    for s in range(num_stages):
        for l in range(num_layers):
            dil = 2 ** l
            w = p[f"{prefix}.stages.{s}.layers.{l}.conv.weight"]
            b = p[f"{prefix}.stages.{s}.layers.{l}.conv.bias"]
            out = F.relu(F.conv1d(out, w, b, padding=dil, dilation=dil)) + out
    return out


def t2_encode(p: Params, actions: torch.Tensor, num_stages: int, num_layers: int) -> torch.Tensor:
    x = F.linear(actions, p["encoder.input_projection.weight"], p["encoder.input_projection.bias"])
    x = _tcn_stages(p, "encoder", x.transpose(1, 2), num_stages, num_layers).transpose(1, 2)
    return F.layer_norm(x, (x.shape[-1],), p["encoder.ln.weight"], p["encoder.ln.bias"])


def t2_decode(p: Params, emb: torch.Tensor, num_stages: int, num_layers: int) -> torch.Tensor:
    x = _tcn_stages(p, "decoder", emb.transpose(1, 2), num_stages, num_layers).transpose(1, 2)
    return F.linear(x, p["decoder.output_projection.weight"], p["decoder.output_projection.bias"])


def _cos_quantize(x_btd: torch.Tensor, codebook: torch.Tensor):
    """heirarchical_vqvae.py:434-465: search on normalised vectors, gather the raw buffer row,
    straight-through on the *normalised* input; distances = 1 - sims."""
    B, T, D = x_btd.shape
    idx, sims = nearest_code_cosine(x_btd.reshape(-1, D), codebook)
    q = codebook[idx].view(B, T, D)
    xn = F.normalize(x_btd, dim=-1)
    q = xn + (q - xn).detach()
    return q, idx.view(B, T), (1.0 - sims).view(B, T, -1)


def t2_forward(p: Params, actions: torch.Tensor, *, training: bool, num_stages: int, num_layers: int,
               ema_decay: float = 0.99, eps: float = 1e-5) -> Dict[str, torch.Tensor]:
    """HierarchicalVQVAE.forward with codebooks already initialised and no dead-code
    replacement (heirarchical_vqvae.py:590-646).  When ``training`` the six buffers in ``p``
    are updated in place, Z level from raw embeddings, Q level from the STE'd quantized_z."""
    emb = t2_encode(p, actions, num_stages, num_layers)
    qz, zi, zd = _cos_quantize(emb, p["codebook_z"])
    qq, qi, qd = _cos_quantize(qz, p["codebook_q"])
    recon = t2_decode(p, qq, num_stages, num_layers)
    if training:
        with torch.no_grad():
            D = emb.shape[-1]
            s, a, c = ema_update_cosine(emb.reshape(-1, D), zi.reshape(-1), p["cluster_size_z"], p["embed_avg_z"],
                                        ema_decay, eps)
            p["cluster_size_z"].copy_(s), p["embed_avg_z"].copy_(a), p["codebook_z"].copy_(c)
            s, a, c = ema_update_cosine(qz.reshape(-1, D), qi.reshape(-1), p["cluster_size_q"], p["embed_avg_q"],
                                        ema_decay, eps)
            p["cluster_size_q"].copy_(s), p["embed_avg_q"].copy_(a), p["codebook_q"].copy_(c)
    return {"embeddings": emb, "quantized_z": qz, "quantized_q": qq, "reconstructed_actions": recon,
            "z_indices": zi, "q_indices": qi, "z_distances": zd, "q_distances": qd}


def t2_loss(out: Dict[str, torch.Tensor], true_actions: torch.Tensor, lambda_rec: float = 1.0,
            commitment_cost: float = 1.0) -> Dict[str, torch.Tensor]:
    """heirarchical_vqvae.py:648-685 (cosine branch)."""
    L_rec = F.mse_loss(out["reconstructed_actions"], true_actions)
    L_cz = F.mse_loss(F.normalize(out["embeddings"], dim=-1), F.normalize(out["quantized_z"].detach(), dim=-1))
    L_cq = F.mse_loss(F.normalize(out["quantized_z"], dim=-1), F.normalize(out["quantized_q"].detach(), dim=-1))
    return {"vqvae_loss": commitment_cost * (L_cz + L_cq) + lambda_rec * L_rec, "L_rec": L_rec,
            "L_commit_z": L_cz, "L_commit_q": L_cq}


def codebook_usage(cluster_size_z: torch.Tensor, cluster_size_q: torch.Tensor, thr_z: float = 1.0,
                   thr_q: float = 1.0) -> Dict[str, float]:
    """heirarchical_vqvae.py:687-706."""
    kz, kq = cluster_size_z.numel(), cluster_size_q.numel()
    z_used = int((cluster_size_z > 0).sum())
    q_used = int((cluster_size_q > 0).sum())
    return {"z_used": z_used, "z_total": kz, "z_usage_pct": 100.0 * z_used / max(1, kz),
            "z_utilization": float((cluster_size_z > 0).float().mean()),
            "q_utilization": float((cluster_size_q > 0).float().mean()),
            "q_used": q_used, "q_total": kq, "q_usage_pct": 100.0 * q_used / max(1, kq),
            "z_dead": int((cluster_size_z < thr_z).sum()), "q_dead": int((cluster_size_q < thr_q).sum())}


# --------------------------------------------------------------------------------------
# Callers / data formats either side of the path
# --------------------------------------------------------------------------------------
def sequence_index_map(lens, seq_length: int, frame_stack: int, pad_seq_length: bool, pad_frame_stack: bool):
    """robomimic/utils/dataset.py:229-260: (demo_of_index, first_index_of_demo) for demos of the given lengths."""
    index_to_demo, demo_seq0, total = [], [], 0
    for d, L in enumerate(lens):
        n = L - (0 if pad_frame_stack else frame_stack - 1) - (0 if pad_seq_length else seq_length - 1)
        if pad_seq_length:
            n = max(n, 1)
        assert n >= 1
        demo_seq0.append(total)
        index_to_demo += [d] * n
        total += n
    return index_to_demo, demo_seq0


def sequence_windows(demos, indices, seq_length: int, frame_stack: int, pad_seq_length: bool, pad_frame_stack: bool,
                     offset=None, scale=None):
    """dataset.py:499-632 for the action key: window [index_in_demo - (fs-1), index_in_demo + T) of the demo, rows outside
    repeated from the first / last frame (TensorUtils.pad_sequence pad_same=True), pad mask, (x - offset) / scale."""
    import numpy as np
    lens = [len(d) for d in demos]
    index_to_demo, demo_seq0 = sequence_index_map(lens, seq_length, frame_stack, pad_seq_length, pad_frame_stack)
    acts, masks = [], []
    for index in indices:
        d = index_to_demo[index]
        L = lens[d]
        i = index - demo_seq0[d] + (0 if pad_frame_stack else frame_stack - 1)
        lo, hi = max(0, i - (frame_stack - 1)), min(L, i + seq_length)
        pad_b, pad_e = max(0, (frame_stack - 1) - i), max(0, i + seq_length - L)
        seq = demos[d][lo:hi]
        seq = np.concatenate([np.repeat(seq[:1], pad_b, 0), seq, np.repeat(seq[-1:], pad_e, 0)], 0)
        if offset is not None:
            seq = (seq - offset) / scale
        acts.append(seq.astype(np.float32))
        masks.append(np.array([0] * pad_b + [1] * (hi - lo) + [0] * pad_e, dtype=bool))
    return np.stack(acts), np.stack(masks)


def linspace_f32(lo: float, hi: float, steps: int):
    """torch.linspace in fp32 (ATen RangeFactories, CPU): step = (hi-lo)/(steps-1) rounded to fp32; fma(step, i, start)
    below the middle, fma(-step, steps-1-i, end) above (ONE rounding per element: the product is exact in double)."""
    import numpy as np
    lo, hi = np.float32(lo), np.float32(hi)
    step = np.float32((hi - lo) / np.float32(steps - 1))
    out = np.empty(steps, np.float32)
    for i in range(steps):
        out[i] = np.float32(float(lo) + float(step) * i) if i < steps // 2 else np.float32(float(hi) - float(step) * (steps - 1 - i))
    return out


def bin_discretize(actions: torch.Tensor, running_min: torch.Tensor, running_max: torch.Tensor, num_bins: int) -> torch.Tensor:
    """robomimic/models/bin_action/backbone.py:41-67: clamp(bucketize(x, linspace(min, max, nb+1)) - 1, 0, nb-1)."""
    cols = []
    for a in range(actions.shape[1]):
        b = torch.from_numpy(linspace_f32(float(running_min[a]), float(running_max[a]), num_bins + 1))
        below = (b[None, :] < actions[:, a:a + 1]).sum(dim=1)  # bucketize(right=False)
        cols.append(torch.clamp(below - 1, 0, num_bins - 1))
    return torch.stack(cols, dim=1)


def bin_embedding_forward(p: Params, actions: torch.Tensor, running_min: torch.Tensor, running_max: torch.Tensor,
                          update: bool = True):
    """AdaptiveBinActionEmbedding.forward (backbone.py:69-89) -> (out, indices, new running_min, new running_max)."""
    if update:
        running_min = torch.minimum(running_min, actions.min(dim=0)[0])
        running_max = torch.maximum(running_max, actions.max(dim=0)[0])
    A = actions.shape[1]
    nb = p["embedding_layers.0.weight"].shape[0]
    idx = bin_discretize(actions, running_min, running_max, nb)
    emb = torch.cat([p[f"embedding_layers.{a}.weight"][idx[:, a]] for a in range(A)], dim=-1)
    h = F.gelu(F.linear(emb, p["output_layer.0.weight"], p["output_layer.0.bias"]))
    out = F.gelu(F.linear(h, p["output_layer.2.weight"], p["output_layer.2.bias"]))
    return out, idx, running_min, running_max



class T1Module(torch.nn.Module):
    """LLFQVAE_V4's parameter tree (backbone_lfqvae_v5.py:52-68) with this oracle's forward: the CPU stand-in for the
    tokenizer inside the restated ICRT policy (tests/test_icrt_cpu.py, bench.py --impl reference --config c5)."""

    def __init__(self, feature_dim, latent_dim, num_codes=1024, hidden_dim=128):
        super().__init__()
        nn = torch.nn
        self.encoder = nn.Sequential(nn.Linear(feature_dim, 64), nn.GELU(), nn.Linear(64, hidden_dim), nn.GELU())
        self.to_latent = nn.Module()
        self.to_latent.W = nn.Parameter(torch.randn(latent_dim, hidden_dim))
        self.to_latent.b = nn.Parameter(torch.zeros(latent_dim))
        self.to_latent.ci = nn.Parameter(torch.ones(latent_dim))
        self.quantizer = nn.Module()
        self.quantizer.codebook = nn.Parameter(torch.randn(num_codes, latent_dim))
        nn.init.kaiming_uniform_(self.quantizer.codebook)
        self.decoder = nn.Sequential(nn.Linear(latent_dim, 64), nn.GELU(), nn.Linear(64, hidden_dim), nn.GELU())
        self.to_output = nn.Linear(hidden_dim, feature_dim)

    def forward(self, x):
        out = t1_forward(dict(self.named_parameters()), x)
        return out["z_latent"], out["loss"]
