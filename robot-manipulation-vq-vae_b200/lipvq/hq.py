"""EMA / k-means / hierarchical Lipschitz quantizers (T3) on liblipvq.so.

Drop-ins for ``LFQQuantizerEMA_KMeans`` (reference robomimic/models/vq_vae/hq_vae.py:19-154) and
``HierarchicalLFQHVQVAE`` (hq_vae.py:289-415): same constructor signatures, parameter/buffer names, return values.
``LFQQuantizerEMA`` (hq_vae.py:224-281) lives in modules.py.

Host-side by design: the one-off sklearn k-means initialisation and the bookkeeping of which codes are dead (mask ->
index lists); the arithmetic -- search, gather, EMA statistics/update, usage statistics, nearest-alive-code lookup --
runs in the CUDA kernels.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _cabi as A
from . import ops
from .hvq import _LinearFn, mse_loss
from .modules import LLFQVAE_V4, LipschitzMLP, _LipschitzFn, _QuantizeFn, _f32c

GELU, NONE = A.ACT_GELU, A.ACT_NONE


class LFQQuantizerEMA_KMeans(nn.Module):
    def __init__(self, num_codes, code_dim, decay=0.99, epsilon=1e-5, dead_threshold=5, replace_strategy="nearest"):
        super().__init__()
        self.num_codes = num_codes
        self.code_dim = code_dim
        self.decay = decay
        self.epsilon = epsilon
        self.dead_threshold = dead_threshold
        self.replace_strategy = replace_strategy
        self.initialized = False  # plain attribute, not saved -- as in the reference (hq_vae.py:48)
        self.codebook = nn.Parameter(torch.randn(num_codes, code_dim))
        nn.init.kaiming_normal_(self.codebook)
        self.ema_cluster_size = nn.Parameter(torch.zeros(num_codes), requires_grad=False)
        self.ema_codebook = nn.Parameter(torch.randn(num_codes, code_dim), requires_grad=False)
        self.register_buffer("usage_counts", torch.zeros(num_codes))
        self.register_buffer("usage_ma", torch.zeros(num_codes))
        self.register_buffer("entropy_ma", torch.tensor(0.0))

    def kmeans_init(self, z_e):
        """hq_vae.py:65-78 -- sklearn on the host, first batch only."""
        from sklearn.cluster import KMeans
        B, D = z_e.shape
        n_samples = min(20000, B)
        sample_idx = torch.randperm(B)[:n_samples]
        sample = z_e.detach()[sample_idx.to(z_e.device)].cpu().numpy()
        centers = KMeans(n_clusters=self.num_codes, n_init="auto", max_iter=50).fit(sample).cluster_centers_
        centers = torch.tensor(centers, dtype=z_e.dtype, device=z_e.device)
        self.codebook.data.copy_(centers)
        self.ema_codebook.data.copy_(centers)
        self.initialized = True

    @torch.no_grad()
    def _replace_dead(self, z_e):
        dead = self.usage_counts < self.dead_threshold
        n_dead = int(dead.sum().item())  # the reference syncs here too (dead.any(), hq_vae.py:136)
        if n_dead == 0:
            return
        dead_idx = dead.nonzero(as_tuple=False).squeeze(1)
        if self.replace_strategy == "nearest":
            alive = (~dead).nonzero(as_tuple=False).squeeze(1)
            if len(alive) > 0:
                cb = self.codebook.data
                alive_codes = cb[alive].contiguous()
                dead_codes = cb[dead_idx].contiguous()
                nearest = ops.nn_search(dead_codes, alive_codes, A.METRIC_L2)  # argmin ||alive - dead||, first-min ties
                sq = torch.zeros(1, dtype=torch.float64, device=cb.device)
                repl, _ = ops.gather_fwd(dead_codes, nearest, alive_codes, sq)
                cb[dead_idx] = repl
        else:
            rand_ids = torch.randint(0, z_e.shape[0], (len(dead_idx),), device=z_e.device)
            self.codebook.data[dead_idx] = z_e.detach()[rand_ids]

    def forward(self, z_e):
        B, D = z_e.shape
        if not self.initialized:
            self.kmeans_init(z_e)
        z_q, indices = _QuantizeFn.apply(z_e, self.codebook)  # rows of the pre-update codebook
        with torch.no_grad():
            z = _f32c(z_e.detach())
            counts = torch.zeros(self.num_codes, dtype=torch.float32, device=z.device)
            sums = torch.zeros(self.num_codes, self.code_dim, dtype=torch.float32, device=z.device)
            ops.ema_stats(z, indices, counts, sums)
            scratch = torch.zeros(1, dtype=torch.float64, device=z.device)
            ops.ema_apply(counts, sums, self.ema_cluster_size.data, self.ema_codebook.data, self.codebook.data, self.decay,
                          self.epsilon, A.EMA_LAPLACE, scratch)  # EMA runs in eval too (no training gate, :103-118)
            ops.usage_stats(counts, self.usage_counts, self.usage_ma, self.entropy_ma.view(1))
            self._replace_dead(z_e)
        return z_q, indices


class _EncMlpFn(torch.autograd.Function):
    """Linear -> GELU -> Linear of the T1 encoder / decoder Sequentials, returning the PRE-activation of the second
    layer (its consumer applies the trailing GELU while loading)."""

    @staticmethod
    def forward(ctx, x, W0, b0, W2, b2):
        x2 = _f32c(x)
        p = [t.detach().contiguous() for t in (W0, b0, W2, b2)]
        a1 = ops.linear_fwd(x2, NONE, p[0], p[1], NONE)
        a2 = ops.linear_fwd(a1, GELU, p[2], p[3], NONE)
        ctx.save_for_backward(x2, a1, p[0], p[2])
        return a2

    @staticmethod
    def backward(ctx, da2):
        x2, a1, W0, W2 = ctx.saved_tensors
        da2 = _f32c(da2)
        dW2, db2 = torch.zeros_like(W2), torch.zeros(W2.shape[0], dtype=torch.float32, device=W2.device)
        ops.linear_bwd_w(da2, a1, GELU, dW2, db2)
        da1 = ops.linear_bwd_x(da2, W2, a1, GELU)
        dW0, db0 = torch.zeros_like(W0), torch.zeros(W0.shape[0], dtype=torch.float32, device=W0.device)
        ops.linear_bwd_w(da1, x2, NONE, dW0, db0)
        dx = ops.linear_bwd_x(da1, W0, None, NONE) if ctx.needs_input_grad[0] else None
        return dx, dW0, db0, dW2, db2


class HierarchicalLFQHVQVAE(nn.Module):
    def __init__(self, feature_dim, z_dim, q_dim, num_z_codes=1024, num_q_codes=512, hidden_dim=128):
        super().__init__()
        # same construction (and RNG draw) order as the reference: four throw-away LLFQVAE_V4 instances around the parts
        self.encoder = LLFQVAE_V4(feature_dim=feature_dim, latent_dim=z_dim, num_codes=num_z_codes,
                                  hidden_dim=hidden_dim).encoder
        self.to_z_latent = LLFQVAE_V4(feature_dim=feature_dim, latent_dim=z_dim, num_codes=num_z_codes,
                                      hidden_dim=hidden_dim).to_latent
        self.z_quantizer = LFQQuantizerEMA_KMeans(num_z_codes, z_dim, dead_threshold=3)
        self.q_quantizer = LFQQuantizerEMA_KMeans(num_q_codes, q_dim, dead_threshold=1)
        self.q_encoder = LipschitzMLP(z_dim, q_dim)
        self.decoder = LLFQVAE_V4(feature_dim=feature_dim, latent_dim=q_dim, num_codes=num_q_codes,
                                  hidden_dim=hidden_dim).decoder
        self.to_output = LLFQVAE_V4(feature_dim=feature_dim, latent_dim=q_dim, num_codes=num_q_codes,
                                    hidden_dim=hidden_dim).to_output

    def vq_loss(self, z_e, z_q):
        return mse_loss(z_e, z_q.detach()), _codebook_mse(z_q, z_e)

    def forward(self, x):
        enc, dec = self.encoder, self.decoder
        h_pre = _EncMlpFn.apply(x, enc[0].weight, enc[0].bias, enc[2].weight, enc[2].bias)
        z_e = _LipschitzFn.apply(h_pre, self.to_z_latent.W, self.to_z_latent.b, self.to_z_latent.ci, GELU)
        z_q, z_idx = self.z_quantizer(z_e)
        commit_z = mse_loss(z_e, z_q.detach())
        codebook_z = _codebook_mse(z_q, z_e)
        q_e = self.q_encoder(z_q.detach())
        q_q, q_idx = self.q_quantizer(q_e)
        commit_q = mse_loss(q_e, q_q.detach())
        codebook_q = _codebook_mse(q_q, q_e)
        d_pre = _EncMlpFn.apply(q_q, dec[0].weight, dec[0].bias, dec[2].weight, dec[2].bias)
        x_recon = _LinearFn.apply(d_pre, self.to_output.weight, self.to_output.bias, GELU)
        recon_loss = mse_loss(x_recon, x)
        loss = recon_loss + 0.25 * (commit_z + codebook_z) + 0.25 * (commit_q + codebook_q)
        return {"loss": loss, "recon_loss": recon_loss, "z_commit": commit_z, "z_codebook": codebook_z,
                "q_commit": commit_q, "q_codebook": codebook_q, "x_recon": x_recon, "z_q": z_q, "q_q": q_q,
                "z_indices": z_idx, "q_indices": q_idx}


def _codebook_mse(z_q, z_e):
    """F.mse_loss(z_q, z_e.detach()): gradient to z_q (and through the gather to the selected code rows)."""
    return mse_loss(z_q, z_e.detach())
