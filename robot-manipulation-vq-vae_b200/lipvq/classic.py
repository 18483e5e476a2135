"""Classic VQ-VAE tokenizer (T4) on liblipvq.so.

Drop-in for ``VQVAE`` of the reference (robomimic/models/vq_vae/backbone.py:6-76): ReLU MLP encoder
F->64->128->D with a trailing ReLU, squared-L2 nearest code over an ``nn.Embedding`` codebook,
commitment (0.25) + embedding MSE, straight-through estimator ``z_e + (z_q - z_e).detach()``, ReLU MLP
decoder D->128->64->F with a trailing ReLU, ``loss = recon + (embedding + commitment)``.
Returns ``(z_latent detached, loss)`` like the reference; same state_dict keys.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _cabi as A
from . import ops

_T4_KEYS = ("encoder.0.weight", "encoder.0.bias", "encoder.2.weight", "encoder.2.bias", "encoder.4.weight",
            "encoder.4.bias", "decoder.0.weight", "decoder.0.bias", "decoder.2.weight", "decoder.2.bias",
            "decoder.4.weight", "decoder.4.bias", "embedding.weight")

R, NONE = A.ACT_RELU, A.ACT_NONE


class _T4Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cc, *params):
        if not x.is_cuda:
            raise RuntimeError("lipvq modules run on CUDA tensors only (no CPU path)")
        x = x.contiguous().float()
        p = {k: v.detach().contiguous() for k, v in zip(_T4_KEYS, params)}
        N, dev = x.shape[0], x.device
        emb = p["embedding.weight"]
        K, D = emb.shape
        a1 = ops.linear_fwd(x, NONE, p["encoder.0.weight"], p["encoder.0.bias"], NONE)
        a2 = ops.linear_fwd(a1, R, p["encoder.2.weight"], p["encoder.2.bias"], NONE)
        z_e = ops.linear_fwd(a2, R, p["encoder.4.weight"], p["encoder.4.bias"], R)  # post-ReLU: the search needs it
        idx = ops.nn_search(z_e, emb, A.METRIC_L2SQ)
        sq = torch.zeros(2, dtype=torch.float64, device=dev)
        z_q, z_st = ops.gather_fwd(z_e, idx, emb, sq[1:], want_ste=True)
        c1 = ops.linear_fwd(z_st, NONE, p["decoder.0.weight"], p["decoder.0.bias"], NONE)
        c2 = ops.linear_fwd(c1, R, p["decoder.2.weight"], p["decoder.2.bias"], NONE)
        xr = ops.linear_fwd(c2, R, p["decoder.4.weight"], p["decoder.4.bias"], R)
        ops.mse_fwd_bwd(xr, x, sq[0:1])
        losses = torch.empty(4, dtype=torch.float32, device=dev)
        # loss = recon + (embedding + cc*commit); both MSE terms share one value
        ops.t1_loss_finalize(sq, float(N) * x.shape[1], float(N) * D, float(cc), 1.0, losses)
        ctx.p, ctx.cc = p, float(cc)
        ctx.save_for_backward(x, a1, a2, z_e, idx, z_q, z_st, c1, c2, xr)
        ctx.mark_non_differentiable(z_st, idx)
        return z_st, losses[0].clone(), idx

    @staticmethod
    def backward(ctx, _dz, dloss, _didx):
        x, a1, a2, z_e, idx, z_q, z_st, c1, c2, xr = ctx.saved_tensors
        p, dev = ctx.p, x.device
        g = {k: torch.zeros_like(v) for k, v in p.items()}
        gs = dloss.detach().to(torch.float32).reshape(1).contiguous()
        scratch = torch.zeros(1, dtype=torch.float64, device=dev)
        dxr = torch.empty_like(xr)
        ops.mse_fwd_bwd(xr, x, scratch, gscale=gs, act_out=R, dY=dxr)  # gradient at decoder.4's pre-activation
        ops.linear_bwd_w(dxr, c2, R, g["decoder.4.weight"], g["decoder.4.bias"])
        dc2 = ops.linear_bwd_x(dxr, p["decoder.4.weight"], c2, R)
        ops.linear_bwd_w(dc2, c1, R, g["decoder.2.weight"], g["decoder.2.bias"])
        dc1 = ops.linear_bwd_x(dc2, p["decoder.2.weight"], c1, R)
        ops.linear_bwd_w(dc1, z_st, NONE, g["decoder.0.weight"], g["decoder.0.bias"])
        dzst = ops.linear_bwd_x(dc1, p["decoder.0.weight"], None, NONE)
        # STE: decoder gradient goes to z_e; codebook rows get the embedding-loss term only
        dze = torch.empty_like(z_e)
        ops.gather_bwd(z_e, z_q, idx, dzst, gs, ctx.cc, 1.0, True, None, p["embedding.weight"].shape[0], dze,
                       g["embedding.weight"])
        da3 = ops.act_bwd_from_output(dze, z_e, R)
        ops.linear_bwd_w(da3, a2, R, g["encoder.4.weight"], g["encoder.4.bias"])
        da2 = ops.linear_bwd_x(da3, p["encoder.4.weight"], a2, R)
        ops.linear_bwd_w(da2, a1, R, g["encoder.2.weight"], g["encoder.2.bias"])
        da1 = ops.linear_bwd_x(da2, p["encoder.2.weight"], a1, R)
        ops.linear_bwd_w(da1, x, NONE, g["encoder.0.weight"], g["encoder.0.bias"])
        return (None, None) + tuple(g[k] for k in _T4_KEYS)


class VQVAE(nn.Module):
    def __init__(self, feature_dim, latent_dim, num_embeddings=128, commitment_cost=0.25):
        super().__init__()
        self.feature_dim = feature_dim
        self.latent_dim = latent_dim
        self.num_embeddings = num_embeddings
        self.commitment_cost = commitment_cost
        self.encoder = nn.Sequential(nn.Linear(feature_dim, 64), nn.ReLU(), nn.Linear(64, 128), nn.ReLU(),
                                     nn.Linear(128, latent_dim), nn.ReLU())
        self.decoder = nn.Sequential(nn.Linear(latent_dim, 128), nn.ReLU(), nn.Linear(128, 64), nn.ReLU(),
                                     nn.Linear(64, feature_dim), nn.ReLU())
        self.embedding = nn.Embedding(num_embeddings, latent_dim)
        self.embedding.weight.data.uniform_(-1 / num_embeddings, 1 / num_embeddings)
        self.last_indices = None

    def forward(self, x):
        if x.dim() != 2 or x.shape[1] != self.feature_dim:
            raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({tuple(x.shape)} vs feature_dim "
                               f"{self.feature_dim})")
        sd = dict(self.named_parameters())
        z_latent, loss, idx = _T4Fn.apply(x, self.commitment_cost, *[sd[k] for k in _T4_KEYS])
        self.last_indices = idx
        return z_latent, loss
