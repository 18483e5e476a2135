"""Drop-in nn.Modules for the reference's tokenizer classes, running on liblipvq.so.

Same constructor signatures, parameter/buffer names (state_dict keys), initialisation draw order
and return conventions as the reference, so ``load_state_dict(strict=True)`` accepts reference
checkpoints and ``ICLObservationGroupEncoder`` (robomimic/models/obs_nets.py:1225-1229,1338) can
instantiate these classes unchanged:

  LipschitzMLP        robomimic/models/vq_vae/backbone_lfqvae_v5.py:15-24
  LFQQuantizer        backbone_lfqvae_v5.py:27-48
  LLFQVAE_V4          backbone_lfqvae_v5.py:51-84      forward(x[N,F]) -> (z_latent[N,D] detached, loss)
  LFQQuantizerEMA     robomimic/models/vq_vae/hq_vae.py:224-281

The nn.Linear / nn.GELU children exist for parameter storage and key names only; arithmetic goes
through hand-written CUDA kernels (autograd.Function wrappers below).  CUDA tensors only -- a CPU
tensor raises, there is no fallback.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch
import torch.nn as nn

from . import _cabi as A
from . import ops, t1


def _f32c(x: torch.Tensor) -> torch.Tensor:
    if not x.is_cuda:
        raise RuntimeError("lipvq modules run on CUDA tensors only (no CPU path)")
    if x.dtype != torch.float32:
        raise TypeError(f"expected float32 input, got {x.dtype}")
    return x.contiguous()


# ----------------------------------------------------------------------------------------------
# LipschitzMLP
# ----------------------------------------------------------------------------------------------
class _LipschitzFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, W, b, ci, act_in=A.ACT_NONE):
        x2 = _f32c(x).reshape(-1, x.shape[-1])
        Wn = ops.lipschitz_scale_fwd(W.detach().contiguous(), ci.detach().contiguous())
        y = ops.linear_fwd(x2, act_in, Wn, b.detach().contiguous(), A.ACT_SIGMOID)
        ctx.save_for_backward(x2, W, ci, Wn, y)
        ctx.xshape, ctx.act_in = x.shape, act_in
        return y.reshape(*x.shape[:-1], W.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, W, ci, Wn, y = ctx.saved_tensors
        dy2 = _f32c(dy).reshape(-1, W.shape[0])
        dpre = ops.act_bwd_from_output(dy2, y, A.ACT_SIGMOID)
        G = torch.zeros_like(W)
        db = torch.zeros(W.shape[0], dtype=torch.float32, device=W.device)
        ops.linear_bwd_w(dpre, x2, ctx.act_in, G, db)
        dW = torch.zeros_like(W)
        dci = torch.zeros_like(ci)
        ops.lipschitz_scale_bwd(G, W.detach().contiguous(), ci.detach().contiguous(), dW, dci)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = ops.linear_bwd_x(dpre, Wn, x2 if ctx.act_in != A.ACT_NONE else None, ctx.act_in).reshape(ctx.xshape)
        return dx, dW, db, dci, None


class LipschitzMLP(nn.Module):
    """sigmoid(x @ (W * min(1, softplus(ci)/||W_i||_1))^T + b); params W~N(0,1), b=0, ci=1."""

    def __init__(self, in_dim, out_dim):
        super().__init__()
        self.W = nn.Parameter(torch.randn(out_dim, in_dim))
        self.b = nn.Parameter(torch.zeros(out_dim))
        self.ci = nn.Parameter(torch.ones(out_dim))

    def forward(self, x):
        return _LipschitzFn.apply(x, self.W, self.b, self.ci)


# ----------------------------------------------------------------------------------------------
# quantizers
# ----------------------------------------------------------------------------------------------
class _QuantizeFn(torch.autograd.Function):
    """(z_q, idx) = nearest code; d z_q flows to the selected codebook rows only (no STE)."""

    @staticmethod
    def forward(ctx, z_e, codebook):
        z = _f32c(z_e)
        cb = codebook.detach().contiguous()
        idx = ops.nn_search(z, cb, A.METRIC_L2)
        sq = torch.zeros(1, dtype=torch.float64, device=z.device)
        z_q, _ = ops.gather_fwd(z, idx, cb, sq)
        ctx.save_for_backward(z, z_q, idx)
        ctx.K = cb.shape[0]
        ctx.mark_non_differentiable(idx)
        return z_q, idx

    @staticmethod
    def backward(ctx, dzq, _didx):
        z, z_q, idx = ctx.saved_tensors
        dcb = torch.zeros(ctx.K, z.shape[1], dtype=torch.float32, device=z.device)
        ops.gather_bwd(z, z_q, idx, _f32c(dzq), None, 0.0, 0.0, False, None, ctx.K, None, dcb)
        return None, dcb


class LFQQuantizer(nn.Module):
    def __init__(self, num_codes, code_dim):
        super().__init__()
        self.num_codes = num_codes
        self.code_dim = code_dim
        self.codebook = nn.Parameter(torch.randn(num_codes, code_dim))
        nn.init.kaiming_uniform_(self.codebook)

    def forward(self, z_e):
        batch_size, latent_dim = z_e.shape  # 2-D only, as the reference unpacks
        return _QuantizeFn.apply(z_e, self.codebook)


class LFQQuantizerEMA(nn.Module):
    """LFQQuantizer + EMA codebook update in training mode (hq_vae.py:224-281)."""

    def __init__(self, num_codes, code_dim, decay=0.99, epsilon=1e-5):
        super().__init__()
        self.num_codes = num_codes
        self.code_dim = code_dim
        self.codebook = nn.Parameter(torch.randn(num_codes, code_dim))
        nn.init.kaiming_uniform_(self.codebook)
        self.register_buffer("ema_cluster_size", torch.zeros(num_codes))
        self.register_buffer("ema_codebook", torch.randn(num_codes, code_dim))
        self.decay = decay
        self.epsilon = epsilon

    @torch.no_grad()
    def ema_update(self, z_e: torch.Tensor, indices: torch.Tensor) -> None:
        z = _f32c(z_e.detach())
        counts = torch.zeros(self.num_codes, dtype=torch.float32, device=z.device)
        sums = torch.zeros(self.num_codes, self.code_dim, dtype=torch.float32, device=z.device)
        ops.ema_stats(z, indices, counts, sums)
        scratch = torch.zeros(1, dtype=torch.float64, device=z.device)
        ops.ema_apply(counts, sums, self.ema_cluster_size, self.ema_codebook, self.codebook.data, self.decay,
                      self.epsilon, A.EMA_SIMPLE, scratch)

    def forward(self, z_e):
        z_q, indices = _QuantizeFn.apply(z_e, self.codebook)  # gathers from the pre-update codebook
        if self.training:
            self.ema_update(z_e, indices)
        return z_q, indices


# ----------------------------------------------------------------------------------------------
# LLFQVAE_V4 -- the LipVQ-VAE tokenizer
# ----------------------------------------------------------------------------------------------
_T1_KEYS = ("encoder.0.weight", "encoder.0.bias", "encoder.2.weight", "encoder.2.bias", "to_latent.W", "to_latent.b",
            "to_latent.ci", "quantizer.codebook", "decoder.0.weight", "decoder.0.bias", "decoder.2.weight",
            "decoder.2.bias", "to_output.weight", "to_output.bias")


class _T1Fn(torch.autograd.Function):
    """Whole-tokenizer forward/backward as one autograd node: (x, 14 params) -> (z_latent, loss)."""

    @staticmethod
    def forward(ctx, x, dims, *params):
        F_, D, K, H = dims
        x2 = _f32c(x)
        p = {k: v.detach().contiguous() for k, v in zip(_T1_KEYS, params)}
        b = t1.T1Buffers(x2.shape[0], F_, D, K, H, x2.device)
        t1.forward(p, x2, b)
        ctx.b, ctx.p, ctx.x = b, p, x2
        z_latent, z_e, idx = b.t["z_q"], b.t["z_e"], b.t["idx"]
        ctx.mark_non_differentiable(z_latent, z_e, idx)
        return z_latent, b.t["losses"][0].clone(), idx, z_e

    @staticmethod
    def backward(ctx, _dz, dloss, _didx, _dze):
        b, p, x2 = ctx.b, ctx.p, ctx.x
        g = {k: torch.zeros_like(v) for k, v in p.items()}
        gs = dloss.detach().to(torch.float32).reshape(1).contiguous()
        t1.backward(p, g, x2, b, gscale=gs)
        return (None, None) + tuple(g[k] for k in _T1_KEYS)


class LLFQVAE_V4(nn.Module):
    def __init__(self, feature_dim, latent_dim, num_codes=1024, hidden_dim=128):
        super().__init__()
        self.encoder = nn.Sequential(nn.Linear(feature_dim, 64), nn.GELU(), nn.Linear(64, hidden_dim), nn.GELU())
        self.to_latent = LipschitzMLP(hidden_dim, latent_dim)
        self.quantizer = LFQQuantizer(num_codes, latent_dim)
        self.decoder = nn.Sequential(nn.Linear(latent_dim, 64), nn.GELU(), nn.Linear(64, hidden_dim), nn.GELU())
        self.to_output = nn.Linear(hidden_dim, feature_dim)
        self._dims = (feature_dim, latent_dim, num_codes, hidden_dim)
        self.last_indices = None  # int64 [N] of the most recent forward (not in the reference; read-only convenience)

    def _params(self) -> Dict[str, torch.Tensor]:
        sd = dict(self.named_parameters())
        return {k: sd[k] for k in _T1_KEYS}

    def forward(self, x):
        if x.dim() != 2 or x.shape[1] != self._dims[0]:
            raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({tuple(x.shape)} vs feature_dim "
                               f"{self._dims[0]})")
        dims = (self._dims[0], self._dims[1], self.quantizer.codebook.shape[0], self._dims[3])
        p = self._params()
        z_latent, loss, idx, z_e = _T1Fn.apply(x, dims, *[p[k] for k in _T1_KEYS])
        self.last_indices = idx
        if isinstance(self.quantizer, LFQQuantizerEMA) and self.training:
            # hq_vae.py:262-279: the gather above read the pre-update codebook; now overwrite it from the EMA buffers
            self.quantizer.ema_update(z_e, idx)
        return z_latent, loss

    @torch.no_grad()
    def encode(self, x) -> Tuple[torch.Tensor, torch.Tensor]:
        """encode + VQ only: x [N,F] -> (z_e [N,D], indices [N] int64).  The bulk-tokenisation path."""
        x2 = _f32c(x)
        dims = (self._dims[0], self._dims[1], self.quantizer.codebook.shape[0], self._dims[3])
        b = t1.T1Buffers(x2.shape[0], *dims, x2.device)
        t1.forward({k: v.detach() for k, v in self._params().items()}, x2, b, encode_only=True)
        return b.t["z_e"], b.t["idx"]
