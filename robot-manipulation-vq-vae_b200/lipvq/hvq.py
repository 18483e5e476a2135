"""Hierarchical VQ-VAE tokenizer (T2) on liblipvq.so.

Drop-in for ``HierarchicalVQVAE`` and its MS-TCN encoder/decoder (reference
robomimic/algo/heirarchical_vqvae.py:8-706): same constructor signature, sub-module / parameter /
buffer names (state_dict keys), forward(actions [B,T,A], training) -> 8-key dict,
compute_vqvae_loss(...) -> 4-key dict, get_codebook_usage() -> 10-key dict.

Each differentiable piece is a torch.autograd.Function over hand-written kernels; torch's autograd
only chains them (plumbing).  Activations stay token-major [B*T, E]: the reference's transposes to
[B,E,T] for cuDNN disappear.  Host-side by design (one-off / RNG-dependent): sklearn
k-means initialisation and dead-code replacement.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from . import _cabi as A
from . import ops

NONE = A.ACT_NONE


def _cuda_f32(x: torch.Tensor) -> torch.Tensor:
    if not x.is_cuda:
        raise RuntimeError("lipvq modules run on CUDA tensors only (no CPU path)")
    if x.dtype != torch.float32:
        raise TypeError(f"expected float32, got {x.dtype}")
    return x.contiguous()


class _LinearFn(torch.autograd.Function):
    """y = act_in(x) W^T + b over the last dim (nn.Linear; ``act_in`` when x is a stored pre-activation, in which
    case the gradient returned for x is the gradient at that pre-activation)."""

    @staticmethod
    def forward(ctx, x, W, b, act_in=NONE):
        x2 = _cuda_f32(x).reshape(-1, x.shape[-1])
        Wc, bc = W.detach().contiguous(), b.detach().contiguous()
        y = ops.linear_fwd(x2, act_in, Wc, bc, NONE)
        ctx.save_for_backward(x2, Wc)
        ctx.xshape, ctx.act_in = x.shape, act_in
        return y.reshape(*x.shape[:-1], W.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, Wc = ctx.saved_tensors
        dy2 = _cuda_f32(dy).reshape(-1, Wc.shape[0])
        dW = torch.zeros_like(Wc)
        db = torch.zeros(Wc.shape[0], dtype=torch.float32, device=Wc.device)
        ops.linear_bwd_w(dy2, x2, ctx.act_in, dW, db)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = ops.linear_bwd_x(dy2, Wc, x2 if ctx.act_in != NONE else None, ctx.act_in).reshape(ctx.xshape)
        return dx, dW, db, None


class _ConvLayerFn(torch.autograd.Function):
    """dropout(relu(conv1d_k3(x, dilation))) + x on token-major x [N,E] (heirarchical_vqvae.py:39-57)."""

    @staticmethod
    def forward(ctx, x, T, W, b, dilation, drop_p, seed):
        x2 = _cuda_f32(x)
        Wc = W.detach().contiguous()
        Wf, Wb = ops.conv_pack_weights(Wc)
        out, mask = ops.dilated_conv_fwd(x2, T, Wf, b.detach().contiguous(), dilation, drop_p, seed)
        ctx.save_for_backward(x2, mask, Wb)
        ctx.cfg = (T, dilation, drop_p, tuple(Wc.shape))
        return out

    @staticmethod
    def backward(ctx, dout):
        x2, mask, Wb = ctx.saved_tensors
        T, dilation, drop_p, wshape = ctx.cfg
        d2 = _cuda_f32(dout)
        Eo, Ei, _ = wshape
        dWf = torch.zeros(Eo, 3, Ei, dtype=torch.float32, device=x2.device)
        db = torch.zeros(Eo, dtype=torch.float32, device=x2.device)
        ops.dilated_conv_bwd_w(d2, mask, x2, T, dilation, dWf, db, drop_p)
        dW = torch.zeros(wshape, dtype=torch.float32, device=x2.device)
        ops.conv_unpack_grad(dWf, dW)
        dx = ops.dilated_conv_bwd_x(d2, mask, T, Wb, dilation, drop_p) if ctx.needs_input_grad[0] else None
        return dx, None, dW, db, None, None, None


class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, eps):
        x2 = _cuda_f32(x)
        wc = w.detach().contiguous()
        y, mean, rstd = ops.layernorm_fwd(x2, wc, b.detach().contiguous(), eps)
        ctx.save_for_backward(x2, mean, rstd, wc)
        return y

    @staticmethod
    def backward(ctx, dy):
        x2, mean, rstd, wc = ctx.saved_tensors
        dw, db = torch.zeros_like(wc), torch.zeros_like(wc)
        dx = ops.layernorm_bwd(_cuda_f32(dy), x2, mean, rstd, wc, dw, db)
        return dx, dw, db, None


class _NormalizeFn(torch.autograd.Function):
    """F.normalize(x, dim=-1)"""

    @staticmethod
    def forward(ctx, x):
        x2 = _cuda_f32(x)
        ctx.save_for_backward(x2)
        return ops.l2_normalize_rows(x2)

    @staticmethod
    def backward(ctx, dy):
        (x2,) = ctx.saved_tensors
        return ops.l2_normalize_rows_bwd(x2, _cuda_f32(dy))


def normalize(x):
    return _NormalizeFn.apply(x)


class _QuantSteFn(torch.autograd.Function):
    """nearest code + straight-through: value base + (codebook[idx] - base), gradient -> base.
    ``query``/``code_search`` are what the search compares (normalised rows, DOT metric, for the cosine branch;
    the raw tensors with the L2 metric otherwise); ``codebook`` is the raw buffer the rows are gathered from."""

    @staticmethod
    def forward(ctx, base, query, codebook, code_search, metric):
        b2, q2 = _cuda_f32(base), _cuda_f32(query)
        idx = ops.nn_search(q2, code_search, metric)
        sq = torch.zeros(1, dtype=torch.float64, device=b2.device)
        _, st = ops.gather_fwd(b2, idx, codebook, sq, want_ste=True)
        ctx.mark_non_differentiable(idx)
        return st, idx

    @staticmethod
    def backward(ctx, dst, _didx):
        return dst, None, None, None, None


class _MseFn(torch.autograd.Function):
    """F.mse_loss(y, t) with gradient to y only (the reference's targets are detached or data)."""

    @staticmethod
    def forward(ctx, y, t):
        y2, t2 = _cuda_f32(y), _cuda_f32(t)
        sq = torch.zeros(1, dtype=torch.float64, device=y2.device)
        ops.mse_fwd_bwd(y2, t2, sq)
        ctx.save_for_backward(y2, t2)
        return (sq / y2.numel()).to(torch.float32).reshape(())

    @staticmethod
    def backward(ctx, g):
        y2, t2 = ctx.saved_tensors
        dy = torch.empty_like(y2)
        scratch = torch.zeros(1, dtype=torch.float64, device=y2.device)
        ops.mse_fwd_bwd(y2, t2, scratch, gscale=g.detach().to(torch.float32).reshape(1).contiguous(), dY=dy)
        return dy, None


def mse_loss(y, t):
    return _MseFn.apply(y, t.detach())


# ----------------------------------------------------------------------------------------------
# MS-TCN encoder / decoder (module tree mirrors the reference so state_dict keys match)
# ----------------------------------------------------------------------------------------------
class DilatedTemporalConvLayer(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size=3, dilation=1, dropout=0.0):
        super().__init__()
        if in_channels != out_channels or kernel_size != 3:
            raise NotImplementedError("the tokenizer path only builds E->E, kernel_size 3 layers "
                                      "(heirarchical_vqvae.py:75-92); the 1x1 residual conv is not on it")
        self.padding = (kernel_size - 1) * dilation // 2
        self.dilation = dilation
        self.conv = nn.Conv1d(in_channels, out_channels, kernel_size, padding=self.padding, dilation=dilation)
        self.relu = nn.ReLU()
        self.dropout = nn.Dropout(dropout) if dropout > 0 else None
        self.residual = None

    def forward_tokens(self, x2d, T):
        """x2d [B*T, E] token-major -> [B*T, E]"""
        p = self.dropout.p if (self.dropout is not None and self.training) else 0.0
        seed = 0
        if p > 0:
            seed = int(torch.randint(0, 2 ** 31 - 1, (1,)).item())  # host RNG: follows torch.manual_seed
        return _ConvLayerFn.apply(x2d, T, self.conv.weight, self.conv.bias, self.dilation, float(p), seed)

    def forward(self, x):
        """x [B, C, T] (the reference's layout) -> [B, C, T]"""
        B, C, T = x.shape
        y = self.forward_tokens(x.transpose(1, 2).reshape(B * T, C), T)
        return y.reshape(B, T, C).transpose(1, 2)


class MSTCNStage(nn.Module):
    def __init__(self, num_layers, in_channels, out_channels, kernel_size=3, dropout=0.0):
        super().__init__()
        self.num_layers = num_layers
        self.layers = nn.ModuleList()
        self.layers.append(DilatedTemporalConvLayer(in_channels, out_channels, kernel_size, dilation=1, dropout=dropout))
        for i in range(1, num_layers):
            self.layers.append(DilatedTemporalConvLayer(out_channels, out_channels, kernel_size, dilation=2 ** i,
                                                        dropout=dropout))

    def forward_tokens(self, x2d, T):
        for layer in self.layers:
            x2d = layer.forward_tokens(x2d, T)
        return x2d

    def forward(self, x):
        B, C, T = x.shape
        y = self.forward_tokens(x.transpose(1, 2).reshape(B * T, C), T)
        return y.reshape(B, T, C).transpose(1, 2)


class ActionEncoder(nn.Module):
    def __init__(self, action_dim, embed_dim, num_stages=2, num_layers_per_stage=10, dropout=0.1):
        super().__init__()
        self.action_dim = action_dim
        self.embed_dim = embed_dim
        self.input_projection = nn.Linear(action_dim, embed_dim)
        self.stages = nn.ModuleList([MSTCNStage(num_layers_per_stage, embed_dim, embed_dim, dropout=dropout)
                                     for _ in range(num_stages)])
        self.ln = nn.LayerNorm(embed_dim)

    def forward(self, actions):
        B, T, _ = actions.shape
        x = _LinearFn.apply(actions, self.input_projection.weight, self.input_projection.bias).reshape(B * T, self.embed_dim)
        for stage in self.stages:
            x = stage.forward_tokens(x, T)
        x = _LayerNormFn.apply(x, self.ln.weight, self.ln.bias, self.ln.eps)
        return x.reshape(B, T, self.embed_dim)


class ActionDecoder(nn.Module):
    def __init__(self, embed_dim, action_dim, num_stages=2, num_layers_per_stage=10, dropout=0.1):
        super().__init__()
        self.embed_dim = embed_dim
        self.action_dim = action_dim
        self.stages = nn.ModuleList([MSTCNStage(num_layers_per_stage, embed_dim, embed_dim, dropout=dropout)
                                     for _ in range(num_stages)])
        self.output_projection = nn.Linear(embed_dim, action_dim)

    def forward(self, embeddings):
        B, T, D = embeddings.shape
        x = embeddings.reshape(B * T, D)
        for stage in self.stages:
            x = stage.forward_tokens(x, T)
        y = _LinearFn.apply(x, self.output_projection.weight, self.output_projection.bias)
        return y.reshape(B, T, self.action_dim)


class _LazyOutputs(dict):
    """forward()'s result dict.  ``z_distances`` / ``q_distances`` ([B,T,K] = 1 - sims) are only ever returned by the
    reference, never consumed (icl.py reads quantized_q, reconstructed_actions, the losses), so they are produced on
    first access instead of on every step."""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self._lazy = {}

    def set_lazy(self, key, fn):
        self._lazy[key] = fn
        dict.__setitem__(self, key, None)

    def __getitem__(self, key):
        if key in self._lazy:
            dict.__setitem__(self, key, self._lazy.pop(key)())
        return dict.__getitem__(self, key)

    def get(self, key, default=None):
        return self[key] if key in self else default

    def items(self):
        return [(k, self[k]) for k in self.keys()]

    def values(self):
        return [self[k] for k in self.keys()]


# ----------------------------------------------------------------------------------------------
# HierarchicalVQVAE
# ----------------------------------------------------------------------------------------------
class HierarchicalVQVAE(nn.Module):
    def __init__(self, action_dim: int, embed_dim: int = 512, num_subclusters: int = 128, num_clusters: int = 64,
                 commitment_cost: float = 1.0, ema_decay: float = 0.99, num_stages=2, beta=0.8, num_layers_per_stage=10,
                 dropout=0.1, eps: float = 1e-5, replace_threshold_z: float = 1.0, replace_threshold_q: float = 1.0,
                 kmeans_init: bool = True, use_cosine: bool = True, normalize_after_ema: bool = True,
                 verbose: bool = False):
        super().__init__()
        self.encoder = ActionEncoder(action_dim=action_dim, embed_dim=embed_dim, num_stages=num_stages,
                                     num_layers_per_stage=num_layers_per_stage, dropout=dropout)
        self.decoder = ActionDecoder(embed_dim=embed_dim, action_dim=action_dim, num_stages=num_stages,
                                     num_layers_per_stage=num_layers_per_stage, dropout=dropout)
        self.action_dim = action_dim
        self.embed_dim = embed_dim
        self.num_subclusters = num_subclusters
        self.num_clusters = num_clusters
        self.commitment_cost = commitment_cost
        self.ema_decay = ema_decay
        self.eps = eps
        self.replace_threshold_z = replace_threshold_z
        self.replace_threshold_q = replace_threshold_q
        self.kmeans_init = kmeans_init
        self.use_cosine = use_cosine
        self.normalize_after_ema = normalize_after_ema
        self.verbose = verbose
        self.beta = beta  # stored, unused -- as in the reference (heirarchical_vqvae.py:286)
        self.register_buffer("codebook_z", torch.randn(num_subclusters, embed_dim))
        self.register_buffer("codebook_q", torch.randn(num_clusters, embed_dim))
        nn.init.uniform_(self.codebook_z, -1.0 / num_subclusters, 1.0 / num_subclusters)
        nn.init.uniform_(self.codebook_q, -1.0 / num_clusters, 1.0 / num_clusters)
        self.register_buffer("cluster_size_z", torch.zeros(num_subclusters))
        self.register_buffer("embed_avg_z", self.codebook_z.clone())
        self.register_buffer("cluster_size_q", torch.zeros(num_clusters))
        self.register_buffer("embed_avg_q", self.codebook_q.clone())
        self.register_buffer("codebooks_initialized", torch.tensor(False, dtype=torch.bool))
        if self.use_cosine:  # construction happens on the host (CPU tensors): plain torch, once
            self.codebook_z.copy_(torch.nn.functional.normalize(self.codebook_z, dim=-1))
            self.codebook_q.copy_(torch.nn.functional.normalize(self.codebook_q, dim=-1))

    # -- search helpers ------------------------------------------------------------------------
    def _search_operands(self, flat, codebook):
        """(query rows, codebook rows as compared, metric): normalised + DOT for cosine, raw + L2 otherwise."""
        if self.use_cosine:
            return ops.l2_normalize_rows(flat), ops.l2_normalize_rows(codebook), A.METRIC_DOT
        return flat, codebook, A.METRIC_L2

    def _distances(self, query, code_search):
        """[N,K] `1 - sims` (cosine) / `1 + cdist` (L2 branch), heirarchical_vqvae.py:461,494 -- lazy output only."""
        if self.use_cosine:
            sims = ops.linear_fwd(query, NONE, code_search, None, NONE)
            return ops.affine(sims, -1.0, 1.0, out=sims)
        return 1.0 + torch.cdist(query, code_search)  # debugging output of the non-default branch

    def _assign_stats(self, flat, codebook, size_buf, avg_buf):
        q, c, metric = self._search_operands(flat, codebook)
        idx = ops.nn_search(q, c, metric)
        counts = torch.zeros_like(size_buf)
        sums = torch.zeros_like(avg_buf)
        ops.ema_stats(flat, idx, counts, sums)
        size_buf.copy_(counts)
        avg_buf.copy_(sums)

    @torch.no_grad()
    def initialize_codebooks_with_current_batch(self, embeddings, sample_limit=10000):
        """One-off host-side k-means init (heirarchical_vqvae.py:321-419): sklearn on <= sample_limit rows, then
        cluster sizes / sums of the sample under the new centres (device kernels)."""
        if self.codebooks_initialized:
            return
        device = embeddings.device
        D = embeddings.shape[-1]
        flat = embeddings.reshape(-1, D).cpu().numpy()
        if flat.shape[0] > sample_limit:
            sample = flat[np.random.choice(flat.shape[0], sample_limit, replace=False)]
        else:
            sample = flat
        if self.kmeans_init:
            from sklearn.cluster import KMeans
            km_z = KMeans(n_clusters=self.num_subclusters, n_init=10, random_state=42, max_iter=300).fit(sample)
            centers_z = torch.from_numpy(km_z.cluster_centers_).float()
            km_q = KMeans(n_clusters=self.num_clusters, n_init=10, random_state=42, max_iter=300).fit(sample)
            centers_q = torch.from_numpy(km_q.cluster_centers_).float()
        else:
            centers_z = torch.randn(self.num_subclusters, self.embed_dim)
            centers_q = torch.randn(self.num_clusters, self.embed_dim)
        centers_z, centers_q = centers_z.to(device).contiguous(), centers_q.to(device).contiguous()
        if self.use_cosine:
            centers_z, centers_q = ops.l2_normalize_rows(centers_z), ops.l2_normalize_rows(centers_q)
        self.codebook_z.copy_(centers_z)
        self.codebook_q.copy_(centers_q)
        flat_t = torch.from_numpy(np.ascontiguousarray(sample)).to(device)
        self._assign_stats(flat_t, self.codebook_z, self.cluster_size_z, self.embed_avg_z)
        self._assign_stats(flat_t, self.codebook_q, self.cluster_size_q, self.embed_avg_q)
        self.codebooks_initialized.fill_(True)
        if self.verbose:
            print(f"[KMeans init] Z centers: {self.num_subclusters}, Q centers: {self.num_clusters}")

    def _quantize(self, x_btd, codebook, use_ste=True):
        B, T, D = x_btd.shape
        flat = x_btd.reshape(-1, D)
        base = normalize(flat) if self.use_cosine else flat
        with torch.no_grad():
            if self.use_cosine:
                query, code_s, metric = base.detach(), ops.l2_normalize_rows(codebook), A.METRIC_DOT
            else:
                query, code_s, metric = flat.detach().contiguous(), codebook.clone(), A.METRIC_L2
        st, idx = _QuantSteFn.apply(base, query, codebook, code_s, metric)
        if not use_ste:
            st = st.detach()  # value codebook[idx] up to the STE expression's rounding; no gradient to the input
        return st.view(B, T, D), idx.view(B, T), (query, code_s)

    def quantize_z(self, embeddings, use_ste=True):
        q, idx, ops_ = self._quantize(embeddings, self.codebook_z, use_ste)
        B, T, _ = embeddings.shape
        return q, idx, self._distances(*ops_).view(B, T, -1)

    def quantize_q(self, subclusters, use_ste=True):
        q, idx, ops_ = self._quantize(subclusters, self.codebook_q, use_ste)
        B, T, _ = subclusters.shape
        return q, idx, self._distances(*ops_).view(B, T, -1)

    @torch.no_grad()
    def _ema_update(self, embeddings_flat, indices_flat, embed_avg_buffer, cluster_size_buffer, codebook_buffer):
        x = _cuda_f32(embeddings_flat.detach())
        counts = torch.zeros_like(cluster_size_buffer)
        sums = torch.zeros_like(embed_avg_buffer)
        ops.ema_stats(x, indices_flat.reshape(-1), counts, sums)
        variant = A.EMA_NORMALIZE if (self.normalize_after_ema and self.use_cosine) else A.EMA_SIMPLE
        scratch = torch.zeros(1, dtype=torch.float64, device=x.device)
        ops.ema_apply(counts, sums, cluster_size_buffer, embed_avg_buffer, codebook_buffer, self.ema_decay, self.eps,
                      variant, scratch)

    @torch.no_grad()
    def replace_dead_codes(self, embeddings, subclusters):
        """heirarchical_vqvae.py:543-588.  Rare and RNG-dependent: the dead counts come from one kernel + one D2H
        copy per level (instead of .any()/.item() pairs); the replacement itself is host-driven torch indexing."""
        for size, thr, cb, avg, src in ((self.cluster_size_z, self.replace_threshold_z, self.codebook_z, self.embed_avg_z,
                                         embeddings),
                                        (self.cluster_size_q, self.replace_threshold_q, self.codebook_q, self.embed_avg_q,
                                         subclusters)):
            num_dead = int(ops.codebook_usage(size, float(thr))[1].item())
            if num_dead == 0:
                continue
            flat = src.reshape(-1, src.shape[-1]).detach()
            if flat.size(0) == 0:
                return
            dead_mask = size < thr
            idx = torch.randint(0, flat.size(0), (num_dead,), device=flat.device)
            repl = flat[idx].contiguous()
            if self.use_cosine:
                repl = ops.l2_normalize_rows(repl)
            cb[dead_mask] = repl
            avg[dead_mask] = repl
            size[dead_mask] = 1.0

    @torch.no_grad()
    def replace_dead_codes_static(self, embeddings, subclusters):
        """replace_dead_codes without a device-to-host copy and with shapes that do not depend on the data (CUDA-graph
        capture, HvqTrainer): EVERY code draws a candidate row, dead codes take theirs.  Same law as the reference's
        (each dead code gets an independent uniform row of the batch, heirarchical_vqvae.py:543-588), not the same
        random stream: the reference draws num_dead indices, this draws K."""
        for size, thr, cb, avg, src in ((self.cluster_size_z, self.replace_threshold_z, self.codebook_z, self.embed_avg_z,
                                         embeddings),
                                        (self.cluster_size_q, self.replace_threshold_q, self.codebook_q, self.embed_avg_q,
                                         subclusters)):
            flat = src.reshape(-1, src.shape[-1]).detach()
            if flat.size(0) == 0:
                return
            dead = size < thr
            idx = torch.randint(0, flat.size(0), (size.numel(),), device=flat.device)
            repl = flat[idx].contiguous()
            if self.use_cosine:
                repl = ops.l2_normalize_rows(repl)
            cb.copy_(torch.where(dead[:, None], repl, cb))
            avg.copy_(torch.where(dead[:, None], repl, avg))
            size.copy_(torch.where(dead, torch.ones_like(size), size))

    def forward(self, actions, training=True):
        embeddings = self.encoder(actions)
        graph_safe = getattr(self, "_graph_safe", False)  # HvqTrainer: initialised beforehand, no host round trips
        if not graph_safe and not self.codebooks_initialized:
            self.initialize_codebooks_with_current_batch(embeddings)
        B, T, D = embeddings.shape
        quantized_z, z_indices, z_ops = self._quantize(embeddings, self.codebook_z)
        quantized_q, q_indices, q_ops = self._quantize(quantized_z, self.codebook_q)
        reconstructed = self.decoder(quantized_q)
        out = _LazyOutputs({"embeddings": embeddings, "quantized_z": quantized_z, "quantized_q": quantized_q,
                            "reconstructed_actions": reconstructed, "z_indices": z_indices, "q_indices": q_indices})
        # the searches used the pre-update codebooks; so must the (lazy) distances
        out.set_lazy("z_distances", lambda o=z_ops: self._distances(*o).view(B, T, -1))
        out.set_lazy("q_distances", lambda o=q_ops: self._distances(*o).view(B, T, -1))
        if training:
            self._ema_update(embeddings.reshape(-1, D), z_indices.reshape(-1), self.embed_avg_z, self.cluster_size_z,
                             self.codebook_z)
            self._ema_update(quantized_z.reshape(-1, D), q_indices.reshape(-1), self.embed_avg_q, self.cluster_size_q,
                             self.codebook_q)
            if graph_safe:
                self.replace_dead_codes_static(embeddings, quantized_z)
            else:
                self.replace_dead_codes(embeddings, quantized_z)
        return out

    # -- losses --------------------------------------------------------------------------------
    def compute_reconstruction_loss(self, pred, target):
        return mse_loss(pred, target)

    def compute_commitment_loss_z(self, embeddings, quantized_z):
        if self.use_cosine:
            return mse_loss(normalize(embeddings), normalize(quantized_z.detach()))
        return mse_loss(embeddings, quantized_z.detach())

    def compute_commitment_loss_q(self, quantized_z, quantized_q):
        if self.use_cosine:
            return mse_loss(normalize(quantized_z), normalize(quantized_q.detach()))
        return mse_loss(quantized_z, quantized_q.detach())

    def compute_vqvae_loss(self, outputs, true_actions, lambda_rec=1.0):
        L_rec = self.compute_reconstruction_loss(outputs["reconstructed_actions"], true_actions)
        L_cz = self.compute_commitment_loss_z(outputs["embeddings"], outputs["quantized_z"])
        L_cq = self.compute_commitment_loss_q(outputs["quantized_z"], outputs["quantized_q"])
        total = self.commitment_cost * (L_cz + L_cq) + lambda_rec * L_rec
        return {"vqvae_loss": total, "L_rec": L_rec, "L_commit_z": L_cz, "L_commit_q": L_cq}

    def get_codebook_usage(self):
        with torch.no_grad():
            uz = ops.codebook_usage(self.cluster_size_z, float(self.replace_threshold_z))
            uq = ops.codebook_usage(self.cluster_size_q, float(self.replace_threshold_q))
            z_used, z_dead, q_used, q_dead = torch.cat([uz, uq]).tolist()  # one D2H copy
            return {"z_used": z_used, "z_total": self.num_subclusters,
                    "z_usage_pct": 100.0 * z_used / max(1, self.num_subclusters),
                    "z_utilization": z_used / self.num_subclusters, "q_utilization": q_used / self.num_clusters,
                    "q_used": q_used, "q_total": self.num_clusters,
                    "q_usage_pct": 100.0 * q_used / max(1, self.num_clusters), "z_dead": z_dead, "q_dead": q_dead}
