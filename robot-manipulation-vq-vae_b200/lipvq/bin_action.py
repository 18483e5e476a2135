"""AdaptiveBinActionEmbedding -- the A/B baseline action tokenizer behind the same boundary.

Drop-in for robomimic/models/bin_action/backbone.py:5-89 (selected by ``bin_enabled`` at obs_nets.py:1214-1217, called at
:1346): same ctor signature, parameter / buffer names (``running_min``, ``running_max``, ``embedding_layers.{i}.weight``,
``output_layer.{0,2}.*``) and the plain-attribute step counter that stops the statistics after ``num_step_stop`` calls.
Running min/max, linspace boundaries, bucketize, the per-dimension embedding gather + concat and its gradient are one
CUDA kernel each (csrc/caller.cu); the two Linear+GELU layers go through the library's linear kernels.  CUDA only.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _cabi as A
from . import ops
from .hvq import _LinearFn
from .modules import _f32c


class _BinEmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, actions, running_min, running_max, update, *tables):
        a = _f32c(actions)
        tab = torch.stack([t.detach() for t in tables]).contiguous()  # [A, num_bins, E]
        out, idx = ops.bin_embed_fwd(a, running_min, running_max, tab, update)
        ctx.save_for_backward(idx)
        ctx.nb, ctx.E = tab.shape[1], tab.shape[2]
        ctx.mark_non_differentiable(idx)
        return out, idx

    @staticmethod
    def backward(ctx, dout, _didx):
        (idx,) = ctx.saved_tensors
        dt = ops.bin_embed_bwd(_f32c(dout), idx, ctx.nb, ctx.E)
        return (None, None, None, None) + tuple(dt[i] for i in range(dt.shape[0]))


class AdaptiveBinActionEmbedding(nn.Module):
    def __init__(self, action_dim, output_dim, num_bins=20, embedding_dim=64, num_step_stop=10000):
        super().__init__()
        self.action_dim, self.num_bins, self.embedding_dim = action_dim, num_bins, embedding_dim
        self.register_buffer("running_min", torch.full((action_dim,), float("inf")))
        self.register_buffer("running_max", torch.full((action_dim,), float("-inf")))
        self.embedding_layers = nn.ModuleList([nn.Embedding(num_embeddings=num_bins, embedding_dim=embedding_dim)
                                               for _ in range(action_dim)])
        self.output_layer = nn.Sequential(nn.Linear(embedding_dim * action_dim, embedding_dim * action_dim // 2), nn.GELU(),
                                          nn.Linear(embedding_dim * action_dim // 2, output_dim), nn.GELU())
        self._num_step, self._num_step_stop, self._update_enabled = 0, num_step_stop, True
        self.last_indices = None

    def discretize(self, actions):
        _, idx = ops.bin_embed_fwd(_f32c(actions), self.running_min, self.running_max,
                                   torch.stack([l.weight.detach() for l in self.embedding_layers]).contiguous(), False)
        return idx

    def forward(self, actions):
        update = self._update_enabled
        if update:
            self._num_step += 1
            if self._num_step >= self._num_step_stop:
                self._update_enabled = False
        emb, idx = _BinEmbedFn.apply(actions, self.running_min, self.running_max, update, *[l.weight for l in self.embedding_layers])
        self.last_indices = idx
        l0, l2 = self.output_layer[0], self.output_layer[2]
        h = _LinearFn.apply(emb, l0.weight, l0.bias, A.ACT_NONE)       # pre-activation of the first GELU
        return _LinearGeluOutFn.apply(h, l2.weight, l2.bias, A.ACT_GELU)  # GELU on the way in (h is a pre-activation) and out


class _LinearGeluOutFn(torch.autograd.Function):
    """y = gelu(act_in(x) W^T + b): the layer's epilogue evaluates GELU and its derivative once (lipvq_linear_fwd_dual_f32);
    backward multiplies the incoming gradient by the stored derivative (one elementwise torch op on this baseline path)."""

    @staticmethod
    def forward(ctx, x, W, b, act_in):
        x2 = _f32c(x)
        Wc, bc = W.detach().contiguous(), b.detach().contiguous()
        y = torch.empty(x2.shape[0], Wc.shape[0], dtype=torch.float32, device=x2.device)
        dy = torch.empty_like(y)
        ops.linear_fwd_dual(x2, act_in, Wc, bc, A.ACT_GELU, y, dy)
        ctx.save_for_backward(x2, Wc, dy)
        ctx.act_in = act_in
        return y

    @staticmethod
    def backward(ctx, g):
        x2, Wc, dydpre = ctx.saved_tensors
        dpre = (_f32c(g) * dydpre).contiguous()
        dW = torch.zeros_like(Wc)
        db = torch.zeros(Wc.shape[0], dtype=torch.float32, device=Wc.device)
        ops.linear_bwd_w(dpre, x2, ctx.act_in, dW, db)
        dx = ops.linear_bwd_x(dpre, Wc, x2 if ctx.act_in != A.ACT_NONE else None, ctx.act_in) if ctx.needs_input_grad[0] else None
        return dx, dW, db, None
