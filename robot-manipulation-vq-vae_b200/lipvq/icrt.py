"""The tokenizer's CALLER: the ICRT in-context transformer policy around ``LLFQVAE_V4`` (BASELINE C5).

Restates, with the reference's module tree (so ``state_dict`` keys match and reference checkpoints load strictly):

  ICLObservationGroupEncoder     robomimic/models/obs_nets.py:1120-1385   low-dim observation groups (flatten + concat, no
                                                                          parameters) + ``action_network`` on the prompt actions
  ICL_MIMO_Transformer           obs_nets.py:2338-2634                    embed x3 + timestep parameter + LayerNorm, 3*T interleave
  SelfAttention(Block)/GPT_Backbone  robomimic/models/transformers.py:80-440
  ObservationDecoder             obs_nets.py:714-771                      one nn.Linear per output key
  ICLTransformerGMMActorNetwork  robomimic/models/policy_nets.py:2375-2610  5-mode GMM head (tanh means, softplus + min_std)

The tokenizer is ours (hand-written CUDA behind ``lipvq.LLFQVAE_V4``); everything AROUND it is stock PyTorch -- nn.Linear,
nn.LayerNorm and ``scaled_dot_product_attention`` -- as the scope table allows (no custom kernel is owed for a 30-token
attention).  ``forward`` of the encoder calls ``action_network(prompt_actions)`` with the flattened ``[B*T, A]`` actions and
stashes ``_vq_vae_loss`` exactly as obs_nets.py:1337-1339 does.  Image observations, the FAST/CLIP, Mamba and spectral-norm
action branches are out of scope; ``bin_enabled`` selects ``AdaptiveBinActionEmbedding`` (lipvq/bin_action.py).
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Callable, Dict, Optional

import numpy as np
import torch
import torch.distributions as D
import torch.nn as nn
import torch.nn.functional as F


class SelfAttention(nn.Module):
    """transformers.py:80-225.  Same parameters (``nets.qkv`` without bias, ``nets.output``) and the ``mask`` buffer; the
    masked softmax(QK^T/sqrt(d))V runs through scaled_dot_product_attention."""

    def __init__(self, embed_dim, num_heads, context_length, causal=True, attn_dropout=0.1, output_dropout=0.1):
        super().__init__()
        assert embed_dim % num_heads == 0, f"num_heads: {num_heads} does not divide embed_dim: {embed_dim} exactly"
        self.embed_dim, self.num_heads, self.context_length = embed_dim, num_heads, context_length
        self.attn_dropout, self.output_dropout = attn_dropout, output_dropout
        self.nets = nn.ModuleDict()
        self.nets["qkv"] = nn.Linear(embed_dim, 3 * embed_dim, bias=False)
        self.nets["attn_dropout"] = nn.Dropout(attn_dropout)
        self.nets["output_dropout"] = nn.Dropout(output_dropout)
        self.nets["output"] = nn.Linear(embed_dim, embed_dim)
        mask = torch.tril(torch.ones(context_length, context_length)).view(1, 1, context_length, context_length)
        if causal is False:
            mask[:] = 1
        self.register_buffer("mask", mask)

    def forward(self, x):
        assert len(x.shape) == 3
        B, T, Dm = x.shape
        assert T <= self.context_length and Dm == self.embed_dim
        NH, DH = self.num_heads, Dm // self.num_heads
        q, k, v = torch.chunk(self.nets["qkv"](x), 3, dim=-1)
        q, k, v = (t.view(B, T, NH, DH).transpose(1, 2) for t in (q, k, v))
        y = F.scaled_dot_product_attention(q, k, v, attn_mask=self.mask[..., :T, :T] != 0,
                                           dropout_p=self.attn_dropout if self.training else 0.0)
        y = y.transpose(1, 2).contiguous().view(B, T, Dm)
        return self.nets["output_dropout"](self.nets["output"](y))


class SelfAttentionBlock(nn.Module):
    """transformers.py:228-318: pre-LN attention + 4x GELU MLP, residuals."""

    def __init__(self, embed_dim, num_heads, context_length, causal=True, attn_dropout=0.1, output_dropout=0.1,
                 activation=None):
        super().__init__()
        self.nets = nn.ModuleDict()
        self.nets["attention"] = SelfAttention(embed_dim, num_heads, context_length, causal, attn_dropout, output_dropout)
        self.nets["mlp"] = nn.Sequential(nn.Linear(embed_dim, 4 * embed_dim), activation if activation is not None else nn.GELU(),
                                         nn.Linear(4 * embed_dim, embed_dim), nn.Dropout(output_dropout))
        self.nets["ln1"] = nn.LayerNorm(embed_dim)
        self.nets["ln2"] = nn.LayerNorm(embed_dim)

    def forward(self, x):
        x = x + self.nets["attention"](self.nets["ln1"](x))
        return x + self.nets["mlp"](self.nets["ln2"](x))


class GPT_Backbone(nn.Module):
    """transformers.py:321-440 (activation 'gelu'; weights N(0, 0.02), biases 0, LayerNorm (1, 0))."""

    def __init__(self, embed_dim, context_length, causal=True, attn_dropout=0.1, block_output_dropout=0.1, num_layers=6,
                 num_heads=8, activation="gelu"):
        super().__init__()
        if activation != "gelu":
            raise NotImplementedError("only the 'gelu' activation of the shipped configs is restated")
        self.embed_dim, self.num_layers, self.num_heads, self.context_length = embed_dim, num_layers, num_heads, context_length
        self.activation = nn.GELU()
        self.nets = nn.ModuleDict()
        self.nets["transformer"] = nn.Sequential(*[
            SelfAttentionBlock(embed_dim, num_heads, context_length, causal, attn_dropout, block_output_dropout, self.activation)
            for _ in range(num_layers)])
        self.nets["output_ln"] = nn.LayerNorm(embed_dim)
        self.apply(self._init_weights)

    @staticmethod
    def _init_weights(module):
        if isinstance(module, (nn.Linear, nn.Embedding)):
            module.weight.data.normal_(mean=0.0, std=0.02)
            if isinstance(module, nn.Linear) and module.bias is not None:
                module.bias.data.zero_()
        elif isinstance(module, nn.LayerNorm):
            module.bias.data.zero_()
            module.weight.data.fill_(1.0)

    def forward(self, inputs):
        assert inputs.shape[1:] == (self.context_length, self.embed_dim), inputs.shape
        return self.nets["output_ln"](self.nets["transformer"](inputs))


class LowDimObservationEncoder(nn.Module):
    """obs_encoder_factory / ObservationEncoder for low-dim keys (obs_nets.py:60-250): flatten each key, concatenate in
    ``obs_shapes`` order; no parameters (``feature_activation=None`` at the ICL call site, obs_nets.py:2419)."""

    def __init__(self, obs_shapes: "OrderedDict[str, tuple]"):
        super().__init__()
        self.obs_shapes = OrderedDict(obs_shapes)

    def output_shape(self):
        return [int(sum(int(np.prod(s)) for s in self.obs_shapes.values()))]

    def forward(self, obs_dict: Dict[str, torch.Tensor]) -> torch.Tensor:
        return torch.cat([obs_dict[k].reshape(obs_dict[k].shape[0], -1) for k in self.obs_shapes], dim=-1)


class ObservationDecoder(nn.Module):
    """obs_nets.py:714-771."""

    def __init__(self, decode_shapes, input_feat_dim):
        super().__init__()
        self.obs_shapes = OrderedDict(decode_shapes)
        self.nets = nn.ModuleDict({k: nn.Linear(input_feat_dim, int(np.prod(s))) for k, s in self.obs_shapes.items()})

    def forward(self, feats):
        return {k: self.nets[k](feats).reshape(-1, *self.obs_shapes[k]) for k in self.obs_shapes}


def _default_action_network(action_input_shape: int, action_output_shape: int) -> nn.Module:
    from .modules import LLFQVAE_V4  # the B200-native tokenizer (CUDA only)
    return LLFQVAE_V4(feature_dim=action_input_shape, latent_dim=action_output_shape)


class ICLObservationGroupEncoder(nn.Module):
    """obs_nets.py:1120-1385, ``vq_vae_enabled`` / ``bin_enabled`` branches.  ``action_network_factory`` lets tests inject
    a stand-in tokenizer with the same (z_latent, loss) contract; the default is ``lipvq.LLFQVAE_V4``."""

    def __init__(self, observation_group_shapes, action_input_shape, bin_enabled=False, vq_vae_enabled=False,
                 action_network_factory: Optional[Callable[[int, int], nn.Module]] = None):
        super().__init__()
        assert isinstance(observation_group_shapes, OrderedDict)
        self.observation_group_shapes = observation_group_shapes
        self.nets = nn.ModuleDict({g: LowDimObservationEncoder(s) for g, s in observation_group_shapes.items()})
        action_output_shape = self.output_shape()[0]
        self.bin_enabled, self.vq_vae_enabled = bin_enabled, vq_vae_enabled
        if bin_enabled:
            from .bin_action import AdaptiveBinActionEmbedding
            self.action_network = AdaptiveBinActionEmbedding(action_dim=action_input_shape, output_dim=action_output_shape)
        elif vq_vae_enabled:
            make = action_network_factory or _default_action_network
            self.action_network = make(action_input_shape, action_output_shape)
        else:
            raise NotImplementedError("only the vq_vae_enabled and bin_enabled action branches are in scope (SURVEY 8f)")
        self._vq_vae_loss = None

    def output_shape(self):
        return [sum(self.nets[g].output_shape()[0] for g in self.observation_group_shapes)]

    def forward(self, **inputs):
        prompt_obs, prompt_actions = inputs["prompt"]["obs"], inputs["prompt"]["action"]
        assert set(self.observation_group_shapes.keys()).issubset(inputs)
        obs = torch.cat([self.nets[g](inputs[g]) for g in self.observation_group_shapes], dim=-1)
        context_obs = self.nets["obs"](prompt_obs)
        if self.vq_vae_enabled:
            # prompt_actions is [B*T, A] fp32 contiguous: the tokenizer's call contract
            context_actions, loss = self.action_network(prompt_actions)
            self._vq_vae_loss = loss
        else:
            context_actions = self.action_network(prompt_actions)
        return obs, context_obs, context_actions


def _join_bt(x):
    """TensorUtils.join_dimensions(x, 0, 1) on a nested dict (tensor_utils.py:1066-1069)."""
    if isinstance(x, dict):
        return {k: _join_bt(v) for k, v in x.items() if v is not None}
    return x.reshape(-1, *x.shape[2:])


class ICL_MIMO_Transformer(nn.Module):
    """obs_nets.py:2338-2634 with ``transformer_nn_parameter_for_timesteps`` or the nn.Embedding timestep table."""

    def __init__(self, input_obs_group_shapes, output_shapes, transformer_embed_dim, transformer_num_layers,
                 transformer_num_heads, transformer_context_length, transformer_causal=True, transformer_emb_dropout=0.1,
                 transformer_attn_dropout=0.1, transformer_block_output_dropout=0.1, transformer_activation="gelu",
                 transformer_bin_enabled=False, transformer_vq_vae_enabled=False,
                 transformer_nn_parameter_for_timesteps=False, action_input_shape=12, action_network_factory=None):
        super().__init__()
        assert isinstance(input_obs_group_shapes, OrderedDict) and isinstance(output_shapes, OrderedDict)
        self.input_obs_group_shapes, self.output_shapes = input_obs_group_shapes, output_shapes
        self.nets = nn.ModuleDict()
        self.params = nn.ParameterDict()
        self.nets["encoder"] = ICLObservationGroupEncoder(input_obs_group_shapes, action_input_shape=action_input_shape,
                                                          bin_enabled=transformer_bin_enabled,
                                                          vq_vae_enabled=transformer_vq_vae_enabled,
                                                          action_network_factory=action_network_factory)
        self.vq_vae_enabled = transformer_vq_vae_enabled
        if transformer_vq_vae_enabled:
            self.vq_vae_model = self.nets["encoder"].action_network  # alias, as obs_nets.py:2422-2424
        transformer_input_dim = self.nets["encoder"].output_shape()[0]
        self.nets["embed_encoder"] = nn.Linear(transformer_input_dim, transformer_embed_dim)
        if transformer_nn_parameter_for_timesteps:
            self.params["embed_timestep"] = nn.Parameter(torch.zeros(1, transformer_context_length, transformer_embed_dim))
        else:
            self.nets["embed_timestep"] = nn.Embedding(transformer_context_length, transformer_embed_dim)
        self.nets["embed_ln"] = nn.LayerNorm(transformer_embed_dim)
        self.nets["embed_drop"] = nn.Dropout(transformer_emb_dropout)
        self.nets["transformer"] = GPT_Backbone(embed_dim=transformer_embed_dim, num_layers=transformer_num_layers,
                                                num_heads=transformer_num_heads,
                                                context_length=transformer_context_length * 3,  # ctx_obs, ctx_act, obs
                                                causal=transformer_causal, attn_dropout=transformer_attn_dropout,
                                                block_output_dropout=transformer_block_output_dropout,
                                                activation=transformer_activation)
        self.nets["decoder"] = ObservationDecoder(decode_shapes=self.output_shapes, input_feat_dim=transformer_embed_dim)
        self.transformer_context_length = transformer_context_length
        self.transformer_embed_dim = transformer_embed_dim
        self.transformer_nn_parameter_for_timesteps = transformer_nn_parameter_for_timesteps
        self._vq_vae_loss = None

    def embed_timesteps(self, embeddings):
        if self.transformer_nn_parameter_for_timesteps:
            return self.params["embed_timestep"]
        t = torch.arange(0, embeddings.shape[1], device=embeddings.device).unsqueeze(0).repeat(embeddings.shape[0], 1)
        return self.nets["embed_timestep"](t.long())

    def input_embedding(self, inputs):
        e = self.nets["embed_encoder"](inputs)
        e = e + self.embed_timesteps(e)
        return self.nets["embed_drop"](self.nets["embed_ln"](e))

    def forward(self, **inputs):
        some = next(iter(inputs["obs"].values()))
        B, T = some.shape[:2]
        flat = _join_bt({k: v for k, v in inputs.items() if v is not None})
        obs, context_obs, context_actions = self.nets["encoder"](**flat)
        obs, context_obs = obs.reshape(B, T, -1), context_obs.reshape(B, T, -1)
        context_actions = context_actions.reshape(B, T, -1)
        if self.vq_vae_enabled:
            self._vq_vae_loss = self.nets["encoder"]._vq_vae_loss
        oe, ce, ae = self.input_embedding(obs), self.input_embedding(context_obs), self.input_embedding(context_actions)
        bs, _, Dm = oe.shape
        interleaved = torch.stack([ce, ae], dim=2).view(bs, -1, Dm)       # [ctx_obs_0, ctx_act_0, ctx_obs_1, ...]
        enc = self.nets["transformer"](torch.cat([interleaved, oe], dim=1))  # [bs, 3T, D]
        out = enc[:, -self.transformer_context_length:]
        dec = self.nets["decoder"](out.reshape(-1, Dm))
        dec = {k: v.reshape(bs, self.transformer_context_length, *v.shape[1:]) for k, v in dec.items()}
        dec["transformer_encoder_outputs"] = enc
        return dec


class ICLTransformerGMMActorNetwork(ICL_MIMO_Transformer):
    """policy_nets.py:2375-2610 (and its base :1590-1700): obs group 'obs' -> GMM over actions."""

    def __init__(self, obs_shapes, ac_dim, transformer_embed_dim, transformer_num_layers, transformer_num_heads,
                 transformer_context_length, transformer_causal=True, transformer_emb_dropout=0.1, transformer_attn_dropout=0.1,
                 transformer_block_output_dropout=0.1, transformer_activation="gelu",
                 transformer_nn_parameter_for_timesteps=False, transformer_bin_enabled=False, transformer_vq_vae_enabled=False,
                 num_modes=5, min_std=0.01, std_activation="softplus", low_noise_eval=True, use_tanh=False,
                 action_network_factory=None):
        assert std_activation in ("softplus", "exp")
        if use_tanh:
            raise NotImplementedError("tanh-wrapped GMM is not used by the shipped configs")
        self.ac_dim, self.num_modes, self.min_std = ac_dim, num_modes, min_std
        self.low_noise_eval, self.std_activation = low_noise_eval, std_activation
        super().__init__(input_obs_group_shapes=OrderedDict(obs=OrderedDict(obs_shapes)),
                         output_shapes=OrderedDict(mean=(num_modes, ac_dim), scale=(num_modes, ac_dim), logits=(num_modes,)),
                         transformer_embed_dim=transformer_embed_dim, transformer_num_layers=transformer_num_layers,
                         transformer_num_heads=transformer_num_heads, transformer_context_length=transformer_context_length,
                         transformer_causal=transformer_causal, transformer_emb_dropout=transformer_emb_dropout,
                         transformer_attn_dropout=transformer_attn_dropout,
                         transformer_block_output_dropout=transformer_block_output_dropout,
                         transformer_activation=transformer_activation, transformer_bin_enabled=transformer_bin_enabled,
                         transformer_vq_vae_enabled=transformer_vq_vae_enabled,
                         transformer_nn_parameter_for_timesteps=transformer_nn_parameter_for_timesteps,
                         action_input_shape=12, action_network_factory=action_network_factory)  # 12: obs_nets.py:2413 ("FIXME")

    def forward_train(self, obs_dict, context_obs, actions=None, goal_dict=None, low_noise_eval=None):
        out = ICL_MIMO_Transformer.forward(self, obs=obs_dict, prompt={"obs": context_obs, "action": actions})
        means, scales, logits = torch.tanh(out["mean"]), out["scale"], out["logits"]
        if low_noise_eval is None:
            low_noise_eval = self.low_noise_eval
        if low_noise_eval and not self.training:
            scales = torch.ones_like(means) * 1e-4
        else:
            scales = (F.softplus(scales) if self.std_activation == "softplus" else torch.exp(scales)) + self.min_std
        comp = D.Independent(D.Normal(loc=means, scale=scales), 1)
        return D.MixtureSameFamily(mixture_distribution=D.Categorical(logits=logits), component_distribution=comp)

    def forward(self, obs_dict, context_obs, actions=None, goal_dict=None):
        return self.forward_train(obs_dict, context_obs, actions, goal_dict).sample()
