"""The ICL algo step around the tokenizer, as a thin trainer.

Restates the optimisation protocol of robomimic/algo/icl.py for the two tokenizer families:

  T1  ``ICLTransformer_GMM`` (icl.py:1437-1526)   a dedicated AdamW(lr=1e-3, wd=1e-4) on ``policy.vq_vae_model``; per step
      vq_optimizer.zero_grad -> forward_train -> ``_vq_vae_loss.backward(); vq_optimizer.step()`` -> policy loss through
      ``TorchUtils.backprop_for_loss`` (torch_utils.py:197-233: zero_grad, backward, clip_grad_norm_, sum of squared grad
      norms, step).  The batch's first half is the in-context prompt, the second half the queries (icl.py:1451-1459).
  T2  ``ICLTransformer._train_step`` (icl.py:1157-1213)   combined = 0.1 * vqvae_loss + 1.0 * action_loss, ONE backward,
      clip-then-step the vqvae optimizer, then the policy optimizer (``cotrain_step`` below).

LR schedule ``constant_with_warmup`` (torch_utils.py:166-172 -> diffusers get_constant_schedule_with_warmup: factor
step / max(1, warmup) below ``num_warmup_steps``, then 1).  Checkpoints use the reference's layout
(utils/train_utils.py:1186-1235: {"model": nets.state_dict() (+ the two non-tensor keys of icl.py:1379-1391),
"config", "algo_name", "env_metadata", "shape_metadata"}) so files written by either side load in the other.

Multi-GPU: one process per GPU, gradients summed with one flat all_reduce per optimizer (the policy's gradients; the
tokenizer's gradients) -- plain NCCL: the transformer is the caller, not the hot path.
"""
from __future__ import annotations

import json
from collections import OrderedDict
from typing import Dict, Optional

import torch
import torch.distributed as dist
import torch.nn as nn


def constant_with_warmup(optimizer: torch.optim.Optimizer, num_warmup_steps: int) -> torch.optim.lr_scheduler.LambdaLR:
    """diffusers.optimization.get_constant_schedule_with_warmup, restated (diffusers is not installed here)."""
    def lr_lambda(step: int) -> float:
        if step < num_warmup_steps:
            return float(step) / float(max(1.0, num_warmup_steps))
        return 1.0
    return torch.optim.lr_scheduler.LambdaLR(optimizer, lr_lambda, last_epoch=-1)


def _allreduce_grads(params, group=None) -> None:
    """sum-all-reduce the gradients of ``params`` as one flat buffer, then average."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    gs = [p.grad for p in params if p.grad is not None]
    if not gs:
        return
    flat = torch._utils._flatten_dense_tensors(gs)
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    flat.div_(dist.get_world_size(group))
    for g, f in zip(gs, torch._utils._unflatten_dense_tensors(flat, gs)):
        g.copy_(f)


def backprop_for_loss(net: nn.Module, optim: torch.optim.Optimizer, loss: torch.Tensor, max_grad_norm=None,
                      group=None) -> float:
    """torch_utils.py:197-233 (the returned value is the SUM of squared per-parameter gradient norms, as there)."""
    optim.zero_grad()
    loss.backward()
    _allreduce_grads(list(net.parameters()), group)
    if max_grad_norm is not None:
        torch.nn.utils.clip_grad_norm_(net.parameters(), max_grad_norm)
    sq = [p.grad.detach().float().pow(2).sum() for p in net.parameters() if p.grad is not None]
    grad_norms = float(torch.stack(sq).sum().item()) if sq else 0.0
    optim.step()
    return grad_norms


def cotrain_step(losses: Dict[str, torch.Tensor], nets: nn.ModuleDict, optimizers: Dict[str, torch.optim.Optimizer],
                 max_grad_norm: float = 100.0, lr_schedulers: Optional[dict] = None, group=None) -> "OrderedDict[str, float]":
    """icl.py:1157-1213 (T2): one backward of 0.1 * vqvae_loss + 1.0 * action_loss, then the two clipped steps."""
    info = OrderedDict()
    if "vqvae" in nets and "vqvae_loss" in losses:
        combined = 0.1 * losses["vqvae_loss"] + 1.0 * losses["action_loss"]
        combined.backward()
        for name in ("vqvae", "policy"):
            _allreduce_grads(list(nets[name].parameters()), group)
            info[name + "_grad_norm"] = float(torch.nn.utils.clip_grad_norm_(nets[name].parameters(), max_grad_norm))
            optimizers[name].step()
            optimizers[name].zero_grad()
    else:
        info["policy_grad_norm"] = backprop_for_loss(nets["policy"], optimizers["policy"], losses["action_loss"],
                                                     max_grad_norm, group)
    for sch in (lr_schedulers or {}).values():
        if sch is not None:
            sch.step()
    return info


class ICLTransformerGMM:
    """T1 algo: in-context GMM policy + LipVQ-VAE tokenizer with its own optimizer (icl.py ``ICLTransformer_GMM``)."""

    algo_name = "icl"

    def __init__(self, policy: nn.Module, lr: float = 1e-4, l2: float = 0.0, vq_lr: float = 1e-3, vq_weight_decay: float = 1e-4,
                 max_grad_norm: Optional[float] = 100.0, num_warmup_steps: int = 10000, supervise_all_steps: bool = False,
                 bf16: bool = False, process_group=None):
        self.nets = nn.ModuleDict({"policy": policy})
        self.device = next(policy.parameters()).device
        self.vq_vae_enabled = bool(getattr(policy, "vq_vae_enabled", False))
        self.optimizers = {"policy": torch.optim.Adam(policy.parameters(), lr=lr, weight_decay=l2)}  # icl_config.py:30-36
        self.lr_schedulers = {"policy": constant_with_warmup(self.optimizers["policy"], num_warmup_steps)}
        if self.vq_vae_enabled:
            self.vq_vae_model = policy.vq_vae_model
            self.vq_optimizer = torch.optim.AdamW(self.vq_vae_model.parameters(), lr=vq_lr, weight_decay=vq_weight_decay)
        self.max_grad_norm, self.supervise_all_steps, self.bf16, self.group = max_grad_norm, supervise_all_steps, bf16, process_group

    # -- one optimisation step (icl.py:1443-1526 + algo.train_on_batch) ---------------------------------------------------
    def train_on_batch(self, batch: Dict) -> "OrderedDict[str, float]":
        policy = self.nets["policy"]
        some = next(iter(batch["obs"].values()))
        mid = some.shape[0] // 2
        context_obs = {k: v[:mid] for k, v in batch["obs"].items()}
        train_obs = {k: v[mid:] for k, v in batch["obs"].items()}
        context_actions, train_actions = batch["actions"][:mid], batch["actions"][mid:]
        if self.vq_vae_enabled:
            self.vq_optimizer.zero_grad()
        # bf16: the transformer (nn.Linear / LayerNorm / attention) under autocast; the tokenizer's kernels take and return
        # fp32 (autocast does not touch a custom autograd.Function's fp32 inputs), distributions and loss in fp32
        with torch.autocast(device_type=self.device.type, dtype=torch.bfloat16, enabled=self.bf16):
            dists = policy.forward_train(obs_dict=train_obs, context_obs=context_obs, actions=context_actions,
                                         low_noise_eval=False)
        info = OrderedDict()
        if self.vq_vae_enabled:
            vq_loss = policy._vq_vae_loss
            info["vq_vae_loss"] = vq_loss.detach()
        assert len(dists.batch_shape) == 2  # [B, T]
        comp = dists.component_distribution.base_dist
        loc, scale, logits = comp.loc.float(), comp.scale.float(), dists.mixture_distribution.logits.float()
        if not self.supervise_all_steps:  # "dists = dists[:, -1]" (icl.py:1486-1500)
            loc, scale, logits = loc[:, -1], scale[:, -1], logits[:, -1]
        import torch.distributions as D
        dists = D.MixtureSameFamily(D.Categorical(logits=logits), D.Independent(D.Normal(loc=loc, scale=scale), 1))
        log_probs = dists.log_prob(train_actions if self.supervise_all_steps or train_actions.dim() == 2 else train_actions[:, -1])
        action_loss = -log_probs.mean()
        if self.vq_vae_enabled:  # icl.py:1520-1522: the tokenizer's own step comes first
            vq_loss.backward()
            _allreduce_grads(list(self.vq_vae_model.parameters()), self.group)
            self.vq_optimizer.step()
        info["policy_grad_norms"] = backprop_for_loss(policy, self.optimizers["policy"], action_loss, self.max_grad_norm,
                                                      self.group)
        for sch in self.lr_schedulers.values():
            sch.step()
        info["action_loss"] = action_loss.detach()
        info["log_probs"] = -action_loss.detach()
        return info

    # -- checkpoints in the reference's layout ------------------------------------------------------------------------------
    def serialize(self) -> dict:
        return self.nets.state_dict()  # algo.py:335-339 (the T1 algo adds nothing; the T2 algo adds the two keys below)

    def deserialize(self, model_dict: dict) -> None:
        cleaned = {k: v for k, v in model_dict.items() if k not in ("vq_vae_enabled", "vqvae_config")}  # algo.py:341-363
        self.nets.load_state_dict(cleaned, strict=True)

    def save_model(self, ckpt_path: str, config: Optional[dict] = None, env_meta: Optional[dict] = None,
                   shape_meta: Optional[dict] = None) -> None:
        torch.save(dict(model=self.serialize(), config=json.dumps(config or {}, indent=4), algo_name=self.algo_name,
                        env_metadata=dict(env_meta or {}), shape_metadata=dict(shape_meta or {})), ckpt_path)

    def load_model(self, ckpt_path: str) -> dict:
        ckpt = torch.load(ckpt_path, map_location=self.device, weights_only=False)
        self.deserialize(ckpt["model"])
        return ckpt


def serialize_with_vqvae(nets: nn.ModuleDict) -> dict:
    """ICLTransformer.serialize of the fork (icl.py:1379-1391): nets.state_dict() + "vq_vae_enabled" + "vqvae_config"."""
    sd = nets.state_dict()
    if "vqvae" in nets:
        v = nets["vqvae"]
        sd["vq_vae_enabled"] = True
        sd["vqvae_config"] = {"num_subclusters": v.num_subclusters, "num_clusters": v.num_clusters, "embed_dim": v.embed_dim,
                              "beta": v.beta}
    return sd
