"""CPU: the restated ICRT caller (lipvq/icrt.py) against the reference's own run (tests/golden/icrt_policy.npz, written by
oracle/gen_golden.py from policy_nets.py / obs_nets.py / transformers.py).  The transformer around the tokenizer is stock
PyTorch, so it runs here; the tokenizer itself is CUDA-only, and is replaced by a torch stand-in built from the ORACLE's
functions with the same parameter names and the same (z_latent, loss) contract -- the -m gpu twin of this test
(tests/test_gpu_icrt.py) runs the real lipvq.LLFQVAE_V4 in that slot."""
from collections import OrderedDict

import torch
import torch.nn as nn

import lipvq_oracle as O
from conftest import load_golden, sub

OBS = OrderedDict((("robot0_eef_pos", (3,)), ("robot0_eef_quat", (4,)), ("robot0_gripper_qpos", (2,)), ("lang_emb", (15,))))
KW = dict(ac_dim=12, transformer_embed_dim=64, transformer_num_layers=2, transformer_num_heads=4, transformer_context_length=10,
          transformer_vq_vae_enabled=True, transformer_nn_parameter_for_timesteps=True, num_modes=5, min_std=0.0001)


OracleTokenizer = O.T1Module


def _build():
    from lipvq import icrt
    g = load_golden("icrt_policy")
    net = icrt.ICLTransformerGMMActorNetwork(obs_shapes=OBS, action_network_factory=lambda a, d: OracleTokenizer(a, d), **KW)
    return g, net


def test_state_dict_keys_match_reference_and_load_strict():
    g, net = _build()
    sd = sub(g, "sd")
    mine = {k for k in net.state_dict() if not k.startswith("vq_vae_model.")}
    assert mine == set(sd), (sorted(mine - set(sd)), sorted(set(sd) - mine))
    # the alias the algo layer uses (icl.py:1437: self.nets["policy"].vq_vae_model) shares storage with the encoder's network
    assert net.vq_vae_model is net.nets["encoder"].action_network
    full = dict(sd)
    full.update({"vq_vae_model." + k[len("nets.encoder.action_network."):]: v for k, v in sd.items()
                 if k.startswith("nets.encoder.action_network.")})
    net.load_state_dict(full, strict=True)


def test_policy_forward_backward_vs_reference():
    g, net = _build()
    sd = sub(g, "sd")
    net.load_state_dict(sd, strict=False)
    net.eval()
    obs, ctx = sub(g, "obs"), sub(g, "ctx")
    d = net.forward_train(obs_dict=obs, context_obs=ctx, actions=g["ctx_actions"], low_noise_eval=False)
    base = d.component_distribution.base_dist
    assert torch.allclose(base.loc, g["means"], rtol=1e-5, atol=1e-6)
    assert torch.allclose(base.scale, g["scales"], rtol=1e-5, atol=1e-7)
    assert torch.allclose(d.mixture_distribution.logits, g["logits"], rtol=1e-5, atol=1e-6)
    assert torch.allclose(net._vq_vae_loss, g["vq_loss"], rtol=1e-6)
    logp = d.log_prob(g["train_actions"])
    assert logp.shape == (3, 10) and torch.allclose(logp, g["log_prob"], rtol=1e-5, atol=1e-4)
    loss = -logp.mean()
    assert torch.allclose(loss, g["action_loss"], rtol=1e-5)
    net.zero_grad()
    loss.backward()
    named = dict(net.named_parameters())
    for k, gr in sub(g, "grad").items():
        assert torch.allclose(named[k].grad, gr, rtol=1e-4, atol=1e-6), k
    # the policy loss cannot reach the tokenizer (z_latent is detached): the two-optimizer protocol relies on it
    assert all(p.grad is None for p in net.vq_vae_model.parameters())


def test_token_order_is_ctx_obs_ctx_act_pairs_then_obs():
    """obs_nets.py:2588-2598: [ctx_obs_0, ctx_act_0, ctx_obs_1, ctx_act_1, ..., obs_0..obs_{T-1}] -- 3*T tokens."""
    g, net = _build()
    net.load_state_dict(sub(g, "sd"), strict=False)
    net.eval()
    seen = {}
    net.nets["transformer"].register_forward_pre_hook(lambda m, a: seen.setdefault("x", a[0].detach()))
    obs, ctx = sub(g, "obs"), sub(g, "ctx")
    net.forward_train(obs_dict=obs, context_obs=ctx, actions=g["ctx_actions"], low_noise_eval=False)
    x = seen["x"]
    assert x.shape == (3, 30, 64)
    flat = lambda dct: torch.cat([dct[k].reshape(30, -1) for k in OBS], dim=-1)
    emb = lambda t: net.input_embedding(t.reshape(3, 10, -1))
    assert torch.allclose(x[:, 0:20:2], emb(flat(ctx)), atol=1e-6)
    assert torch.allclose(x[:, 1:20:2], emb(g["z_latent"]), atol=1e-6)
    assert torch.allclose(x[:, 20:], emb(flat(obs)), atol=1e-6)


def test_algo_step_protocol_and_checkpoint(tmp_path):
    """T1 protocol (icl.py:1437-1441,1466,1520-1522 + torch_utils.py:197-233): the tokenizer moves by ITS AdamW only, the
    policy by Adam on the clipped policy-loss gradient under the warm-up factor; a reference-layout checkpoint (weights
    from the reference's own state_dict) loads strictly and continues identically."""
    import copy
    from lipvq import icl_algo
    g, net = _build()
    net.load_state_dict(sub(g, "sd"), strict=False)
    net.train()
    for m in net.modules():
        if isinstance(m, nn.Dropout):
            m.p = 0.0  # deterministic comparison
    for blk in net.nets["transformer"].nets["transformer"]:
        blk.nets["attention"].attn_dropout = 0.0
    ref = copy.deepcopy(net)
    algo = icl_algo.ICLTransformerGMM(net, num_warmup_steps=4, max_grad_norm=0.05)
    obs, ctx = sub(g, "obs"), sub(g, "ctx")
    batch = {"obs": {k: torch.cat([ctx[k], obs[k]]) for k in obs}, "actions": torch.cat([g["ctx_actions"], g["train_actions"]])}
    info = algo.train_on_batch(batch)
    assert set(info) >= {"vq_vae_loss", "action_loss", "log_probs", "policy_grad_norms"}
    # manual protocol on the copy
    vq_opt = torch.optim.AdamW(ref.vq_vae_model.parameters(), lr=1e-3, weight_decay=1e-4)
    pol_opt = torch.optim.Adam(ref.parameters(), lr=1e-4)
    sched = torch.optim.lr_scheduler.LambdaLR(pol_opt, lambda s: s / 4 if s < 4 else 1.0)
    vq_opt.zero_grad()
    d = ref.forward_train(obs_dict=obs, context_obs=ctx, actions=g["ctx_actions"], low_noise_eval=False)
    ref._vq_vae_loss.backward()
    vq_opt.step()
    base = d.component_distribution.base_dist
    import torch.distributions as D
    last = D.MixtureSameFamily(D.Categorical(logits=d.mixture_distribution.logits[:, -1]),
                               D.Independent(D.Normal(base.loc[:, -1], base.scale[:, -1]), 1))
    loss = -last.log_prob(g["train_actions"][:, -1]).mean()
    pol_opt.zero_grad()  # clears the tokenizer grads too: the policy step never moves the tokenizer
    loss.backward()
    torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.05)
    pol_opt.step()
    sched.step()
    assert torch.allclose(info["action_loss"], loss.detach(), rtol=1e-6)
    for (k, a), (_, b) in zip(net.state_dict().items(), ref.state_dict().items()):
        assert torch.allclose(a, b, rtol=1e-6, atol=1e-8), k
    assert algo.optimizers["policy"].param_groups[0]["lr"] == pol_opt.param_groups[0]["lr"] == 1e-4 * 0.25
    # step 1 ran at factor 0 (LambdaLR starts at step 0): policy weights outside the tokenizer are unchanged, tokenizer moved
    sd0 = sub(g, "sd")
    assert torch.equal(net.state_dict()["nets.embed_encoder.weight"], sd0["nets.embed_encoder.weight"])
    assert not torch.equal(net.state_dict()["nets.encoder.action_network.to_latent.W"], sd0["nets.encoder.action_network.to_latent.W"])
    # checkpoint: reference layout, reference key names ("policy." + the reference module tree)
    path = str(tmp_path / "model_epoch_1.pth")
    algo.save_model(path, config={"algo_name": "icl"}, shape_meta={"ac_dim": 12})
    ck = torch.load(path, weights_only=False)
    assert set(ck) == {"model", "config", "algo_name", "env_metadata", "shape_metadata"}
    assert all(k.startswith("policy.") for k in ck["model"])
    ref_ckpt = {"policy." + k: v for k, v in sd0.items()}
    ref_ckpt.update({"policy.vq_vae_model." + k[len("nets.encoder.action_network."):]: v for k, v in sd0.items()
                     if k.startswith("nets.encoder.action_network.")})
    ref_ckpt.update({"vq_vae_enabled": True, "vqvae_config": {"num_subclusters": 1}})  # the fork's two non-tensor keys
    algo.deserialize(ref_ckpt)
    assert torch.equal(net.state_dict()["nets.embed_encoder.weight"], sd0["nets.embed_encoder.weight"])


def test_constant_with_warmup_matches_lambda():
    from lipvq.icl_algo import constant_with_warmup
    p = nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=1.0)
    s = constant_with_warmup(opt, 5)
    lrs = []
    for _ in range(8):
        lrs.append(opt.param_groups[0]["lr"])
        opt.step()
        s.step()
    assert lrs == [0.0, 0.2, 0.4, 0.6, 0.8, 1.0, 1.0, 1.0]
