"""-m gpu: the callers and data formats either side of the tokenizer, through liblipvq.so, against the
reference-generated fixtures: action-chunk windows (dataset.py:499-632), AdaptiveBinActionEmbedding (bin_action/backbone.py),
the ICRT policy with the REAL lipvq.LLFQVAE_V4 in the action_network slot (policy_nets.py / obs_nets.py) and the T1 / T2
optimisation protocols of algo/icl.py."""
from collections import OrderedDict

import numpy as np
import pytest
import torch

import lipvq_oracle as O
from conftest import load_golden, sub
from test_icrt_cpu import KW, OBS

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_window_gather_bit_exact_vs_reference_get_item():
    import lipvq as L
    g = load_golden("dataset_windows")
    demos_all = [g[f"demo.demo_{i}"].numpy() for i in range(len(g["lens"]))]
    for ci in range(4):
        T, fs, pad_seq, pad_fs, norm = (int(v) for v in g[f"c{ci}.cfg"])
        demos = [demos_all[int(d)] for d in g[f"c{ci}.demos"]]
        stats = {"offset": g["stats_offset"].numpy(), "scale": g["stats_scale"].numpy()} if norm else None
        ds = L.ActionSequenceDataset(demos, seq_length=T, frame_stack=fs, pad_seq_length=bool(pad_seq),
                                     pad_frame_stack=bool(pad_fs), action_normalization_stats=stats, device=DEV)
        ref = g[f"c{ci}.actions"]
        assert len(ds) == ref.shape[0]
        idx = torch.randperm(len(ds), generator=torch.Generator().manual_seed(ci))
        acts, mask = ds.get_batch(idx, want_mask=True)
        assert torch.equal(acts.cpu(), ref[idx]) and torch.equal(mask.cpu(), g[f"c{ci}.pad_mask"][idx])
    # full-size property check: 4096 chunks of T=16 from long demos, every window row is a row of its demo, clamped
    rng = np.random.RandomState(3)
    demos = [rng.rand(int(n), 12).astype(np.float32) for n in rng.randint(50, 400, size=64)]
    ds = L.ActionSequenceDataset(demos, seq_length=16, frame_stack=1, device=DEV)
    idx = torch.randint(0, len(ds), (4096,), generator=torch.Generator().manual_seed(0))
    acts = ds.get_batch(idx)
    ref, _ = O.sequence_windows(demos, idx.tolist(), 16, 1, True, True)
    assert torch.equal(acts.cpu(), torch.from_numpy(ref))


def test_bin_action_embedding_vs_reference_fixture():
    import lipvq as L
    g = load_golden("icrt_policy")
    m = L.AdaptiveBinActionEmbedding(action_dim=12, output_dim=24, embedding_dim=8)
    m.load_state_dict(sub(g, "bin_sd"), strict=True)
    m.to(DEV)
    a = g["ctx_actions"].reshape(-1, 12).to(DEV)
    out1 = m(a)
    assert torch.allclose(out1.detach().cpu(), g["bin_out1"], rtol=1e-4, atol=1e-6)
    out2 = m((a * 0.5).contiguous())
    assert torch.allclose(out2.detach().cpu(), g["bin_out2"], rtol=1e-4, atol=1e-6)
    post = sub(g, "bin_post")
    assert torch.equal(m.running_min.cpu(), post["running_min"]) and torch.equal(m.running_max.cpu(), post["running_max"])
    assert torch.equal(m.discretize(a).cpu(), g["bin_idx"])
    # gradients against torch autograd of the oracle's forward (same statistics: no further update)
    m._update_enabled = False
    m.zero_grad()
    m(a).square().sum().backward()
    p = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in m.state_dict().items() if k not in ("running_min", "running_max")}
    ref, _, _, _ = O.bin_embedding_forward(p, a.cpu(), m.running_min.cpu(), m.running_max.cpu(), update=False)
    ref.square().sum().backward()
    for k, v in m.named_parameters():
        assert torch.allclose(v.grad.cpu(), p[k].grad, rtol=1e-3, atol=1e-6), k


def _policy(g, **kw):
    from lipvq import icrt
    net = icrt.ICLTransformerGMMActorNetwork(obs_shapes=OBS, **{**KW, **kw})
    return net


def test_icrt_policy_with_real_tokenizer_vs_reference():
    """The reference's ICLTransformerGMMActorNetwork run (fixture) vs the restated caller with lipvq.LLFQVAE_V4 (CUDA
    kernels) in the action_network slot: strict load of the reference state_dict, tokenizer output bit-exact, GMM
    parameters / log-likelihood / policy gradients within 1e-4."""
    g = load_golden("icrt_policy")
    net = _policy(g)
    sd = dict(sub(g, "sd"))
    sd.update({"vq_vae_model." + k[len("nets.encoder.action_network."):]: v for k, v in sd.items()
               if k.startswith("nets.encoder.action_network.")})
    net.load_state_dict(sd, strict=True)
    net.to(DEV).eval()
    dev = lambda d: {k: v.to(DEV) for k, v in d.items()}
    d = net.forward_train(obs_dict=dev(sub(g, "obs")), context_obs=dev(sub(g, "ctx")), actions=g["ctx_actions"].to(DEV),
                          low_noise_eval=False)
    base = d.component_distribution.base_dist
    assert torch.allclose(base.loc.cpu(), g["means"], rtol=1e-4, atol=1e-5)
    assert torch.allclose(base.scale.cpu(), g["scales"], rtol=1e-4, atol=1e-6)
    assert torch.allclose(d.mixture_distribution.logits.cpu(), g["logits"], rtol=1e-4, atol=1e-5)
    assert abs(net._vq_vae_loss.item() - g["vq_loss"].item()) <= 1e-4 * abs(g["vq_loss"].item())
    z_latent, _ = net.vq_vae_model(g["ctx_actions"].reshape(30, 12).to(DEV))
    assert z_latent.grad_fn is None and torch.equal(z_latent.cpu(), g["z_latent"])  # gathered codewords: bit-exact
    logp = d.log_prob(g["train_actions"].to(DEV))
    assert torch.allclose(logp.cpu(), g["log_prob"], rtol=1e-4, atol=1e-3)
    net.zero_grad()
    (-logp.mean()).backward()
    named = dict(net.named_parameters())
    for k, gr in sub(g, "grad").items():
        assert torch.allclose(named[k].grad.cpu(), gr, rtol=1e-3, atol=1e-6), k
    assert all(p.grad is None for p in net.vq_vae_model.parameters())


@pytest.mark.parametrize("bf16", [False, True])
def test_icl_algo_step_on_gpu(bf16, tmp_path):
    """T1 protocol end to end on the device (tokenizer = CUDA kernels, transformer = torch, optionally under bf16
    autocast): losses finite and decreasing on a fixed batch, tokenizer indices unaffected by autocast, checkpoint
    round trip continues bit-identically."""
    from lipvq import icl_algo
    g = load_golden("icrt_policy")
    torch.manual_seed(0)
    net = _policy(g, transformer_emb_dropout=0.0, transformer_attn_dropout=0.0, transformer_block_output_dropout=0.0)
    net.load_state_dict(sub(g, "sd"), strict=False)
    net.to(DEV).train()
    algo = icl_algo.ICLTransformerGMM(net, num_warmup_steps=2, lr=1e-3, bf16=bf16)
    obs, ctx = sub(g, "obs"), sub(g, "ctx")
    batch = {"obs": {k: torch.cat([ctx[k], obs[k]]).to(DEV) for k in obs},
             "actions": torch.cat([g["ctx_actions"], g["train_actions"]]).to(DEV)}
    infos = [algo.train_on_batch(batch) for _ in range(12)]
    al = [float(i["action_loss"]) for i in infos]
    vl = [float(i["vq_vae_loss"]) for i in infos]
    assert all(np.isfinite(al)) and all(np.isfinite(vl))
    assert al[-1] < al[1] and vl[-1] < vl[0]
    if not bf16:
        assert abs(vl[0] - g["vq_loss"].item()) <= 1e-4 * abs(g["vq_loss"].item())
    path = str(tmp_path / "ck.pth")
    algo.save_model(path)
    net2 = _policy(g, transformer_emb_dropout=0.0, transformer_attn_dropout=0.0, transformer_block_output_dropout=0.0).to(DEV)
    algo2 = icl_algo.ICLTransformerGMM(net2, bf16=bf16)
    algo2.load_model(path)
    for (k, a), (_, b) in zip(net.state_dict().items(), net2.state_dict().items()):
        assert torch.equal(a, b), k


def test_t2_cotrain_step_protocol():
    """icl.py:1157-1213: one backward of 0.1 * vqvae_loss + 1.0 * action_loss, clip 100, vqvae step then policy step --
    with lipvq.HierarchicalVQVAE (CUDA kernels) as nets['vqvae'] and a linear policy head on its quantised output."""
    import lipvq as L
    from lipvq import icl_algo
    torch.manual_seed(0)
    vq = L.HierarchicalVQVAE(7, embed_dim=32, num_subclusters=16, num_clusters=8, num_stages=1, num_layers_per_stage=2,
                             dropout=0.0, kmeans_init=False).to(DEV)
    vq.codebooks_initialized.fill_(True)
    policy = torch.nn.Linear(32, 7).to(DEV)
    nets = torch.nn.ModuleDict({"vqvae": vq, "policy": policy})
    opts = {"vqvae": torch.optim.AdamW(vq.parameters(), lr=1e-4, weight_decay=1e-4),
            "policy": torch.optim.Adam(policy.parameters(), lr=1e-4)}
    actions = (torch.rand(8, 16, 7, generator=torch.Generator().manual_seed(1)) * 2 - 1).to(DEV)
    w0 = {k: v.clone() for k, v in nets.state_dict().items()}
    out = vq(actions, training=True)
    losses = dict(vq.compute_vqvae_loss(out, actions, lambda_rec=0.002))
    losses["action_loss"] = torch.nn.functional.mse_loss(policy(out["quantized_z"]), actions)
    info = icl_algo.cotrain_step(losses, nets, opts, max_grad_norm=100.0)
    assert set(info) == {"vqvae_grad_norm", "policy_grad_norm"} and all(np.isfinite(list(info.values())))
    sd = nets.state_dict()
    assert not torch.equal(sd["policy.weight"], w0["policy.weight"])
    assert not torch.equal(sd["vqvae.encoder.input_projection.weight"], w0["vqvae.encoder.input_projection.weight"])
    assert all(p.grad is None or float(p.grad.abs().sum()) == 0.0 for p in nets.parameters())  # zero_grad after each step
    ck = icl_algo.serialize_with_vqvae(nets)
    assert ck["vq_vae_enabled"] is True and set(ck["vqvae_config"]) == {"num_subclusters", "num_clusters", "embed_dim", "beta"}
