"""Generate tests/golden/*.npz by running the REFERENCE's own modules (imported by file path
from /root/reference -- the package __init__ chain needs mamba_ssm/clip/h5py which are absent).

Run in the build container only (the GPU box has no /root/reference):
    python oracle/gen_golden.py
The fixtures are committed; tests/test_oracle_golden.py pins oracle/lipvq_oracle.py to them and
the -m gpu tests pin the CUDA path to the same files.
"""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np
import torch

REF = os.environ.get("LIPVQ_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def _load(name: str, rel: str):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _np(d):
    return {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in d.items()}


def _save(name: str, **arrs):
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **_np(arrs))
    print(f"wrote {path}  ({os.path.getsize(path) / 1024:.0f} KiB)")


def _actions(n, f, seed=123):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(n, f, generator=g) * 2 - 1


def gen_t1(v5, tag, F_, D, K, N, steps):
    torch.manual_seed(0)
    model = v5.LLFQVAE_V4(F_, D, K)
    x = _actions(N, F_)
    sd0 = {k: v.clone() for k, v in model.state_dict().items()}
    # forward pieces through the reference's own submodules
    h = model.encoder(x)
    z_e = model.to_latent(h)
    z_q, idx = model.quantizer(z_e)
    z_latent, loss = model(x)
    model.zero_grad()
    loss.backward()
    grads = {"grad." + k: p.grad.clone() for k, p in model.named_parameters()}
    # AdamW trajectory with the reference protocol (icl.py:1437-1441, 1466, 1520-1522)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=1e-4)
    losses = []
    for _ in range(steps):
        opt.zero_grad()
        _, l = model(x)
        l.backward()
        opt.step()
        losses.append(l.detach().clone())
    sdT = {"final." + k: v.clone() for k, v in model.state_dict().items()}
    _save(tag, x=x, z_e=z_e, indices=idx, z_latent=z_latent, loss=loss, traj_losses=torch.stack(losses),
          **{"sd." + k: v for k, v in sd0.items()}, **grads, **sdT)


def gen_t1_ema(v5, hq, tag, F_, D, K, N, steps):
    """T1 body with the EMA quantizer of hq_vae.py:224-281 swapped in (train mode)."""
    torch.manual_seed(0)
    model = v5.LLFQVAE_V4(F_, D, K)
    model.quantizer = hq.LFQQuantizerEMA(K, D)
    model.train()
    x = _actions(N, F_)
    sd0 = {k: v.clone() for k, v in model.state_dict().items()}
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=1e-4)
    losses = []
    for _ in range(steps):
        opt.zero_grad()
        _, l = model(x)
        l.backward()
        opt.step()
        losses.append(l.detach().clone())
    sdT = {"final." + k: v.clone() for k, v in model.state_dict().items()}
    _save(tag, x=x, traj_losses=torch.stack(losses), **{"sd." + k: v for k, v in sd0.items()}, **sdT)


def gen_quantizers(hq, tag, N, K, D):
    torch.manual_seed(1)
    z = torch.rand(N, D)
    q = hq.LFQQuantizerEMA(K, D)
    q.train()
    pre = {k: v.clone() for k, v in q.state_dict().items()}
    z_q, idx = q(z)
    post = {k: v.clone() for k, v in q.state_dict().items()}
    torch.manual_seed(2)
    ql = hq.LFQQuantizerEMA_KMeans(K, D, dead_threshold=0)
    ql.initialized = True  # k-means init is host-side sklearn: excluded
    prel = {k: v.clone() for k, v in ql.state_dict().items()}
    z_ql, idxl = ql(z)
    postl = {k: v.clone() for k, v in ql.state_dict().items()}
    _save(tag, z=z, z_q=z_q, indices=idx, **{"pre." + k: v for k, v in pre.items()},
          **{"post." + k: v for k, v in post.items()}, lap_z_q=z_ql, lap_indices=idxl,
          **{"lap_pre." + k: v for k, v in prel.items()}, **{"lap_post." + k: v for k, v in postl.items()})


def gen_t4(bb, tag, F_, D, K, N):
    torch.manual_seed(0)
    model = bb.VQVAE(F_, D, num_embeddings=K)
    x = _actions(N, F_)
    sd0 = {k: v.clone() for k, v in model.state_dict().items()}
    z_e = model.encoder(x)
    d = (z_e.unsqueeze(1) - model.embedding.weight).pow(2).sum(-1)
    idx = torch.argmin(d, dim=1)
    z_latent, loss = model(x)
    model.zero_grad()
    loss.backward()
    grads = {"grad." + k: p.grad.clone() for k, p in model.named_parameters()}
    _save(tag, x=x, z_e=z_e, indices=idx, z_latent=z_latent, loss=loss,
          **{"sd." + k: v for k, v in sd0.items()}, **grads)


def gen_t2(hv, tag, A, E, Kz, Kq, stages, layers, B, T):
    torch.manual_seed(0)
    model = hv.HierarchicalVQVAE(A, embed_dim=E, num_subclusters=Kz, num_clusters=Kq, num_stages=stages,
                                 num_layers_per_stage=layers, dropout=0.0, replace_threshold_z=0.0,
                                 replace_threshold_q=0.0, kmeans_init=False)
    model.codebooks_initialized.fill_(True)  # skip RNG/k-means init; constructor codebooks are used
    g = torch.Generator().manual_seed(123)
    actions = torch.rand(B, T, A, generator=g) * 2 - 1
    sd0 = {k: v.clone() for k, v in model.state_dict().items()}
    out = model(actions, training=True)
    losses = model.compute_vqvae_loss(out, actions, lambda_rec=0.002)
    model.zero_grad()
    losses["vqvae_loss"].backward()
    grads = {"grad." + k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}
    post = {"post." + k: model.state_dict()[k].clone() for k in
            ("codebook_z", "codebook_q", "cluster_size_z", "cluster_size_q", "embed_avg_z", "embed_avg_q")}
    usage = model.get_codebook_usage()
    _save(tag, actions=actions, **{"out." + k: v for k, v in out.items()},
          **{"loss." + k: v for k, v in losses.items()}, **{"sd." + k: v for k, v in sd0.items()}, **grads, **post,
          usage_z_used=usage["z_used"], usage_q_used=usage["q_used"])


def _blobs(n_super, n_sub, per, dim, seed):
    """Well-separated hierarchical blobs: n_super super-clusters of n_sub sub-blobs each.  k-means with k = n_super or
    k = n_super * n_sub then has one unambiguous optimum, so the fixture does not depend on sklearn's thread count."""
    g = torch.Generator().manual_seed(seed)
    sup = torch.randn(n_super, dim, generator=g) * 4.0
    pts = []
    for s in range(n_super):
        for _ in range(n_sub):
            c = sup[s] + torch.randn(dim, generator=g) * 0.5
            pts.append(c + torch.randn(per, dim, generator=g) * 0.02)
    x = torch.cat(pts)
    return x[torch.randperm(x.shape[0], generator=g)]


def gen_t3_hier(hq, tag):
    """HierarchicalLFQHVQVAE.forward (hq_vae.py:289-415) run by the reference itself: 11-key dict, all parameter
    gradients, EMA / usage buffers of both quantizers after the step (k-means skipped, dead-code replacement off)."""
    torch.manual_seed(4)
    m = hq.HierarchicalLFQHVQVAE(7, 32, 16, num_z_codes=64, num_q_codes=32)
    for qz in (m.z_quantizer, m.q_quantizer):
        qz.initialized, qz.dead_threshold = True, 0
    with torch.no_grad():
        m.z_quantizer.codebook.copy_(torch.rand(64, 32))
        m.q_quantizer.codebook.copy_(torch.rand(32, 16))
    sd0 = {k: v.clone() for k, v in m.state_dict().items()}
    x = torch.rand(200, 7, generator=torch.Generator().manual_seed(123)) * 2 - 1
    out = m(x)
    m.zero_grad()
    out["loss"].backward()
    grads = {"grad." + k: p.grad.clone() for k, p in m.named_parameters() if p.grad is not None}
    post = {"post." + k: v.clone() for k, v in m.state_dict().items()}
    _save(tag, x=x, **{"out." + k: v for k, v in out.items()}, **{"sd." + k: v for k, v in sd0.items()}, **grads, **post)


def gen_kmeans(hq, hv, tag):
    """The two k-means initialisers run by the reference: HierarchicalVQVAE.initialize_codebooks_with_current_batch
    (heirarchical_vqvae.py:321-419, sklearn random_state=42) and LFQQuantizerEMA_KMeans.kmeans_init (hq_vae.py:65-78,
    numpy global RNG) on blob data whose optimum is unambiguous."""
    emb = _blobs(8, 2, 16, 16, 7).view(16, 16, 16)  # [B=16, T=16, E=16]: 8 super-clusters x 2 sub-blobs x 16 points
    torch.manual_seed(0)
    m = hv.HierarchicalVQVAE(7, embed_dim=16, num_subclusters=16, num_clusters=8, num_stages=1, num_layers_per_stage=1,
                             dropout=0.0, kmeans_init=True)
    np.random.seed(0)
    m.initialize_codebooks_with_current_batch(emb)
    t2 = {"t2." + k: m.state_dict()[k].clone() for k in
          ("codebook_z", "codebook_q", "cluster_size_z", "cluster_size_q", "embed_avg_z", "embed_avg_q",
           "codebooks_initialized")}
    z = torch.sigmoid(_blobs(8, 1, 32, 8, 9))  # [256, 8] latents in (0,1), 8 blobs
    torch.manual_seed(1)
    q = hq.LFQQuantizerEMA_KMeans(8, 8, dead_threshold=0)
    torch.manual_seed(5)
    np.random.seed(5)
    q.kmeans_init(z)
    _save(tag, emb=emb, z=z, **t2, hq_codebook=q.codebook.data.clone(), hq_ema_codebook=q.ema_codebook.data.clone())


def _stub_missing_packages():
    """The reference's policy files import packages that are absent here (mamba_ssm, clip, h5py, matplotlib, ...) at
    module level although the low-dim / vq_vae path never calls them: give them empty stand-ins so obs_nets.py and
    policy_nets.py can be imported from /root/reference unmodified."""
    import types

    class _Any(types.ModuleType):
        def __getattr__(self, k):
            if k.startswith("__"):
                raise AttributeError(k)
            return object

    for name in ("termcolor", "mamba_ssm", "clip", "h5py", "tensorboardX", "matplotlib", "matplotlib.pyplot", "matplotlib.cm",
                 "imageio", "cv2"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = _Any(name)
    sys.modules["termcolor"].colored = lambda s, *a, **k: s
    if REF not in sys.path:
        sys.path.insert(0, REF)


ICRT_OBS = (("robot0_eef_pos", (3,)), ("robot0_eef_quat", (4,)), ("robot0_gripper_qpos", (2,)), ("lang_emb", (15,)))
ICRT_KW = dict(ac_dim=12, transformer_embed_dim=64, transformer_num_layers=2, transformer_num_heads=4,
               transformer_context_length=10, transformer_vq_vae_enabled=True, transformer_nn_parameter_for_timesteps=True,
               num_modes=5, min_std=0.0001)


def gen_icrt(tag):
    """The tokenizer's caller run by the reference itself: ICLTransformerGMMActorNetwork (policy_nets.py:2375-2610) over
    ICL_MIMO_Transformer / ICLObservationGroupEncoder / GPT_Backbone with the vq_vae action branch, low-dim observations,
    eval mode (dropout off).  Saves weights, inputs, the tokenizer's output, GMM parameters, log-likelihood and the
    gradients of the policy loss, plus AdaptiveBinActionEmbedding (bin_action/backbone.py) on the same actions."""
    _stub_missing_packages()
    from collections import OrderedDict
    import robomimic.utils.obs_utils as ObsUtils
    import robomimic.models.policy_nets as PN
    from robomimic.models.bin_action.backbone import AdaptiveBinActionEmbedding
    obs_shapes = OrderedDict(ICRT_OBS)
    ObsUtils.initialize_obs_utils_with_obs_specs({"obs": {"low_dim": list(obs_shapes.keys())}})
    enc_kwargs = {"low_dim": {"core_class": None, "core_kwargs": {}, "obs_randomizer_class": None, "obs_randomizer_kwargs": {}}}
    torch.manual_seed(0)
    net = PN.ICLTransformerGMMActorNetwork(obs_shapes=obs_shapes, encoder_kwargs=enc_kwargs, **ICRT_KW)
    with torch.no_grad():
        net.params["embed_timestep"].normal_(0, 0.02)  # zeros at init: give the parity test something to see
    net.eval()
    B, T = 3, 10
    g = torch.Generator().manual_seed(1)
    obs = {k: torch.randn(B, T, *s, generator=g) for k, s in obs_shapes.items()}
    ctx = {k: torch.randn(B, T, *s, generator=g) for k, s in obs_shapes.items()}
    ctx_actions = torch.rand(B, T, 12, generator=g) * 2 - 1
    train_actions = torch.rand(B, T, 12, generator=g) * 2 - 1
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    z_latent, _ = net.vq_vae_model(ctx_actions.reshape(B * T, 12))
    dists = net.forward_train(obs_dict=obs, context_obs=ctx, actions=ctx_actions, low_noise_eval=False)
    logp = dists.log_prob(train_actions)
    action_loss = -logp.mean()
    vq_loss = net._vq_vae_loss
    net.zero_grad()
    action_loss.backward()
    keep = ("embed_encoder", "decoder.nets.mean", "transformer.0.nets.attention.nets.qkv", "embed_timestep", "output_ln",
            "transformer.1.nets.mlp.2")
    grads = {"grad." + k: p.grad.clone() for k, p in net.named_parameters() if p.grad is not None and any(t in k for t in keep)}
    sd = {k: v for k, v in sd.items() if not k.startswith("vq_vae_model.")}  # alias of nets.encoder.action_network.*
    base = dists.component_distribution.base_dist
    torch.manual_seed(3)
    bin_net = AdaptiveBinActionEmbedding(action_dim=12, output_dim=24, embedding_dim=8)
    bin_sd = {k: v.clone() for k, v in bin_net.state_dict().items()}
    a2 = ctx_actions.reshape(B * T, 12)
    bin_out1 = bin_net(a2)
    bin_out2 = bin_net(a2 * 0.5)  # second call: running min/max already set, narrower batch
    bin_idx = bin_net.discretize(a2)
    _save(tag, z_latent=z_latent, means=base.loc, scales=base.scale, logits=dists.mixture_distribution.logits, log_prob=logp,
          action_loss=action_loss, vq_loss=vq_loss, ctx_actions=ctx_actions, train_actions=train_actions,
          **{"obs." + k: v for k, v in obs.items()}, **{"ctx." + k: v for k, v in ctx.items()},
          **{"sd." + k: v for k, v in sd.items()}, **grads, **{"bin_sd." + k: v for k, v in bin_sd.items()},
          bin_out1=bin_out1, bin_out2=bin_out2, bin_idx=bin_idx,
          **{"bin_post." + k: v.clone() for k, v in bin_net.state_dict().items()})


def gen_dataset(tag):
    """Action-chunk input side: the reference's own SequenceDataset.get_item (robomimic/utils/dataset.py:499-586) run on
    in-memory demos.  The object is built without __init__ (which needs an HDF5 file and a language encoder); the index
    map follows dataset.py:229-260; window, padding, pad mask and normalisation are the reference's code."""
    _stub_missing_packages()
    from collections import OrderedDict
    import robomimic.utils.dataset as DS
    rng = np.random.RandomState(0)
    lens = [7, 23, 1, 12, 40]
    demos = {f"demo_{i}": (rng.rand(L, 12).astype(np.float32) * 2 - 1) for i, L in enumerate(lens)}
    stats = {"actions": {"offset": (rng.rand(1, 12).astype(np.float32) - 0.5), "scale": (rng.rand(1, 12).astype(np.float32) + 0.5)}}
    out = {"lens": np.asarray(lens), **{"demo." + k: v for k, v in demos.items()},
           "stats_offset": stats["actions"]["offset"], "stats_scale": stats["actions"]["scale"]}
    cfgs = [(10, 1, True, True, False), (4, 3, True, True, True), (5, 2, True, False, False), (3, 2, False, True, True)]
    for ci, (T, fs, pad_seq, pad_fs, norm) in enumerate(cfgs):
        ds = DS.SequenceDataset.__new__(DS.SequenceDataset)
        ds._hdf5_file = None
        ds.seq_length, ds.n_frame_stack, ds.pad_seq_length, ds.pad_frame_stack = T, fs, pad_seq, pad_fs
        ds.dataset_keys, ds.obs_keys, ds.action_keys = ("actions",), (), ("actions",)
        ds.goal_mode, ds.load_next_obs, ds.get_pad_mask = None, False, True
        ds.hdf5_cache_mode, ds.shuffled_obs_key_groups = None, []
        ident = {"actions": {"offset": np.zeros((1, 12), np.float32), "scale": np.ones((1, 12), np.float32)}}
        ds.action_normalization_stats = stats if norm else ident  # normalization=null in the shipped config -> identity stats
        ds.get_action_normalization_stats = lambda ds=ds: ds.action_normalization_stats
        ds._demo_id_to_demo_lang_emb = {}
        ds.get_dataset_for_ep = lambda ep, key: demos[ep][:] if key == "actions" else None
        ds._index_to_demo_id, ds._demo_id_to_start_indices, ds._demo_id_to_demo_length = {}, {}, {}
        total = 0
        for ep, L in zip(demos, lens):
            n = L - (0 if pad_fs else fs - 1) - (0 if pad_seq else T - 1)
            if pad_seq:
                n = max(n, 1)
            if n < 1 or (not pad_fs and L - (fs - 1) < 1):
                continue  # the reference asserts here; such a demo cannot be in the dataset
            ds._demo_id_to_start_indices[ep], ds._demo_id_to_demo_length[ep] = total, L
            for _ in range(n):
                ds._index_to_demo_id[total] = ep
                total += 1
        acts, masks = [], []
        for index in range(total):
            item = ds.get_item(index)
            acts.append(np.asarray(item["actions"], dtype=np.float32))
            masks.append(np.asarray(item["obs"]["pad_mask"])[:, 0])
        out.update({f"c{ci}.cfg": np.asarray([T, fs, int(pad_seq), int(pad_fs), int(norm)]), f"c{ci}.actions": np.stack(acts),
                    f"c{ci}.pad_mask": np.stack(masks), f"c{ci}.demos": np.asarray([int(e.split("_")[1]) for e in ds._demo_id_to_demo_length])})
    _save(tag, **out)


def main():
    v5 = _load("ref_v5", "robomimic/models/vq_vae/backbone_lfqvae_v5.py")
    hq = _load("ref_hq", "robomimic/models/vq_vae/hq_vae.py")
    bb = _load("ref_bb", "robomimic/models/vq_vae/backbone.py")
    hv = _load("ref_hv", "robomimic/algo/heirarchical_vqvae.py")
    torch.set_num_threads(1)
    gen_t1(v5, "t1_c1", 7, 64, 512, 64, steps=10)           # BASELINE C1 shape (rows reduced)
    gen_t1(v5, "t1_real", 12, 208, 128, 80, steps=5)        # reference __main__ shape (backbone_lfqvae_v5.py:89-93)
    gen_t1_ema(v5, hq, "t1_ema", 12, 64, 128, 256, steps=6)
    gen_quantizers(hq, "quantizers", 256, 64, 32)
    gen_t4(bb, "t4_classic", 7, 32, 128, 64)
    gen_t2(hv, "t2_hvq", 7, 64, 64, 32, 2, 3, 4, 16)
    gen_t3_hier(hq, "t3_hier")
    gen_kmeans(hq, hv, "kmeans_init")
    gen_icrt("icrt_policy")
    gen_dataset("dataset_windows")


if __name__ == "__main__":
    sys.exit(main())
