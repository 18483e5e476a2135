"""-m gpu: the LipVQ-VAE tokenizer end to end (module API and trainer) against the golden fixtures produced by the
reference's own modules (oracle/gen_golden.py) and against the oracle on fresh inputs."""
import pytest
import torch

import lipvq_oracle as O
from conftest import load_golden, sub
from test_gpu_kernels import assert_indices

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _model(L, name, F_, D, K, ema=False):
    g = load_golden(name)
    m = L.LLFQVAE_V4(F_, D, K)
    if ema:
        m.quantizer = L.LFQQuantizerEMA(K, D)
    m.load_state_dict(sub(g, "sd"), strict=True)
    return g, m.to(DEV)


@pytest.mark.parametrize("name,F_,D,K", [("t1_c1", 7, 64, 512), ("t1_real", 12, 208, 128)])
def test_t1_forward_backward_vs_reference_fixture(name, F_, D, K):
    import lipvq as L
    g, m = _model(L, name, F_, D, K)
    x = g["x"].to(DEV)
    before = L.launch_count()
    z_latent, loss = m(x)
    assert L.launch_count() > before  # our kernels ran
    assert z_latent.grad_fn is None and not z_latent.requires_grad and loss.dim() == 0 and loss.requires_grad
    nt = assert_indices(g["z_e"], sub(g, "sd")["quantizer.codebook"], m.last_indices, g["indices"])
    if nt == 0:
        assert torch.equal(z_latent.cpu(), g["z_latent"])  # gathered codewords bit-exact
    assert abs(loss.item() - g["loss"].item()) <= 1e-4 * abs(g["loss"].item())
    loss.backward()
    for k, gr in sub(g, "grad").items():
        got = dict(m.named_parameters())[k].grad.cpu()
        assert torch.allclose(got, gr, rtol=1e-4, atol=1e-7), f"{k}: max abs diff {(got - gr).abs().max():.3e}"
    # encode-only path returns the same tokens
    z_e, idx = m.encode(x)
    assert torch.equal(idx, m.last_indices)
    assert torch.allclose(z_e.cpu(), g["z_e"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("name,F_,D,K,use_graph", [("t1_c1", 7, 64, 512, True), ("t1_real", 12, 208, 128, False)])
def test_t1_adamw_trajectory_vs_reference_fixture(name, F_, D, K, use_graph):
    import lipvq as L
    from lipvq.trainer import TokenizerTrainer
    g, m = _model(L, name, F_, D, K)
    x = g["x"].to(DEV)
    tr = TokenizerTrainer(m, x.shape[0], use_graph=use_graph)
    losses = []
    for _ in range(g["traj_losses"].numel()):
        losses.append(tr.step(x)[0].item())
    ref = g["traj_losses"]
    assert torch.allclose(torch.tensor(losses), ref, rtol=1e-4, atol=1e-6), (losses, ref.tolist())
    sd = m.state_dict()  # parameters live in the trainer's slab but state_dict still sees them
    for k, v in sub(g, "final").items():
        assert torch.allclose(sd[k].cpu(), v, rtol=1e-3, atol=2e-5), k


def test_t1_ema_trajectory_vs_reference_fixture():
    import lipvq as L
    from lipvq.trainer import TokenizerTrainer
    g, m = _model(L, "t1_ema", 12, 64, 128, ema=True)
    m.train()
    x = g["x"].to(DEV)
    tr = TokenizerTrainer(m, x.shape[0], use_graph=True)
    losses = [tr.step(x)[0].item() for _ in range(g["traj_losses"].numel())]
    assert torch.allclose(torch.tensor(losses), g["traj_losses"], rtol=1e-4, atol=1e-6), (losses, g["traj_losses"].tolist())
    sd = m.state_dict()
    fin = sub(g, "final")
    for k in ("quantizer.ema_cluster_size", "quantizer.ema_codebook", "quantizer.codebook"):
        assert torch.allclose(sd[k].cpu(), fin[k], rtol=1e-3, atol=2e-5), k


def test_t1_ema_module_path_matches_trainer_path():
    """nn.Module + torch.optim.AdamW (what a reference user writes) == TokenizerTrainer (flat slabs, graphs)."""
    import lipvq as L
    from lipvq.trainer import TokenizerTrainer
    g, m1 = _model(L, "t1_ema", 12, 64, 128, ema=True)
    _, m2 = _model(L, "t1_ema", 12, 64, 128, ema=True)
    m1.train(), m2.train()
    x = g["x"].to(DEV)
    opt = torch.optim.AdamW(m1.parameters(), lr=1e-3, weight_decay=1e-4)
    tr = TokenizerTrainer(m2, x.shape[0], use_graph=False)
    for _ in range(3):
        opt.zero_grad()
        _, loss = m1(x)
        loss.backward()
        opt.step()
        l2 = tr.step(x)[0].item()
        assert abs(loss.item() - l2) <= 1e-5 * abs(l2)
    for (k, a), (_, b) in zip(m1.state_dict().items(), m2.state_dict().items()):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-6), k


def test_t1_training_step_at_baseline_batch():
    """BASELINE config 2 shape (4096 chunks x 16 steps = 65536 tokens, 12-DoF, K=512, D=64): one step against the
    oracle, plus size-independent properties (losses decrease over steps, counts sum to N)."""
    import lipvq as L
    from lipvq.trainer import TokenizerTrainer
    torch.manual_seed(0)
    F_, D, K, N = 12, 64, 512, 65536
    m = L.LLFQVAE_V4(F_, D, K)
    m.quantizer = L.LFQQuantizerEMA(K, D)
    p0 = {k: v.clone() for k, v in m.state_dict().items()}
    m.to(DEV).train()
    x = torch.rand(N, F_, generator=torch.Generator().manual_seed(123)) * 2 - 1
    tr = TokenizerTrainer(m, N, use_graph=True)
    l0 = tr.step(x.to(DEV)).clone()
    idx = tr.indices.clone()
    # oracle: same step on CPU
    ema = {"ema_cluster_size": p0.pop("quantizer.ema_cluster_size"), "ema_codebook": p0.pop("quantizer.ema_codebook")}
    cb0 = p0["quantizer.codebook"].clone()  # the codebook the search saw (the step overwrites it)
    z_gpu = tr.bufs.t["z_e"].clone().cpu()
    ref = O.t1_train_step(p0, x, {}, ema=ema)
    # (i) the search itself: on the GPU's OWN latents every index is the oracle's, or a near-tie (fp64 gap < 1e-6)
    assert_indices(z_gpu, cb0, idx, O.nearest_code_l2(z_gpu, cb0, row_chunk=2048), max_near_tie_frac=1e-4)
    # (ii) end to end: the encoder's fp32 noise (~1e-6 on z_e) may flip a token only where the oracle's own two best
    # codes are closer than that noise -- every mismatch is classified, none is just counted
    assert torch.allclose(z_gpu, ref["z_e"], rtol=1e-4, atol=2e-6)
    bad = (idx.cpu() != ref["indices"]).nonzero().flatten()
    if bad.numel():
        zz = ref["z_e"][bad].double()
        dg = (zz - cb0[idx.cpu()[bad]].double()).norm(dim=-1)
        dr = (zz - cb0[ref["indices"][bad]].double()).norm(dim=-1)
        gap = ((dg - dr).abs() / dr).max().item()
        assert gap < 2e-5, f"{bad.numel()} token mismatches vs oracle, worst relative distance gap {gap:.3e}"
    assert bad.numel() <= 0.001 * N, f"{bad.numel()} token mismatches vs oracle"
    assert abs(l0[0].item() - ref["loss"].item()) <= 1e-4 * abs(ref["loss"].item())
    sd = m.state_dict()
    assert abs(sd["quantizer.ema_cluster_size"].sum().item() - 0.01 * N) <= 1e-3 * 0.01 * N
    # gradients of this step are still in the trainer's reduction slab (AdamW's first update is ~lr*sign(g), so
    # comparing post-step weights would amplify fp32 noise on near-zero gradients; compare the gradients themselves)
    for k in ("encoder.0.weight", "encoder.2.weight", "to_latent.W", "to_latent.ci", "decoder.0.weight", "to_output.bias"):
        gr = ref["grads"][k]
        assert torch.allclose(tr.g[k].cpu(), gr, rtol=1e-3, atol=1e-3 * gr.abs().max().item()), k
    losses = [l0[0].item()] + [tr.step(x.to(DEV))[0].item() for _ in range(5)]
    assert losses[-1] < losses[0]
    toks = tr.tokenize(x.to(DEV))
    assert toks.shape == (N,) and int(toks.min()) >= 0 and int(toks.max()) < K


def test_trainer_prefetch_matches_plain_steps():
    """step(x, prefetch=x_next) (next batch's H2D copy under this step's kernels) gives the same losses as feeding the
    same pinned batches one call at a time (to fp32 reduction-order noise: dW/db/codebook sums use atomics)."""
    import lipvq as L
    from lipvq.trainer import TokenizerTrainer
    F_, D, K, N = 12, 64, 512, 8192
    xs = [(torch.rand(N, F_, generator=torch.Generator().manual_seed(50 + i)) * 2 - 1).pin_memory() for i in range(4)]
    out = []
    for use_prefetch in (False, True):
        torch.manual_seed(0)
        m = L.LLFQVAE_V4(F_, D, K)
        m.quantizer = L.LFQQuantizerEMA(K, D)
        m.to(DEV).train()
        tr = TokenizerTrainer(m, N, use_graph=True)
        ls = []
        for i in range(6):
            nxt = xs[(i + 1) % 4] if use_prefetch else None
            ls.append(tr.step(xs[i % 4], prefetch=nxt).cpu().clone())
        out.append(torch.stack(ls))
    assert torch.allclose(out[0], out[1], rtol=2e-5, atol=1e-7), (out[0] - out[1]).abs().max()
