"""-m gpu: lipvq_mlp3_fwd_f32 (three linear layers in one launch, csrc/mlp_chain.cu) through the C-ABI against the CPU
oracle's encoder / decoder (backbone_lfqvae_v5.py:54-68, 70-76) and against three lipvq_linear_fwd_f32 launches."""
import pytest
import torch

import lipvq_oracle as O
from test_gpu_kernels import DEV

pytestmark = pytest.mark.gpu


def _three(L, x, W1, b1, a1, W2, b2, a2, W3, b3, a3):
    h = L.ops.linear_fwd(x, L.ACT_NONE, W1, b1, a1)
    h = L.ops.linear_fwd(h, L.ACT_NONE, W2, b2, a2)
    return L.ops.linear_fwd(h, L.ACT_NONE, W3, b3, a3)


@pytest.mark.parametrize("N,dims", [(65536, (12, 64, 128, 64)), (4096, (7, 64, 128, 64)), (9000, (64, 64, 128, 12)),
                                    (5000, (12, 64, 128, 48)), (4100, (30, 50, 100, 20)), (6000, (100, 64, 72, 16))])
def test_mlp3_vs_three_launches_and_torch(N, dims):
    import lipvq as L
    d0, d1, d2, d3 = dims
    assert L.ops.mlp3_supported(N, d0, d1, d2, d3)
    g = torch.Generator().manual_seed(N + d0 + d3)
    x = torch.rand(N, d0, generator=g) * 2 - 1
    Ws = [torch.randn(o, i, generator=g) / i ** 0.5 for i, o in ((d0, d1), (d1, d2), (d2, d3))]
    bs = [torch.randn(o, generator=g) * 0.1 for o in (d1, d2, d3)]
    acts = (L.ACT_GELU, L.ACT_GELU, L.ACT_SIGMOID)
    xd = x.to(DEV)
    Wd, bd = [w.to(DEV) for w in Ws], [b.to(DEV) for b in bs]
    y = L.ops.mlp3_fwd(xd, Wd[0], bd[0], acts[0], Wd[1], bd[1], acts[1], Wd[2], bd[2], acts[2])
    y3 = _three(L, xd, Wd[0], bd[0], acts[0], Wd[1], bd[1], acts[1], Wd[2], bd[2], acts[2])
    torch.cuda.synchronize()
    h = torch.nn.functional.gelu(x.double() @ Ws[0].double().t() + bs[0].double())
    h = torch.nn.functional.gelu(h @ Ws[1].double().t() + bs[1].double())
    ref = torch.sigmoid(h @ Ws[2].double().t() + bs[2].double())
    assert torch.allclose(y.cpu().double(), ref, rtol=1e-5, atol=2e-6), (y.cpu().double() - ref).abs().max()
    assert torch.allclose(y, y3, rtol=1e-6, atol=1e-6), (y - y3).abs().max()
    # no bias, other activations (the decoder ends without one)
    y = L.ops.mlp3_fwd(xd, Wd[0], None, L.ACT_RELU, Wd[1], bd[1], L.ACT_GELU, Wd[2], None, L.ACT_NONE)
    h = torch.relu(x.double() @ Ws[0].double().t())
    h = torch.nn.functional.gelu(h @ Ws[1].double().t() + bs[1].double())
    ref = h @ Ws[2].double().t()
    assert torch.allclose(y.cpu().double(), ref, rtol=1e-5, atol=5e-6), (y.cpu().double() - ref).abs().max()


def test_tokenize_uses_the_fused_encoder_and_matches_the_oracle():
    import lipvq as L
    from lipvq.trainer import TokenizerTrainer
    torch.manual_seed(1)
    N = 8192
    m = L.LLFQVAE_V4(12, 64, 512)
    m.quantizer = L.LFQQuantizerEMA(512, 64)
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    m.to(DEV).train()
    tr = TokenizerTrainer(m, N)
    x = torch.rand(N, 12) * 2 - 1
    c0 = L.launch_count()
    idx = tr.tokenize(x.to(DEV)).cpu()
    z_e = tr.bufs.t["z_e"].cpu()
    p = {k: sd[k] for k in ("encoder.0.weight", "encoder.0.bias", "encoder.2.weight", "encoder.2.bias", "to_latent.W",
                            "to_latent.b", "to_latent.ci", "quantizer.codebook")}
    ref_z = O.t1_encode(p, x)
    assert torch.allclose(z_e, ref_z, rtol=1e-5, atol=2e-6), (z_e - ref_z).abs().max()
    from test_gpu_kernels import assert_indices
    assert_indices(ref_z, p["quantizer.codebook"], idx, O.nearest_code_l2(ref_z, p["quantizer.codebook"]), max_near_tie_frac=0.02)


def test_mlp3_bf16_mode_within_the_reduced_precision_gate():
    """north_star: losses and reconstructions within 1e-2 for bf16.  The bf16 mode of the fused encoder (plain bf16 operands,
    fp32 accumulation / bias / activation) against the fp64 evaluation of the same layers, and tokenize(bf16=True) against
    the fp32 path: same token except where two codes are about equally near."""
    import lipvq as L
    from lipvq.trainer import TokenizerTrainer
    g = torch.Generator().manual_seed(5)
    N, dims = 16384, (12, 64, 128, 64)
    x = torch.rand(N, dims[0], generator=g) * 2 - 1
    Ws = [torch.randn(o, i, generator=g) / i ** 0.5 for i, o in zip(dims[:-1], dims[1:])]
    bs = [torch.randn(o, generator=g) * 0.1 for o in dims[1:]]
    y = L.ops.mlp3_fwd(x.to(DEV), Ws[0].to(DEV), bs[0].to(DEV), L.ACT_GELU, Ws[1].to(DEV), bs[1].to(DEV), L.ACT_GELU,
                       Ws[2].to(DEV), bs[2].to(DEV), L.ACT_SIGMOID, bf16=True).cpu().double()
    h = torch.nn.functional.gelu(x.double() @ Ws[0].double().t() + bs[0].double())
    h = torch.nn.functional.gelu(h @ Ws[1].double().t() + bs[1].double())
    ref = torch.sigmoid(h @ Ws[2].double().t() + bs[2].double())
    err = (y - ref).abs().max().item()
    assert 1e-6 < err < 1e-2, err      # (bf16 operands: not the fp32-accurate path, and inside the gate)
    torch.manual_seed(2)
    m = L.LLFQVAE_V4(12, 64, 512)
    m.quantizer = L.LFQQuantizerEMA(512, 64)
    m.to(DEV).train()
    tr = TokenizerTrainer(m, 8192)
    xb = (torch.rand(8192, 12) * 2 - 1).to(DEV)
    for _ in range(3):
        tr.step(xb)
    a = tr.tokenize(xb).clone()
    z32 = tr.bufs.t["z_e"].clone()
    b = tr.tokenize(xb, bf16=True).clone()
    z16 = tr.bufs.t["z_e"].clone()
    assert (z16 - z32).abs().max().item() < 1e-2
    cb = tr.p["quantizer.codebook"]
    diff = (a != b).nonzero().flatten()
    if diff.numel():   # a different token only where the fp32 latent is about as near to the other code
        da = (z32[diff] - cb[a[diff]]).norm(dim=-1)
        db = (z32[diff] - cb[b[diff]]).norm(dim=-1)
        assert ((db - da) <= 2e-2 * (1 + da)).all()
    assert diff.numel() <= 0.2 * a.numel()


def test_mlp3_train_variant_stores_what_two_dual_layers_store():
    """lipvq_mlp3_fwd_train_f32 against lipvq_linear_fwd_dual_f32 x 2 + lipvq_linear_fwd_f32: hidden activations, their
    derivatives and the output, at the encoder's and the decoder's widths."""
    import lipvq as L
    for N, dims, act3 in ((65536, (12, 64, 128, 64), L.ACT_SIGMOID), (9000, (64, 64, 128, 12), L.ACT_NONE)):
        d0, d1, d2, d3 = dims
        g = torch.Generator().manual_seed(N + d3)
        x = (torch.rand(N, d0, generator=g) * 2 - 1).to(DEV)
        Ws = [(torch.randn(o, i, generator=g) / i ** 0.5).to(DEV) for i, o in ((d0, d1), (d1, d2), (d2, d3))]
        bs = [(torch.randn(o, generator=g) * 0.1).to(DEV) for o in (d1, d2, d3)]
        e = lambda *s_: torch.empty(*s_, device=DEV)
        y, h1, g1, h2, g2 = e(N, d3), e(N, d1), e(N, d1), e(N, d2), e(N, d2)
        L.ops.mlp3_fwd_train(x, Ws[0], bs[0], L.ACT_GELU, Ws[1], bs[1], L.ACT_GELU, Ws[2], bs[2], act3, y, h1, g1, h2, g2)
        r1, q1, r2, q2 = e(N, d1), e(N, d1), e(N, d2), e(N, d2)
        L.ops.linear_fwd_dual(x, L.ACT_NONE, Ws[0], bs[0], L.ACT_GELU, r1, q1)
        L.ops.linear_fwd_dual(r1, L.ACT_NONE, Ws[1], bs[1], L.ACT_GELU, r2, q2)
        ry = L.ops.linear_fwd(r2, L.ACT_NONE, Ws[2], bs[2], act3)
        torch.cuda.synchronize()
        for a, b, what in ((h1, r1, "h1"), (g1, q1, "g1"), (h2, r2, "h2"), (g2, q2, "g2"), (y, ry, "y")):
            assert torch.allclose(a, b, rtol=1e-6, atol=1e-6), (what, (a - b).abs().max().item())
