"""-m gpu: the MS-TCN layer's convolution on the tensor cores (csrc/conv_tc.cu, behind lipvq_dilated_conv_fwd_f32 for wide
layers with many tokens) against torch's Conv1d evaluation of DilatedTemporalConvLayer.forward
(heirarchical_vqvae.py:39-57) in fp64 and against the fp32 FFMA tiles of conv.cu."""
import os

import pytest
import torch

from test_gpu_kernels import DEV

pytestmark = pytest.mark.gpu


def _ref_layer(x_btE, W, b, dil):
    xb = x_btE.double().transpose(1, 2)                     # [B, E, T]
    pre = torch.nn.functional.conv1d(xb, W.double(), b.double(), dilation=dil, padding=dil)
    return (torch.relu(pre) + xb).transpose(1, 2), (pre > 0).transpose(1, 2)


@pytest.mark.parametrize("B,T,E,dil", [(512, 16, 512, 1), (512, 16, 512, 4), (300, 16, 128, 2), (1024, 10, 256, 8),
                                       (512, 8, 128, 16)])   # (the last: dilation >= T, only the centre tap is alive)
def test_conv_tc_forward_vs_torch_and_ffma(B, T, E, dil):
    import lipvq as L
    g = torch.Generator().manual_seed(B + T + E + dil)
    x = torch.randn(B, T, E, generator=g)
    W = torch.randn(E, E, 3, generator=g) / (3 * E) ** 0.5
    b = torch.randn(E, generator=g) * 0.1
    Wf, _ = L.ops.conv_pack_weights(W.to(DEV), True, False)
    xd = x.reshape(B * T, E).to(DEV)
    out, mask = L.ops.dilated_conv_fwd(xd, T, Wf, b.to(DEV), dil)
    os.environ["LIPVQ_CONV_ALGO"] = "simt"
    try:
        out2, mask2 = L.ops.dilated_conv_fwd(xd, T, Wf, b.to(DEV), dil)
    finally:
        del os.environ["LIPVQ_CONV_ALGO"]
    torch.cuda.synchronize()
    ref, keep = _ref_layer(x, W, b, dil)
    ref = ref.reshape(B * T, E)
    assert torch.allclose(out.cpu().double(), ref, rtol=1e-5, atol=2e-5), (out.cpu().double() - ref).abs().max()
    assert torch.allclose(out, out2, rtol=1e-5, atol=2e-5), (out - out2).abs().max()
    # keep-mask: equal except where the pre-activation is within rounding of zero
    dm = (mask.cpu().bool() != keep.reshape(B * T, E))
    assert dm.float().mean().item() < 1e-4
    assert (mask != mask2).float().mean().item() < 1e-4


def test_conv_tc_dropout_mask_is_the_ffma_path_s():
    import lipvq as L
    g = torch.Generator().manual_seed(3)
    B, T, E = 512, 16, 256
    x = torch.randn(B * T, E, generator=g).to(DEV)
    W = (torch.randn(E, E, 3, generator=g) / (3 * E) ** 0.5).to(DEV)
    b = (torch.randn(E, generator=g) * 0.1).to(DEV)
    Wf, _ = L.ops.conv_pack_weights(W, True, False)
    out, mask = L.ops.dilated_conv_fwd(x, T, Wf, b, 2, drop_p=0.3, seed=1234)
    os.environ["LIPVQ_CONV_ALGO"] = "simt"
    try:
        out2, mask2 = L.ops.dilated_conv_fwd(x, T, Wf, b, 2, drop_p=0.3, seed=1234)
    finally:
        del os.environ["LIPVQ_CONV_ALGO"]
    assert (mask != mask2).float().mean().item() < 1e-4     # same counter-based stream, same element numbering
    same = (mask == mask2)
    assert torch.allclose(out[same], out2[same], rtol=1e-5, atol=2e-5)
    assert 0.3 < mask.float().mean().item() / (out2 - x != 0).float().mean().item() < 3.0


def _simt(fn):
    os.environ["LIPVQ_CONV_ALGO"] = "simt"
    try:
        return fn()
    finally:
        del os.environ["LIPVQ_CONV_ALGO"]


@pytest.mark.parametrize("B,T,E,dil,drop", [(512, 16, 256, 1, 0.0), (300, 16, 128, 4, 0.0), (1024, 10, 256, 2, 0.25),
                                            (512, 8, 128, 16, 0.0), (4100, 1, 128, 1, 0.0)])
def test_conv_tc_backward_vs_autograd_and_ffma(B, T, E, dil, drop):
    """dX and dWf / db of the layer (heirarchical_vqvae.py:39-57 under autograd) from the tensor-core kernels: against torch
    autograd in fp64 with the SAME keep-mask, and against the FFMA tiles.  Tolerances: 2e-5 of the gradient's scale (three
    bf16 terms per operand, fp32 accumulators that truncate per MMA; measured 1e-6 .. 5e-6)."""
    import lipvq as L
    g = torch.Generator().manual_seed(B + T + E + dil)
    x = torch.randn(B, T, E, generator=g)
    W = torch.randn(E, E, 3, generator=g) / (3 * E) ** 0.5
    b = torch.randn(E, generator=g) * 0.1
    go = torch.randn(B, T, E, generator=g)
    N = B * T
    Wd = W.to(DEV)
    Wf, Wb = L.ops.conv_pack_weights(Wd, True, True)
    xd, god = x.reshape(N, E).to(DEV), go.reshape(N, E).to(DEV)
    out, mask = L.ops.dilated_conv_fwd(xd, T, Wf, b.to(DEV), dil, drop_p=drop, seed=77)
    dx = L.ops.dilated_conv_bwd_x(god, mask, T, Wb, dil, drop_p=drop)
    dWf, db = torch.zeros(E, 3, E, device=DEV), torch.zeros(E, device=DEV)
    L.ops.dilated_conv_bwd_w(god, mask, xd, T, dil, dWf, db, drop_p=drop)
    dx2 = _simt(lambda: L.ops.dilated_conv_bwd_x(god, mask, T, Wb, dil, drop_p=drop))
    dWf2, db2 = torch.zeros(E, 3, E, device=DEV), torch.zeros(E, device=DEV)
    _simt(lambda: L.ops.dilated_conv_bwd_w(god, mask, xd, T, dil, dWf2, db2, drop_p=drop))
    torch.cuda.synchronize()
    # fp64 autograd with the kernel's keep-mask as the ReLU / dropout gate
    xr = x.double().to(DEV).requires_grad_(True)
    Wr = W.double().to(DEV).requires_grad_(True)
    br = b.double().to(DEV).requires_grad_(True)
    pre = torch.nn.functional.conv1d(xr.transpose(1, 2), Wr, br, dilation=dil, padding=dil).transpose(1, 2)
    y = pre * mask.reshape(B, T, E).double() / (1.0 - drop) + xr
    y.backward(go.double().to(DEV))
    dxr = xr.grad.reshape(N, E)
    dWr = Wr.grad.permute(0, 2, 1).contiguous()          # [Eo, 3, Ei]

    def close(a, r, what):
        scale = r.abs().max().item()
        err = (a.double() - r).abs().max().item()
        assert err <= 2e-5 * scale, (what, err, scale)
    close(dx, dxr, "dx"); close(dx2, dxr, "dx ffma")
    close(dWf, dWr, "dWf"); close(dWf2, dWr, "dWf ffma")
    close(db, br.grad, "db"); close(db2, br.grad, "db ffma")


def test_conv_tc_bwd_w_accumulates_into_dwf():
    import lipvq as L
    g = torch.Generator().manual_seed(9)
    N, T, E = 4096, 16, 128
    x = torch.randn(N, E, generator=g).to(DEV)
    go = torch.randn(N, E, generator=g).to(DEV)
    mask = (torch.rand(N, E, generator=g) > 0.5).to(torch.uint8).to(DEV)
    base = torch.randn(E, 3, E, generator=g).to(DEV)
    d0, b0 = torch.zeros(E, 3, E, device=DEV), torch.zeros(E, device=DEV)
    L.ops.dilated_conv_bwd_w(go, mask, x, T, 2, d0, b0)
    d1, b1 = base.clone(), torch.ones(E, device=DEV)
    L.ops.dilated_conv_bwd_w(go, mask, x, T, 2, d1, b1)
    assert torch.allclose(d1 - base, d0, rtol=1e-4, atol=1e-3 * d0.abs().max().item())
    assert torch.allclose(b1 - 1.0, b0, rtol=1e-4, atol=1e-3 * b0.abs().max().item())


def test_conv_misaligned_mask_takes_the_ffma_tiles():
    """A keep-mask that is only 4-byte aligned (all the C-ABI asks for) cannot be read in 8-byte words: the entry points must
    fall back to the FFMA tiles instead of faulting, with the same results."""
    import lipvq as L
    g = torch.Generator().manual_seed(11)
    N, T, E = 2048, 16, 128
    x = torch.randn(N, E, generator=g).to(DEV)
    go = torch.randn(N, E, generator=g).to(DEV)
    W = (torch.randn(E, E, 3, generator=g) / (3 * E) ** 0.5).to(DEV)
    b = torch.zeros(E, device=DEV)
    Wf, Wb = L.ops.conv_pack_weights(W, True, True)
    _, mask = L.ops.dilated_conv_fwd(x, T, Wf, b, 2)
    store = torch.zeros(N * E + 16, dtype=torch.uint8, device=DEV)
    m4 = store[4:4 + N * E].view(N, E)
    assert m4.data_ptr() % 8 == 4
    m4.copy_(mask)
    dx = L.ops.dilated_conv_bwd_x(go, m4, T, Wb, 2)
    dx0 = L.ops.dilated_conv_bwd_x(go, mask, T, Wb, 2)
    d1, d0 = torch.zeros(E, 3, E, device=DEV), torch.zeros(E, 3, E, device=DEV)
    b1, b0 = torch.zeros(E, device=DEV), torch.zeros(E, device=DEV)
    L.ops.dilated_conv_bwd_w(go, m4, x, T, 2, d1, b1)
    L.ops.dilated_conv_bwd_w(go, mask, x, T, 2, d0, b0)
    torch.cuda.synchronize()
    assert torch.allclose(dx, dx0, rtol=1e-5, atol=2e-5 * dx0.abs().max().item())
    assert torch.allclose(d1, d0, rtol=1e-5, atol=2e-5 * d0.abs().max().item())
    assert torch.allclose(b1, b0, rtol=1e-5, atol=2e-5 * b0.abs().max().item())


@pytest.mark.parametrize("drop", [0.0, 0.2])
def test_conv_tc_few_tokens_split_by_tap(drop):
    """160 tokens (the reference's own batch: 16 chunks of 10) at E = 512: the live taps of a tile run in different CTAs and add
    their partial sums into OUT / dX.  Same values (fp32 atomic order aside) and the same mask as the unsplit tensor path and
    as the FFMA tiles."""
    import lipvq as L
    g = torch.Generator().manual_seed(21)
    B, T, E, dil = 16, 10, 512, 2
    N = B * T
    x = torch.randn(N, E, generator=g).to(DEV)
    go = torch.randn(N, E, generator=g).to(DEV)
    W = (torch.randn(E, E, 3, generator=g) / (3 * E) ** 0.5).to(DEV)
    b = (torch.randn(E, generator=g) * 0.1).to(DEV)
    Wf, Wb = L.ops.conv_pack_weights(W, True, True)

    def run():
        out, mask = L.ops.dilated_conv_fwd(x, T, Wf, b, dil, drop_p=drop, seed=99)
        return out, mask, L.ops.dilated_conv_bwd_x(go, mask, T, Wb, dil, drop_p=drop)
    out, mask, dx = run()
    os.environ["LIPVQ_CONV_SPLIT"] = "0"
    try:
        out1, mask1, dx1 = run()
    finally:
        del os.environ["LIPVQ_CONV_SPLIT"]
    out2, mask2, dx2 = _simt(run)
    torch.cuda.synchronize()
    for o, m_, d in ((out1, mask1, dx1), (out2, mask2, dx2)):
        assert (mask != m_).float().mean().item() < 1e-4
        same = mask == m_
        assert torch.allclose(out[same], o[same], rtol=1e-5, atol=2e-5)
    # the gradients use their own run's mask: compare where the masks agree everywhere on the rows involved
    if torch.equal(mask, mask1):
        assert torch.allclose(dx, dx1, rtol=1e-5, atol=2e-5 * dx1.abs().max().item())
    if torch.equal(mask, mask2):
        assert torch.allclose(dx, dx2, rtol=1e-5, atol=2e-5 * dx2.abs().max().item())


def _bf16(fn):
    os.environ["LIPVQ_CONV_PRECISION"] = "bf16"
    try:
        return fn()
    finally:
        del os.environ["LIPVQ_CONV_PRECISION"]


@pytest.mark.parametrize("B,T,E", [(512, 16, 512), (16, 10, 512)])
def test_conv_bf16_mode_within_the_bf16_gate(B, T, E):
    """LIPVQ_CONV_PRECISION=bf16: plain bf16 operands, one MMA chain, fp32 accumulation and epilogue.  north_star's gate for
    bf16 is 1e-2: forward values and both gradients within 1e-2 of the fp32-accurate path, relative to the tensor's scale."""
    import lipvq as L
    g = torch.Generator().manual_seed(31)
    N = B * T
    x = torch.randn(N, E, generator=g).to(DEV)
    go = torch.randn(N, E, generator=g).to(DEV)
    W = (torch.randn(E, E, 3, generator=g) / (3 * E) ** 0.5).to(DEV)
    b = (torch.randn(E, generator=g) * 0.1).to(DEV)
    Wf, Wb = L.ops.conv_pack_weights(W, True, True)
    out, mask = L.ops.dilated_conv_fwd(x, T, Wf, b, 2)

    def grads():
        dx = L.ops.dilated_conv_bwd_x(go, mask, T, Wb, 2)
        dW, db = torch.zeros(E, 3, E, device=DEV), torch.zeros(E, device=DEV)
        L.ops.dilated_conv_bwd_w(go, mask, x, T, 2, dW, db)
        return dx, dW, db
    dx, dW, db = grads()
    out_b, mask_b = _bf16(lambda: L.ops.dilated_conv_fwd(x, T, Wf, b, 2))
    dx_b, dW_b, db_b = _bf16(grads)
    torch.cuda.synchronize()

    def gate(a, r, what):
        err = (a - r).abs().max().item() / r.abs().max().item()
        assert err < 1e-2, (what, err)
        assert err > 1e-6 or what == "db", (what, err, "the bf16 mode did not run")   # (db is summed in fp32 in both modes)
    gate(out_b, out, "out"); gate(dx_b, dx, "dx"); gate(dW_b, dW, "dW"); gate(db_b, db, "db")
    assert (mask_b != mask).float().mean().item() < 5e-3     # keep-mask flips only where the pre-activation is ~0


def test_t2_embeddings_in_bf16_conv_mode_within_gate():
    import lipvq as L
    torch.manual_seed(7)
    m = L.HierarchicalVQVAE(12, embed_dim=512, num_subclusters=1024, num_clusters=512, num_stages=2, num_layers_per_stage=4,
                            dropout=0.0, kmeans_init=False).to(DEV).eval()
    x = torch.rand(64, 16, 12, device=DEV) * 2 - 1
    with torch.no_grad():
        e32 = m.encoder(x)
        e16 = _bf16(lambda: m.encoder(x))
    err = (e16 - e32).abs().max().item() / e32.abs().max().item()
    assert 1e-6 < err < 1e-2, err
