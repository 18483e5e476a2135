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
