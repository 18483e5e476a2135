"""Kernel schedule of the LipVQ-VAE tokenizer (T1): forward, backward, EMA -- as launches of liblipvq.so.

Follows LLFQVAE_V4.forward (reference robomimic/models/vq_vae/backbone_lfqvae_v5.py:70-84) and, with
``ema``, LFQQuantizerEMA.forward (robomimic/models/vq_vae/hq_vae.py:240-281).  Gradient topology is
the reference's: z_q = codebook[idx] WITHOUT straight-through, so the reconstruction loss trains the
decoder and the selected code rows, and the encoder learns through the commitment term only.

Used by both front-ends: the nn.Module / autograd path (modules.py) and the graph-captured trainer
(trainer.py).  All buffers are caller-visible tensors; nothing here allocates when ``bufs`` is reused.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

from . import _cabi as A
from . import ops

PARAM_SHAPES = lambda F, D, K, H=128: {  # state_dict order/names of LLFQVAE_V4 (backbone_lfqvae_v5.py:52-68)
    "encoder.0.weight": (64, F), "encoder.0.bias": (64,),
    "encoder.2.weight": (H, 64), "encoder.2.bias": (H,),
    "to_latent.W": (D, H), "to_latent.b": (D,), "to_latent.ci": (D,),
    "quantizer.codebook": (K, D),
    "decoder.0.weight": (64, D), "decoder.0.bias": (64,),
    "decoder.2.weight": (H, 64), "decoder.2.bias": (H,),
    "to_output.weight": (F, H), "to_output.bias": (F,),
}


@dataclass
class T1Buffers:
    """Activations of one step.  Hidden layers store GELU(pre) (a1, a2, c1, c2) and, for training, GELU'(pre) (the d*
    buffers): the producing layer's epilogue evaluates erf once and every consumer (next layer, weight gradient, input
    gradient) reads the result."""
    N: int
    F: int
    D: int
    K: int
    H: int
    device: torch.device
    t: Dict[str, torch.Tensor] = field(default_factory=dict)

    def __post_init__(self):
        N, F, D, K, H, dev = self.N, self.F, self.D, self.K, self.H, self.device
        e = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
        self.t = {
            "Wn": e(D, H), "a1": e(N, 64), "a2": e(N, H), "z_e": e(N, D),
            "idx": torch.empty(N, dtype=torch.int64, device=dev), "z_q": e(N, D),
            "c1": e(N, 64), "c2": e(N, H), "x_recon": e(N, F),
            "sq": torch.zeros(2, dtype=torch.float64, device=dev),   # [sum (x_recon-x)^2, sum (z_q-z_e)^2]
            "losses": e(4),                                          # loss, recon, commit, codebook
            "stats": torch.zeros(4, dtype=torch.int32, device=dev),
            "ws": torch.empty(max(1, ops.search_workspace_bytes(N, D, K)), dtype=torch.uint8, device=dev),
        }
        # deterministic per-code sums (stable counting sort + ordered segment sums, csrc/scatter_det.cu): opt-in with
        # LIPVQ_SCATTER=sorted -- bit-identical run to run, but measured slower than the shared-memory atomics of
        # quantize.cu (spread codes: 70 vs 47 us for the two sums; a collapsed codebook: 240 vs 48 us), so not the default
        import os
        nbytes = ops.code_sort_workspace_bytes(N, D, K) if (N >= 4096 and os.environ.get("LIPVQ_SCATTER", "atomic") == "sorted") else 0
        self.sort_ws = torch.empty(nbytes, dtype=torch.uint8, device=dev) if nbytes else None
        self.sorted_ok = False  # sort_ws holds the order of the CURRENT t["idx"]
        # the search can run on a codebook prepared once per codebook version (ops.nn_search_prepare -> t["ws"])
        self.search_prepared_ok = ops.search_prepared_supported(N, D, K)

    def alloc_derivatives(self):
        N, H, dev = self.N, self.H, self.device
        if "da1_act" not in self.t:
            e = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
            self.t.update({"da1_act": e(N, 64), "da2_act": e(N, H), "dc1_act": e(N, 64), "dc2_act": e(N, H)})

    def alloc_backward(self):
        N, F, D, H, dev = self.N, self.F, self.D, self.H, self.device
        e = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
        if "dxr" not in self.t:
            self.t.update({"dxr": e(N, F), "dc2": e(N, H), "dc1": e(N, 64), "dzq": e(N, D), "dzpre": e(N, D),
                           "da2": e(N, H), "da1": e(N, 64), "G": e(D, H),
                           "scratch": torch.zeros(1, dtype=torch.float64, device=dev)})


def _fused_train() -> bool:
    import os
    # opt-in: measured SLOWER than the per-layer launches (0.497 vs 0.455 ms per step) -- in the fused kernel the four
    # hidden tiles of a row block leave as one 32-byte piece per lane and array from the serial chain of a pipeline,
    # while linear_tc_kernel stores them through TMA under the next block's MMAs
    return os.environ.get("LIPVQ_MLP3_TRAIN", "0") == "1"


def prepare_search(p: Dict[str, torch.Tensor], b: T1Buffers, center: Optional[torch.Tensor] = None) -> None:
    """Codebook-dependent half of the search, once per codebook version (see include/lipvq.h, K3)."""
    ops.nn_search_prepare(p["quantizer.codebook"], b.t["ws"], A.METRIC_L2, center=center)


def forward(p: Dict[str, torch.Tensor], x: torch.Tensor, b: T1Buffers, *, algo: int = A.ALGO_AUTO,
            encode_only: bool = False, search_prepared: bool = False, weights_prepared: bool = False,
            prepare_search_here: bool = False, encoder_bf16: bool = False) -> None:
    """x [N,F] -> b.t[...]: z_e, idx, z_q, x_recon, losses.  ``encode_only`` stops after the search
    (the bulk-tokenisation path: encode + VQ).  ``search_prepared``: b.t["ws"] was prepared for the current codebook
    (prepare_search); ``weights_prepared``: b.t["Wn"] already holds the clamped to_latent weight of the current
    parameters (bulk tokenisation with fixed weights); ``prepare_search_here``: prepare the codebook inside this call,
    centred on this batch's latents, right before the search (then ``search_prepared`` applies); ``encoder_bf16``
    (bulk tokenisation only): the fused encoder in its bf16 mode."""
    t = b.t
    if not weights_prepared:
        ops.lipschitz_scale_fwd(p["to_latent.W"], p["to_latent.ci"], t["Wn"])
    fused_train = False
    fused_enc = encode_only and ops.mlp3_supported(b.N, b.F, 64, b.H, b.D)
    if fused_enc:  # bulk tokenisation: the three encoder layers in one launch, hidden activations never leave the SM
        ops.mlp3_fwd(x, p["encoder.0.weight"], p["encoder.0.bias"], A.ACT_GELU, p["encoder.2.weight"], p["encoder.2.bias"],
                     A.ACT_GELU, t["Wn"], p["to_latent.b"], A.ACT_SIGMOID, out=t["z_e"], bf16=encoder_bf16)
    elif encode_only:
        ops.linear_fwd(x, A.ACT_NONE, p["encoder.0.weight"], p["encoder.0.bias"], A.ACT_GELU, out=t["a1"])
        ops.linear_fwd(t["a1"], A.ACT_NONE, p["encoder.2.weight"], p["encoder.2.bias"], A.ACT_GELU, out=t["a2"])
    else:
        b.alloc_derivatives()
        # training forward: the three encoder layers in one launch that also stores the hidden activations and their
        # derivatives (opt-in, LIPVQ_MLP3_TRAIN=1; default: three launches per chain)
        fused_train = _fused_train() and ops.mlp3_supported(b.N, b.F, 64, b.H, b.D) and ops.mlp3_supported(b.N, b.D, 64, b.H, b.F)
        if fused_train:
            ops.mlp3_fwd_train(x, p["encoder.0.weight"], p["encoder.0.bias"], A.ACT_GELU, p["encoder.2.weight"],
                               p["encoder.2.bias"], A.ACT_GELU, t["Wn"], p["to_latent.b"], A.ACT_SIGMOID, t["z_e"],
                               t["a1"], t["da1_act"], t["a2"], t["da2_act"])
            fused_enc = True
        else:
            ops.linear_fwd_dual(x, A.ACT_NONE, p["encoder.0.weight"], p["encoder.0.bias"], A.ACT_GELU, t["a1"], t["da1_act"])
            ops.linear_fwd_dual(t["a1"], A.ACT_NONE, p["encoder.2.weight"], p["encoder.2.bias"], A.ACT_GELU, t["a2"], t["da2_act"])
    if not fused_enc:
        ops.linear_fwd(t["a2"], A.ACT_NONE, t["Wn"], p["to_latent.b"], A.ACT_SIGMOID, out=t["z_e"])
    if prepare_search_here:
        prepare_search(p, b, center=t["z_e"])
    if search_prepared or prepare_search_here:
        ops.nn_search_prepared(t["z_e"], p["quantizer.codebook"], t["ws"], A.METRIC_L2, out=t["idx"])
    else:
        ops.nn_search(t["z_e"], p["quantizer.codebook"], A.METRIC_L2, algo, out=t["idx"], workspace=t["ws"])
    b.sorted_ok = False
    if encode_only:
        return
    if not getattr(b, "zeroed_by_caller", False):  # the trainer clears sq and G with one fill at the start of the step
        t["sq"].zero_()
    ops.gather_fwd(t["z_e"], t["idx"], p["quantizer.codebook"], t["sq"][1:], z_q=t["z_q"])
    if fused_train:
        ops.mlp3_fwd_train(t["z_q"], p["decoder.0.weight"], p["decoder.0.bias"], A.ACT_GELU, p["decoder.2.weight"],
                           p["decoder.2.bias"], A.ACT_GELU, p["to_output.weight"], p["to_output.bias"], A.ACT_NONE,
                           t["x_recon"], t["c1"], t["dc1_act"], t["c2"], t["dc2_act"])
    else:
        ops.linear_fwd_dual(t["z_q"], A.ACT_NONE, p["decoder.0.weight"], p["decoder.0.bias"], A.ACT_GELU, t["c1"], t["dc1_act"])
        ops.linear_fwd_dual(t["c1"], A.ACT_NONE, p["decoder.2.weight"], p["decoder.2.bias"], A.ACT_GELU, t["c2"], t["dc2_act"])
        ops.linear_fwd(t["c2"], A.ACT_NONE, p["to_output.weight"], p["to_output.bias"], A.ACT_NONE, out=t["x_recon"])
    # when the backward buffers already exist (a training step) the same pass also writes d recon_loss / d x_recon,
    # which backward() then finds in place as long as the loss gradient is 1 (the trainer's case)
    b.dxr_from_forward = "dxr" in t
    ops.mse_fwd_bwd(t["x_recon"], x, t["sq"][0:1], dY=t["dxr"] if b.dxr_from_forward else None)
    ops.t1_loss_finalize(t["sq"], float(b.N) * b.F, float(b.N) * b.D, 0.25, 0.25, t["losses"])


def backward(p: Dict[str, torch.Tensor], g: Dict[str, torch.Tensor], x: torch.Tensor, b: T1Buffers,
             gscale: Optional[torch.Tensor] = None) -> None:
    """Accumulates d loss / d param into ``g`` (same keys as ``p``; caller zeroes them).  ``gscale``:
    optional 1-element device tensor multiplying the loss gradient (autograd's grad_output)."""
    b.alloc_backward()
    t = b.t
    scratch = t["scratch"]
    # d recon_loss / d x_recon (the forward pass already produced the loss value; sq is not reused)
    if gscale is not None or not getattr(b, "dxr_from_forward", False):
        ops.mse_fwd_bwd(t["x_recon"], x, scratch, gscale=gscale, dY=t["dxr"])
    # decoder
    ops.linear_bwd_w(t["dxr"], t["c2"], A.ACT_NONE, g["to_output.weight"], g["to_output.bias"])
    ops.linear_bwd_x(t["dxr"], p["to_output.weight"], t["dc2_act"], A.ACT_GIVEN, out=t["dc2"])
    ops.linear_bwd_w(t["dc2"], t["c1"], A.ACT_NONE, g["decoder.2.weight"], g["decoder.2.bias"])
    ops.linear_bwd_x(t["dc2"], p["decoder.2.weight"], t["dc1_act"], A.ACT_GIVEN, out=t["dc1"])
    ops.linear_bwd_w(t["dc1"], t["z_q"], A.ACT_NONE, g["decoder.0.weight"], g["decoder.0.bias"])
    ops.linear_bwd_x(t["dc1"], p["decoder.0.weight"], None, A.ACT_NONE, out=t["dzq"])
    # quantizer: codebook rows get recon grad + 0.25*codebook-loss grad; encoder gets 0.25*commitment grad,
    # already multiplied by sigmoid'(.) so it is the gradient at the Lipschitz layer's pre-activation
    if b.sort_ws is not None:
        if not b.sorted_ok:
            ops.code_sort(t["idx"], b.D, b.K, b.sort_ws)
            b.sorted_ok = True
        ops.gather_bwd_sorted(t["z_e"], t["z_q"], t["dzq"], gscale, 0.25, 0.25, False, t["z_e"], b.K, b.sort_ws, t["dzpre"],
                              g["quantizer.codebook"])
    else:
        ops.gather_bwd(t["z_e"], t["z_q"], t["idx"], t["dzq"], gscale, 0.25, 0.25, False, t["z_e"], b.K, t["dzpre"],
                       g["quantizer.codebook"])
    # Lipschitz layer: G = dL/dW_norm, then through min(1, softplus(ci)/||W_i||_1)
    if not getattr(b, "zeroed_by_caller", False):
        t["G"].zero_()
    ops.linear_bwd_w(t["dzpre"], t["a2"], A.ACT_NONE, t["G"], g["to_latent.b"])
    ops.lipschitz_scale_bwd(t["G"], p["to_latent.W"], p["to_latent.ci"], g["to_latent.W"], g["to_latent.ci"])
    ops.linear_bwd_x(t["dzpre"], t["Wn"], t["da2_act"], A.ACT_GIVEN, out=t["da2"])
    # encoder
    ops.linear_bwd_w(t["da2"], t["a1"], A.ACT_NONE, g["encoder.2.weight"], g["encoder.2.bias"])
    ops.linear_bwd_x(t["da2"], p["encoder.2.weight"], t["da1_act"], A.ACT_GIVEN, out=t["da1"])
    ops.linear_bwd_w(t["da1"], x, A.ACT_NONE, g["encoder.0.weight"], g["encoder.0.bias"])


def ema_stats(b: T1Buffers, counts: torch.Tensor, sums: torch.Tensor) -> None:
    """counts[K] += bincount(idx); sums[K,D] += onehot^T z_e  (hq_vae.py:264-271)."""
    if b.sort_ws is not None:
        if not b.sorted_ok:
            ops.code_sort(b.t["idx"], b.D, b.K, b.sort_ws)
            b.sorted_ok = True
        ops.ema_stats_sorted(b.t["z_e"], b.sort_ws, counts, sums)
    else:
        ops.ema_stats(b.t["z_e"], b.t["idx"], counts, sums)
