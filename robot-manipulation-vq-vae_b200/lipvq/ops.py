"""Tensor-level wrappers over the C ABI: validate device/dtype/contiguity, hand raw pointers and the
current CUDA stream to liblipvq.so.  PyTorch is used for device memory and streams only -- every
arithmetic op here is a hand-written sm_100a kernel; nothing falls back to torch ops or the CPU.
"""
from __future__ import annotations

import threading
from typing import Optional, Tuple

import torch

from . import _cabi as A

lib = A.lib


class _CallState(threading.local):
    dev = None  # device of the operands validated by _chk since the last C call on this thread


_cs = _CallState()


def _stream() -> int:
    """The current stream of the OPERANDS' device (torch's current device may be another one: the reference's
    nn.Modules work on any device after .to(), so do these)."""
    dev = _cs.dev
    return torch.cuda.current_stream(dev).cuda_stream if dev is not None else torch.cuda.current_stream().cuda_stream


def _chk(t: Optional[torch.Tensor], dtype, name: str, allow_none: bool = False) -> int:
    if t is None:
        if allow_none:
            return 0
        _cs.dev = None
        raise ValueError(f"{name} is required")
    err = None
    if not t.is_cuda:
        err = RuntimeError(f"{name} must be a CUDA tensor: lipvq has no CPU path")
    elif t.dtype != dtype:
        err = TypeError(f"{name} must be {dtype}, got {t.dtype}")
    elif not t.is_contiguous():
        err = ValueError(f"{name} must be contiguous")
    elif _cs.dev is not None and t.device != _cs.dev:
        err = RuntimeError(f"{name} is on {t.device} but the other operands are on {_cs.dev}: all operands of a lipvq "
                           "call must live on one device")
    if err is not None:
        _cs.dev = None
        raise err
    _cs.dev = t.device
    return t.data_ptr()


def _call(fn, what: str, *args) -> None:
    """Run one C-ABI entry point with the operands' device current (kernel launches, per-device attribute caches and
    the stream handle all refer to it) and raise on a non-zero status."""
    dev, _cs.dev = _cs.dev, None
    if dev is None or dev.index == torch.cuda.current_device():
        A.check(fn(*args), what)
    else:
        with torch.cuda.device(dev):
            A.check(fn(*args), what)


F32, I64, I32, F64 = torch.float32, torch.int64, torch.int32, torch.float64

_ws_cache = {}


def _workspace(nbytes: int, device) -> torch.Tensor:
    """Grow-only byte workspace per (device, stream): stable addresses for CUDA-graph capture."""
    key = (device.index, _stream())
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


def nn_search(z: torch.Tensor, codebook: torch.Tensor, metric: int = A.METRIC_L2, algo: int = A.ALGO_AUTO,
              want_best: bool = False, want_stats: bool = False, out: Optional[torch.Tensor] = None,
              workspace: Optional[torch.Tensor] = None):
    """argmin_k dist(z_n, e_k) -> int64 [N] (+ exact winning value, + int32[4] stats)."""
    N, D = z.shape
    K, D2 = codebook.shape
    if D != D2:
        raise RuntimeError(f"shape mismatch: z has dim {D}, codebook {D2}")
    zp, cp = _chk(z, F32, "z"), _chk(codebook, F32, "codebook")
    idx = out if out is not None else torch.empty(N, dtype=I64, device=z.device)
    best = torch.empty(N, dtype=F32, device=z.device) if want_best else None
    stats = torch.empty(4, dtype=I32, device=z.device) if want_stats else None
    need = lib.lipvq_nn_search_workspace_bytes(N, D, K, metric, algo)
    ws = workspace if workspace is not None else _workspace(need, z.device)
    if ws.numel() * ws.element_size() < need:
        _cs.dev = None
        raise ValueError(f"workspace too small: {ws.numel() * ws.element_size()} < {need}")
    _call(lib.lipvq_nn_search_f32, "nn_search", zp, N, D, cp, K, metric, algo, _chk(idx, I64, "idx"),
                                    _chk(best, F32, "best", True), _chk(stats, I32, "stats", True),
                                    ws.data_ptr(), ws.numel() * ws.element_size(), _stream())
    res = (idx,)
    if want_best:
        res += (best,)
    if want_stats:
        res += (stats,)
    return res[0] if len(res) == 1 else res


def search_prepared_supported(N: int, D: int, K: int, metric: int = A.METRIC_L2) -> bool:
    """True when (N, D, K) takes the prepare / prepared pair (codebook resident in shared memory, tensor path)."""
    return bool(lib.lipvq_nn_search_prepared_supported(N, D, K, metric))


def nn_search_prepare(codebook: torch.Tensor, workspace: torch.Tensor, metric: int = A.METRIC_L2,
                      center: Optional[torch.Tensor] = None) -> None:
    """Prepare ``workspace`` for nn_search_prepared on this codebook (once per codebook VERSION).  ``center``: rows
    [n, D] whose mean centres the contraction (a batch of latents); None: the codebook's own rows."""
    K, D = codebook.shape
    if center is not None and center.shape[-1] != D:
        raise RuntimeError(f"shape mismatch: center has dim {center.shape[-1]}, codebook {D}")
    n = 0 if center is None else center.numel() // D
    _chk(workspace, torch.uint8, "workspace")
    _call(lib.lipvq_nn_search_prepare_f32, "nn_search_prepare", _chk(codebook, F32, "codebook"), K, D, metric,
          _chk(center, F32, "center", True) if n > 0 else None, n, workspace.data_ptr(),
          workspace.numel() * workspace.element_size(), _stream())


def nn_search_prepared(z: torch.Tensor, codebook: torch.Tensor, workspace: torch.Tensor, metric: int = A.METRIC_L2,
                       out: Optional[torch.Tensor] = None, want_stats: bool = False):
    """nn_search on a workspace that nn_search_prepare filled for this codebook (which must not have changed since)."""
    N, D = z.shape
    K, D2 = codebook.shape
    if D != D2:
        raise RuntimeError(f"shape mismatch: z has dim {D}, codebook {D2}")
    idx = out if out is not None else torch.empty(N, dtype=I64, device=z.device)
    stats = torch.empty(4, dtype=I32, device=z.device) if want_stats else None
    _chk(workspace, torch.uint8, "workspace")
    _call(lib.lipvq_nn_search_prepared_f32, "nn_search_prepared", _chk(z, F32, "z"), N, D, _chk(codebook, F32, "codebook"),
          K, metric, _chk(idx, I64, "idx"), _chk(stats, I32, "stats", True), workspace.data_ptr(),
          workspace.numel() * workspace.element_size(), _stream())
    return (idx, stats) if want_stats else idx


def search_workspace_bytes(N: int, D: int, K: int, metric: int = A.METRIC_L2, algo: int = A.ALGO_AUTO) -> int:
    return int(lib.lipvq_nn_search_workspace_bytes(N, D, K, metric, algo))


def pick_algo(N: int, D: int, K: int, metric: int = A.METRIC_L2) -> int:
    return int(lib.lipvq_nn_search_pick_algo(N, D, K, metric))


def l2_normalize_rows(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    D = x.shape[-1]
    rows = x.numel() // D
    y = out if out is not None else torch.empty_like(x)
    _call(lib.lipvq_l2_normalize_rows_f32, "l2_normalize_rows", _chk(x, F32, "x"), rows, D, _chk(y, F32, "y"), _stream())
    return y


def gather_fwd(z, idx, codebook, sqerr: torch.Tensor, z_q=None, z_st=None, want_ste=False):
    N, D = z.shape
    K = codebook.shape[0]
    if z_q is None:
        z_q = torch.empty_like(z)
    if want_ste and z_st is None:
        z_st = torch.empty_like(z)
    _call(lib.lipvq_gather_ste_loss_fwd_f32, "gather_fwd", _chk(z, F32, "z"), _chk(idx, I64, "idx"), _chk(codebook, F32, "codebook"),
                                              N, D, K, _chk(z_q, F32, "z_q"), _chk(z_st, F32, "z_st", True),
                                              _chk(sqerr, F64, "sqerr"), _stream())
    return z_q, z_st


def gather_bwd(z, z_q, idx, dzq_up, gscale, w_commit: float, w_codebook: float, ste: bool, sigmoid_out, K: int,
               dz_out, dcodebook):
    N, D = z.shape
    _call(lib.lipvq_gather_ste_loss_bwd_f32, "gather_bwd", _chk(z, F32, "z"), _chk(z_q, F32, "z_q"), _chk(idx, I64, "idx"),
                                              _chk(dzq_up, F32, "dzq_up", True), _chk(gscale, F32, "gscale", True),
                                              w_commit, w_codebook, int(ste), _chk(sigmoid_out, F32, "sigmoid_out", True),
                                              N, D, K, _chk(dz_out, F32, "dz_out", True),
                                              _chk(dcodebook, F32, "dcodebook", True), _stream())


def ema_stats(z, idx, counts, sums):
    N, D = z.shape
    K = counts.numel()
    _call(lib.lipvq_ema_stats_f32, "ema_stats", _chk(z, F32, "z"), _chk(idx, I64, "idx"), N, D, K, _chk(counts, F32, "counts"),
                                    _chk(sums, F32, "sums"), _stream())


def code_sort_workspace_bytes(N: int, D: int, K: int) -> int:
    """0 when the shape is outside the deterministic path (D % 4, D <= 128, K <= 8192)."""
    return int(lib.lipvq_code_sort_workspace_bytes(N, D, K))


def code_sort(idx, D: int, K: int, workspace):
    """Stable counting sort of the tokens by code into ``workspace`` (uint8, code_sort_workspace_bytes): the input of
    ema_stats_sorted / gather_bwd_sorted.  Re-run whenever ``idx`` changes."""
    _call(lib.lipvq_code_sort_i64, "code_sort", _chk(idx, I64, "idx"), idx.numel(), D, K, _chk(workspace, U8, "workspace"),
          workspace.numel(), _stream())


def ema_stats_sorted(z, workspace, counts, sums):
    N, D = z.shape
    K = counts.numel()
    _call(lib.lipvq_ema_stats_sorted_f32, "ema_stats_sorted", _chk(z, F32, "z"), N, D, K, _chk(workspace, U8, "workspace"),
          workspace.numel(), _chk(counts, F32, "counts"), _chk(sums, F32, "sums"), _stream())


def gather_bwd_sorted(z, z_q, dzq_up, gscale, w_commit: float, w_codebook: float, ste: bool, sigmoid_out, K: int, workspace,
                      dz_out, dcodebook):
    N, D = z.shape
    _call(lib.lipvq_gather_ste_loss_bwd_sorted_f32, "gather_bwd_sorted", _chk(z, F32, "z"), _chk(z_q, F32, "z_q"),
          _chk(dzq_up, F32, "dzq_up", True), _chk(gscale, F32, "gscale", True), w_commit, w_codebook, int(ste),
          _chk(sigmoid_out, F32, "sigmoid_out", True), N, D, K, _chk(workspace, U8, "workspace"), workspace.numel(),
          _chk(dz_out, F32, "dz_out", True), _chk(dcodebook, F32, "dcodebook"), _stream())


def ema_apply(counts, sums, ema_size, ema_sum, codebook, decay: float, eps: float, variant: int, scratch):
    K, D = ema_sum.shape
    _call(lib.lipvq_ema_apply_f32, "ema_apply", _chk(counts, F32, "counts"), _chk(sums, F32, "sums"), K, D, decay, 1.0 - decay, eps,
                                    variant, _chk(ema_size, F32, "ema_size"), _chk(ema_sum, F32, "ema_sum"),
                                    _chk(codebook, F32, "codebook"), _chk(scratch, F64, "scratch"), _stream())


def usage_stats(counts, usage_counts, usage_ma, entropy_ma):
    _call(lib.lipvq_usage_stats_f32, "usage_stats", _chk(counts, F32, "counts"), counts.numel(), _chk(usage_counts, F32, "usage_counts"),
                                      _chk(usage_ma, F32, "usage_ma"), _chk(entropy_ma, F32, "entropy_ma"), _stream())


def codebook_usage(cluster_size, dead_thr: float) -> torch.Tensor:
    out = torch.empty(2, dtype=I32, device=cluster_size.device)
    _call(lib.lipvq_codebook_usage_f32, "codebook_usage", _chk(cluster_size, F32, "cluster_size"), cluster_size.numel(), dead_thr,
                                         _chk(out, I32, "out"), _stream())
    return out


def linear_fwd(x, act_in: int, W, bias, act_out: int, out=None):
    in_dim = x.shape[-1]
    N = x.numel() // in_dim
    out_dim = W.shape[0]
    if W.shape[1] != in_dim:
        raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({N}x{in_dim} and {W.shape[1]}x{out_dim})")
    y = out if out is not None else torch.empty(*x.shape[:-1], out_dim, dtype=F32, device=x.device)
    _call(lib.lipvq_linear_fwd_f32, "linear_fwd", _chk(x, F32, "x"), N, in_dim, act_in, _chk(W, F32, "W"), _chk(bias, F32, "bias", True),
                                     out_dim, act_out, _chk(y, F32, "y"), _stream())
    return y


def linear_fwd_dual(x, act_in: int, W, bias, act_out: int, out, dout):
    """out = act_out(pre), dout = act_out'(pre) with pre = act_in(x) W^T + b: the activation is evaluated once, in the
    producing layer (training forward).  Consumers read ``out`` with act_in = NONE; linear_bwd_x takes ``dout`` with
    act_in = ACT_GIVEN."""
    in_dim = x.shape[-1]
    N = x.numel() // in_dim
    out_dim = W.shape[0]
    if W.shape[1] != in_dim:
        raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({N}x{in_dim} and {W.shape[1]}x{out_dim})")
    _call(lib.lipvq_linear_fwd_dual_f32, "linear_fwd_dual", _chk(x, F32, "x"), N, in_dim, act_in, _chk(W, F32, "W"), _chk(bias, F32, "bias", True),
                                          out_dim, act_out, _chk(out, F32, "y"), _chk(dout, F32, "dy_dpre"), _stream())
    return out, dout


def mlp3_supported(N: int, d0: int, d1: int, d2: int, d3: int) -> bool:
    """True when a three-layer forward chain of these widths runs as ONE launch (mlp3_fwd)."""
    return bool(lib.lipvq_mlp3_fwd_supported(N, d0, d1, d2, d3))


def mlp3_fwd(x, W1, b1, act1: int, W2, b2, act2: int, W3, b3, act3: int, out=None, bf16: bool = False):
    """act3(act2(act1(x W1^T + b1) W2^T + b2) W3^T + b3) in one launch (inference: no hidden activations are stored).
    ``bf16``: activations and weights enter the tensor core as plain bf16 (fp32 accumulation, bias, activation, I/O)."""
    d0 = x.shape[-1]
    N = x.numel() // d0
    d1, d2, d3 = W1.shape[0], W2.shape[0], W3.shape[0]
    if W1.shape[1] != d0 or W2.shape[1] != d1 or W3.shape[1] != d2:
        raise RuntimeError(f"mlp3_fwd: widths do not chain: x {d0}, W1 {tuple(W1.shape)}, W2 {tuple(W2.shape)}, W3 {tuple(W3.shape)}")
    y = out if out is not None else torch.empty(*x.shape[:-1], d3, dtype=F32, device=x.device)
    _call(lib.lipvq_mlp3_fwd_bf16_f32 if bf16 else lib.lipvq_mlp3_fwd_f32, "mlp3_fwd", _chk(x, F32, "x"), N, d0, _chk(W1, F32, "W1"), _chk(b1, F32, "b1", True), d1, act1,
          _chk(W2, F32, "W2"), _chk(b2, F32, "b2", True), d2, act2, _chk(W3, F32, "W3"), _chk(b3, F32, "b3", True), d3, act3,
          _chk(y, F32, "y"), _stream())
    return y


def mlp3_fwd_train(x, W1, b1, act1: int, W2, b2, act2: int, W3, b3, act3: int, out, h1, g1, h2, g2):
    """mlp3_fwd as the forward of a training step: the hidden layers also store their activation (h1, h2) and its
    derivative (g1, g2) -- what two linear_fwd_dual calls store."""
    d0 = x.shape[-1]
    N = x.numel() // d0
    d1, d2, d3 = W1.shape[0], W2.shape[0], W3.shape[0]
    if W1.shape[1] != d0 or W2.shape[1] != d1 or W3.shape[1] != d2:
        raise RuntimeError(f"mlp3_fwd_train: widths do not chain: x {d0}, W1 {tuple(W1.shape)}, W2 {tuple(W2.shape)}, W3 {tuple(W3.shape)}")
    _call(lib.lipvq_mlp3_fwd_train_f32, "mlp3_fwd_train", _chk(x, F32, "x"), N, d0, _chk(W1, F32, "W1"), _chk(b1, F32, "b1", True), d1,
          act1, _chk(W2, F32, "W2"), _chk(b2, F32, "b2", True), d2, act2, _chk(W3, F32, "W3"), _chk(b3, F32, "b3", True), d3, act3,
          _chk(out, F32, "y"), _chk(h1, F32, "h1"), _chk(g1, F32, "g1"), _chk(h2, F32, "h2"), _chk(g2, F32, "g2"), _stream())
    return out


def linear_bwd_x(dY, W, Xpre, act_in: int, out=None):
    out_dim, in_dim = W.shape
    N = dY.numel() // out_dim
    dX = out if out is not None else torch.empty(*dY.shape[:-1], in_dim, dtype=F32, device=dY.device)
    _call(lib.lipvq_linear_bwd_x_f32, "linear_bwd_x", _chk(dY, F32, "dY"), N, out_dim, _chk(W, F32, "W"), in_dim,
                                       _chk(Xpre, F32, "Xpre", act_in == A.ACT_NONE), act_in, _chk(dX, F32, "dX"),
                                       _stream())
    return dX


def linear_bwd_w(dY, X, act_in: int, dW, db):
    """dW += dY^T act_in(X); db += colsum(dY).  dW/db are accumulated into."""
    out_dim, in_dim = dW.shape
    N = dY.numel() // out_dim
    _call(lib.lipvq_linear_bwd_w_f32, "linear_bwd_w", _chk(dY, F32, "dY"), N, out_dim, _chk(X, F32, "X"), in_dim, act_in,
                                       _chk(dW, F32, "dW"), _chk(db, F32, "db", True), _stream())


def act_bwd_from_output(dY, Y, act: int, out=None):
    dX = out if out is not None else torch.empty_like(dY)
    _call(lib.lipvq_act_bwd_from_output_f32, "act_bwd_from_output", _chk(dY, F32, "dY"), _chk(Y, F32, "Y"), act, dY.numel(), _chk(dX, F32, "dX"),
                                              _stream())
    return dX


def lipschitz_scale_fwd(W, ci, W_norm=None, scale=None):
    out_dim, in_dim = W.shape
    if W_norm is None:
        W_norm = torch.empty_like(W)
    _call(lib.lipvq_lipschitz_scale_fwd_f32, "lipschitz_scale_fwd", _chk(W, F32, "W"), _chk(ci, F32, "ci"), out_dim, in_dim,
                                              _chk(scale, F32, "scale", True), _chk(W_norm, F32, "W_norm"), _stream())
    return W_norm


def lipschitz_scale_bwd(G, W, ci, dW, dci):
    out_dim, in_dim = W.shape
    _call(lib.lipvq_lipschitz_scale_bwd_f32, "lipschitz_scale_bwd", _chk(G, F32, "G"), _chk(W, F32, "W"), _chk(ci, F32, "ci"), out_dim, in_dim,
                                              _chk(dW, F32, "dW"), _chk(dci, F32, "dci"), _stream())


def mse_fwd_bwd(y, target, sqerr, gscale=None, act_out: int = A.ACT_NONE, dY=None):
    if y.shape != target.shape:
        raise RuntimeError(f"mse: shape mismatch {tuple(y.shape)} vs {tuple(target.shape)}")
    _call(lib.lipvq_mse_fwd_bwd_f32, "mse_fwd_bwd", _chk(y, F32, "y"), _chk(target, F32, "target"), y.numel(),
                                      _chk(gscale, F32, "gscale", True), act_out, _chk(sqerr, F64, "sqerr"),
                                      _chk(dY, F32, "dY", True), _stream())


def t1_loss_finalize(sq2, n_recon: float, n_lat: float, w_commit: float, w_cb: float, out4):
    _call(lib.lipvq_t1_loss_finalize, "t1_loss_finalize", _chk(sq2, F64, "sq2"), float(n_recon), float(n_lat), w_commit, w_cb,
                                       _chk(out4, F32, "out4"), _stream())


def adamw_step(p, g, m, v, step: int, step_dev=None, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-4,
               grad_scale: float = 1.0):
    _call(lib.lipvq_adamw_step_f32, "adamw_step", _chk(p, F32, "p"), _chk(g, F32, "g"), _chk(m, F32, "m"), _chk(v, F32, "v"), p.numel(),
                                     int(step), _chk(step_dev, I32, "step_dev", True), lr, betas[0], betas[1], eps,
                                     weight_decay, grad_scale, _stream())


def counter_add(ctr, delta: int = 1):
    _call(lib.lipvq_counter_add_i32, "counter_add", _chk(ctr, I32, "ctr"), delta, _stream())


# ---- K8: MS-TCN layers, LayerNorm, normalize backward -------------------------------------------------
U8 = torch.uint8


def conv_pack_weights(W, want_fwd=True, want_bwd=True):
    """W [Eo,Ei,3] -> (Wf [Eo,3,Ei] | None, Wb [3,Eo,Ei] | None)"""
    Eo, Ei, k = W.shape
    if k != 3:
        raise RuntimeError(f"dilated conv kernels support kernel_size 3, got {k}")
    Wf = torch.empty(Eo, 3, Ei, dtype=F32, device=W.device) if want_fwd else None
    Wb = torch.empty(3, Eo, Ei, dtype=F32, device=W.device) if want_bwd else None
    _call(lib.lipvq_conv_pack_weights_f32, "conv_pack_weights", _chk(W, F32, "W"), Eo, Ei, _chk(Wf, F32, "Wf", True), _chk(Wb, F32, "Wb", True),
                                            _stream())
    return Wf, Wb


def conv_unpack_grad(dWf, dW):
    Eo, Ei, _ = dW.shape
    _call(lib.lipvq_conv_unpack_grad_f32, "conv_unpack_grad", _chk(dWf, F32, "dWf"), Eo, Ei, _chk(dW, F32, "dW"), _stream())


def dilated_conv_fwd(x, T: int, Wf, bias, dilation: int, drop_p: float = 0.0, seed: int = 0):
    """x [N,E] token-major (row n has time index n % T) -> (out [N,E], keep-mask uint8 [N,E])"""
    N, Ei = x.shape
    Eo = Wf.shape[0]
    out = torch.empty(N, Eo, dtype=F32, device=x.device)
    mask = torch.empty(N, Eo, dtype=U8, device=x.device)
    _call(lib.lipvq_dilated_conv_fwd_f32, "dilated_conv_fwd", _chk(x, F32, "x"), N, T, Ei, _chk(Wf, F32, "Wf"), _chk(bias, F32, "bias"), Eo,
                                           dilation, drop_p, seed & 0xFFFFFFFF, _chk(out, F32, "out"), _chk(mask, U8, "mask"),
                                           _stream())
    return out, mask


def dilated_conv_bwd_x(dout, mask, T: int, Wb, dilation: int, drop_p: float = 0.0):
    N, Eo = dout.shape
    Ei = Wb.shape[2]
    dx = torch.empty(N, Ei, dtype=F32, device=dout.device)
    _call(lib.lipvq_dilated_conv_bwd_x_f32, "dilated_conv_bwd_x", _chk(dout, F32, "dout"), _chk(mask, U8, "mask"), N, T, Eo, _chk(Wb, F32, "Wb"),
                                             Ei, dilation, drop_p, _chk(dx, F32, "dx"), _stream())
    return dx


def dilated_conv_bwd_w(dout, mask, x, T: int, dilation: int, dWf, db, drop_p: float = 0.0):
    N, Eo = dout.shape
    Ei = x.shape[1]
    _call(lib.lipvq_dilated_conv_bwd_w_f32, "dilated_conv_bwd_w", _chk(dout, F32, "dout"), _chk(mask, U8, "mask"), _chk(x, F32, "x"), N, T, Eo, Ei,
                                             dilation, drop_p, _chk(dWf, F32, "dWf"), _chk(db, F32, "db", True), _stream())


def layernorm_fwd(x, w, b, eps: float):
    E = x.shape[-1]
    rows = x.numel() // E
    y = torch.empty_like(x)
    mean = torch.empty(rows, dtype=F32, device=x.device)
    rstd = torch.empty(rows, dtype=F32, device=x.device)
    _call(lib.lipvq_layernorm_fwd_f32, "layernorm_fwd", _chk(x, F32, "x"), rows, E, _chk(w, F32, "w"), _chk(b, F32, "b"), eps,
                                        _chk(y, F32, "y"), _chk(mean, F32, "mean"), _chk(rstd, F32, "rstd"), _stream())
    return y, mean, rstd


def layernorm_bwd(dy, x, mean, rstd, w, dw, db):
    E = x.shape[-1]
    rows = x.numel() // E
    dx = torch.empty_like(x)
    _call(lib.lipvq_layernorm_bwd_f32, "layernorm_bwd", _chk(dy, F32, "dy"), _chk(x, F32, "x"), _chk(mean, F32, "mean"), _chk(rstd, F32, "rstd"),
                                        _chk(w, F32, "w"), rows, E, _chk(dx, F32, "dx"), _chk(dw, F32, "dw", True),
                                        _chk(db, F32, "db", True), _stream())
    return dx


def l2_normalize_rows_bwd(x, dy):
    D = x.shape[-1]
    dx = torch.empty_like(x)
    _call(lib.lipvq_l2_normalize_rows_bwd_f32, "l2_normalize_rows_bwd", _chk(x, F32, "x"), _chk(dy, F32, "dy"), x.numel() // D, D, _chk(dx, F32, "dx"),
                                                _stream())
    return dx


def affine(x, a: float, b: float, out=None):
    y = out if out is not None else torch.empty_like(x)
    _call(lib.lipvq_affine_f32, "affine", _chk(x, F32, "x"), a, b, x.numel(), _chk(y, F32, "y"), _stream())
    return y


def window_gather(data, demo_row0, demo_len, demo_seq0, sample_demo, sample_index, n_frame_stack: int, seq_length: int,
                  pad_frame_stack: bool, offset=None, scale=None, want_mask: bool = False):
    """[B, T, A] windows (+ [B, T] bool pad mask) of the HBM-resident demonstrations ``data`` [L_total, A]."""
    B, A = sample_index.numel(), data.shape[1]
    T = n_frame_stack - 1 + seq_length
    out = torch.empty(B, T, A, dtype=F32, device=data.device)
    mask = torch.empty(B, T, dtype=U8, device=data.device) if want_mask else None
    _call(lib.lipvq_window_gather_f32, "window_gather", _chk(data, F32, "data"), _chk(demo_row0, I64, "demo_row0"),
          _chk(demo_len, I32, "demo_len"), _chk(demo_seq0, I64, "demo_seq0"), _chk(sample_demo, I32, "sample_demo"),
          _chk(sample_index, I64, "sample_index"), B, A, n_frame_stack, seq_length, int(pad_frame_stack),
          _chk(offset, F32, "offset", True), _chk(scale, F32, "scale", True), _chk(out, F32, "out"), _chk(mask, U8, "mask", True),
          _stream())
    return (out, mask.bool()) if want_mask else out


def bin_embed_fwd(actions, running_min, running_max, tables, update_stats: bool, want_idx: bool = True):
    N, A = actions.shape
    _, nb, E = tables.shape
    out = torch.empty(N, A * E, dtype=F32, device=actions.device)
    idx = torch.empty(N, A, dtype=I64, device=actions.device) if want_idx else None
    _call(lib.lipvq_bin_embed_fwd_f32, "bin_embed_fwd", _chk(actions, F32, "actions"), N, A, nb, E, int(update_stats),
          _chk(running_min, F32, "running_min"), _chk(running_max, F32, "running_max"), _chk(tables, F32, "tables"),
          _chk(out, F32, "out"), _chk(idx, I64, "idx", True), _stream())
    return out, idx


def bin_embed_bwd(dout, idx, num_bins: int, E: int):
    N, A = idx.shape
    dt = torch.zeros(A, num_bins, E, dtype=F32, device=dout.device)
    _call(lib.lipvq_bin_embed_bwd_f32, "bin_embed_bwd", _chk(dout, F32, "dout"), _chk(idx, I64, "idx"), N, A, num_bins, E,
          _chk(dt, F32, "dtables"), _stream())
    return dt
