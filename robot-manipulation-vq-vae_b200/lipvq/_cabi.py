"""ctypes binding of liblipvq.so (the C ABI declared in include/lipvq.h).

The library is the product: there is NO fallback.  If the shared object is missing or a symbol is
absent this module raises at import time, and every wrapper raises RuntimeError when a call
returns a non-zero status.  Build it with ``python -c 'import __graft_entry__ as g; g.build()'``
(or ``make -C robot-manipulation-vq-vae_b200/csrc``).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LIPVQ_LIB", os.path.join(_HERE, "liblipvq.so"))

# enums of include/lipvq.h
OK, ERR_INVALID, ERR_CUDA, ERR_UNSUPPORTED = 0, 1, 2, 3
METRIC_L2, METRIC_L2SQ, METRIC_DOT = 0, 1, 2
ALGO_AUTO, ALGO_SIMT_F32, ALGO_TCGEN05 = 0, 1, 2
ACT_NONE, ACT_RELU, ACT_GELU, ACT_SIGMOID, ACT_GIVEN = 0, 1, 2, 3, 4
EMA_SIMPLE, EMA_LAPLACE, EMA_NORMALIZE = 0, 1, 2
ABI_VERSION = 9

_p, _i, _i64, _f, _d, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double, C.c_size_t

# name -> (restype, argtypes); must list every symbol include/lipvq.h declares (tests check this)
SIGNATURES = {
    "lipvq_abi_version": (_i, []),
    "lipvq_last_error": (C.c_char_p, []),
    "lipvq_launch_count": (C.c_uint64, []),
    "lipvq_nn_search_workspace_bytes": (_sz, [_i64, _i, _i, _i, _i]),
    "lipvq_nn_search_f32": (_i, [_p, _i64, _i, _p, _i, _i, _i, _p, _p, _p, _p, _sz, _p]),
    "lipvq_nn_search_pick_algo": (_i, [_i64, _i, _i, _i]),
    "lipvq_nn_search_prepared_supported": (_i, [_i64, _i, _i, _i]),
    "lipvq_nn_search_prepare_f32": (_i, [_p, _i, _i, _i, _p, _i64, _p, _sz, _p]),
    "lipvq_nn_search_prepared_f32": (_i, [_p, _i64, _i, _p, _i, _i, _p, _p, _p, _sz, _p]),
    "lipvq_l2_normalize_rows_f32": (_i, [_p, _i64, _i, _p, _p]),
    "lipvq_gather_ste_loss_fwd_f32": (_i, [_p, _p, _p, _i64, _i, _i, _p, _p, _p, _p]),
    "lipvq_gather_ste_loss_bwd_f32": (_i, [_p, _p, _p, _p, _p, _f, _f, _i, _p, _i64, _i, _i, _p, _p, _p]),
    "lipvq_ema_stats_f32": (_i, [_p, _p, _i64, _i, _i, _p, _p, _p]),
    "lipvq_code_sort_workspace_bytes": (_sz, [_i64, _i, _i]),
    "lipvq_code_sort_i64": (_i, [_p, _i64, _i, _i, _p, _sz, _p]),
    "lipvq_ema_stats_sorted_f32": (_i, [_p, _i64, _i, _i, _p, _sz, _p, _p, _p]),
    "lipvq_gather_ste_loss_bwd_sorted_f32": (_i, [_p, _p, _p, _p, _f, _f, _i, _p, _i64, _i, _i, _p, _sz, _p, _p, _p]),
    "lipvq_ema_stats_static_max_bytes": (_sz, []),
    "lipvq_ema_stats_static_f32": (_i, [_p, _p, _i64, _i, _i, _p, _p, _p]),
    "lipvq_ema_apply_f32": (_i, [_p, _p, _i, _i, _f, _f, _f, _i, _p, _p, _p, _p, _p]),
    "lipvq_usage_stats_f32": (_i, [_p, _i, _p, _p, _p, _p]),
    "lipvq_codebook_usage_f32": (_i, [_p, _i, _f, _p, _p]),
    "lipvq_linear_fwd_f32": (_i, [_p, _i64, _i, _i, _p, _p, _i, _i, _p, _p]),
    "lipvq_linear_fwd_dual_f32": (_i, [_p, _i64, _i, _i, _p, _p, _i, _i, _p, _p, _p]),
    "lipvq_mlp3_fwd_supported": (_i, [_i64, _i, _i, _i, _i]),
    "lipvq_mlp3_fwd_f32": (_i, [_p, _i64, _i, _p, _p, _i, _i, _p, _p, _i, _i, _p, _p, _i, _i, _p, _p]),
    "lipvq_mlp3_fwd_train_f32": (_i, [_p, _i64, _i, _p, _p, _i, _i, _p, _p, _i, _i, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p]),
    "lipvq_mlp3_fwd_bf16_f32": (_i, [_p, _i64, _i, _p, _p, _i, _i, _p, _p, _i, _i, _p, _p, _i, _i, _p, _p]),
    "lipvq_linear_bwd_x_f32": (_i, [_p, _i64, _i, _p, _i, _p, _i, _p, _p]),
    "lipvq_linear_bwd_w_f32": (_i, [_p, _i64, _i, _p, _i, _i, _p, _p, _p]),
    "lipvq_act_bwd_from_output_f32": (_i, [_p, _p, _i, _i64, _p, _p]),
    "lipvq_lipschitz_scale_fwd_f32": (_i, [_p, _p, _i, _i, _p, _p, _p]),
    "lipvq_lipschitz_scale_bwd_f32": (_i, [_p, _p, _p, _i, _i, _p, _p, _p]),
    "lipvq_conv_pack_weights_f32": (_i, [_p, _i, _i, _p, _p, _p]),
    "lipvq_conv_unpack_grad_f32": (_i, [_p, _i, _i, _p, _p]),
    "lipvq_dilated_conv_fwd_f32": (_i, [_p, _i64, _i, _i, _p, _p, _i, _i, _f, C.c_uint32, _p, _p, _p]),
    "lipvq_dilated_conv_bwd_x_f32": (_i, [_p, _p, _i64, _i, _i, _p, _i, _i, _f, _p, _p]),
    "lipvq_dilated_conv_bwd_w_f32": (_i, [_p, _p, _p, _i64, _i, _i, _i, _i, _f, _p, _p, _p]),
    "lipvq_layernorm_fwd_f32": (_i, [_p, _i64, _i, _p, _p, _f, _p, _p, _p, _p]),
    "lipvq_layernorm_bwd_f32": (_i, [_p, _p, _p, _p, _p, _i64, _i, _p, _p, _p, _p]),
    "lipvq_l2_normalize_rows_bwd_f32": (_i, [_p, _p, _i64, _i, _p, _p]),
    "lipvq_affine_f32": (_i, [_p, _f, _f, _i64, _p, _p]),
    "lipvq_mse_fwd_bwd_f32": (_i, [_p, _p, _i64, _p, _i, _p, _p, _p]),
    "lipvq_t1_loss_finalize": (_i, [_p, _d, _d, _f, _f, _p, _p]),
    "lipvq_adamw_step_f32": (_i, [_p, _p, _p, _p, _i64, _i, _p, _f, _f, _f, _f, _f, _f, _p]),
    "lipvq_counter_add_i32": (_i, [_p, _i, _p]),
    "lipvq_allreduce_oneshot_ctas": (_i, []),
    "lipvq_allreduce_oneshot_flag_bytes": (C.c_size_t, [_i]),
    "lipvq_allreduce_oneshot_f32": (_i, [_p, _i, _i, _i64, _i64, _p, _p, _p]),
    "lipvq_allreduce_oneshot_deferred_f32": (_i, [_p, _i, _i, _i64, _i64, _p, _p, _p]),
    "lipvq_allreduce_oneshot_clear_f32": (_i, [_p, _i, _i, _i64, _i64, _p, _p]),
    "lipvq_window_gather_f32": (_i, [_p, _p, _p, _p, _p, _p, _i64, _i, _i, _i, _i, _p, _p, _p, _p, _p]),
    "lipvq_bin_embed_fwd_f32": (_i, [_p, _i64, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p]),
    "lipvq_bin_embed_bwd_f32": (_i, [_p, _p, _i64, _i, _i, _i, _p, _p]),
}


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"liblipvq.so not found at {LIB_PATH}: the CUDA extension is required (no CPU fallback). "
            "Build it with `python -c 'import __graft_entry__ as g; g.build()'`.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:  # pragma: no cover
            raise ImportError(f"{LIB_PATH} does not export {name}; rebuild the extension") from e
        fn.restype = res
        fn.argtypes = args
    v = lib.lipvq_abi_version()
    if v != ABI_VERSION:
        raise ImportError(f"{LIB_PATH}: ABI version {v}, binding expects {ABI_VERSION}; rebuild the extension")
    return lib


lib = _load()


def last_error() -> str:
    return (lib.lipvq_last_error() or b"").decode("utf-8", "replace")


def launch_count() -> int:
    return int(lib.lipvq_launch_count())


def check(status: int, what: str) -> None:
    if status != OK:
        raise RuntimeError(f"{what} failed (status {status}): {last_error()}")
