"""lipvq_ema_stats_static_f32: the allocation-free entry point (scratch in a __device__ array inside the library).

No counterpart in the reference.  Called through the raw C-ABI (ctypes,
plain pointers) and compared with the existing entry points on two calls in a row with DIFFERENT inputs, so a stale
scratch from the first call would show in the second.  Collected last on purpose
(file name).
"""
import ctypes as C
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "robot-manipulation-vq-vae_b200"))

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def test_ema_stats_static_matches_existing_entry_points_twice():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from lipvq import _cabi as A
    lib = A.lib
    dev = torch.device("cuda:0")
    assert int(lib.lipvq_ema_stats_static_max_bytes()) == 8 << 20
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    # call 1: the bench shape, spread codes; call 2: another shape, collapsed codes + out-of-range indices (clamped)
    cases = [(65536, 64, 512, "spread", 11), (8192 + 37, 32, 100, "collapsed", 12)]
    with torch.cuda.device(dev):
        for N, D, K, kind, seed in cases:
            g = torch.Generator().manual_seed(seed)
            z = torch.rand(N, D, generator=g).to(dev)
            if kind == "spread":
                idx = torch.randint(0, K, (N,), generator=g)
            else:
                idx = torch.randint(0, 5, (N,), generator=g) * 20
                idx[:7] = torch.tensor([-3, K, K + 9, -1, 0, K - 1, 2 * K])
            idc = idx.clamp(0, K - 1)
            idx = idx.to(dev)
            assert int(lib.lipvq_code_sort_workspace_bytes(N, D, K)) <= 8 << 20
            # the new entry point
            counts, sums = torch.zeros(K, device=dev), torch.zeros(K, D, device=dev)
            rc = lib.lipvq_ema_stats_static_f32(_ptr(z), _ptr(idx), N, D, K, _ptr(counts), _ptr(sums), stream)
            assert rc == 0, A.last_error()
            # existing entry point 1: lipvq_ema_stats_f32 (atomic accumulation) -- the tolerance of test_gpu_kernels.py
            c_ref, s_ref = torch.zeros(K, device=dev), torch.zeros(K, D, device=dev)
            rc = lib.lipvq_ema_stats_f32(_ptr(z), _ptr(idx.clamp(0, K - 1)), N, D, K, _ptr(c_ref), _ptr(s_ref), stream)
            assert rc == 0, A.last_error()
            torch.cuda.synchronize(dev)
            assert torch.equal(counts.cpu(), c_ref.cpu()), kind
            assert torch.equal(counts.cpu(), torch.bincount(idc, minlength=K).float()), kind
            assert torch.allclose(sums.cpu(), s_ref.cpu(), rtol=1e-4, atol=1e-3), kind
            ref64 = torch.zeros(K, D, dtype=torch.float64).index_add_(0, idc, z.cpu().double())
            assert torch.allclose(sums.cpu().double(), ref64, rtol=1e-4, atol=1e-3), kind
            # existing entry point 2: the caller-workspace sequence computes the same ordered sums -> bit-identical
            ws = torch.empty(int(lib.lipvq_code_sort_workspace_bytes(N, D, K)), dtype=torch.uint8, device=dev)
            c2, s2 = torch.zeros(K, device=dev), torch.zeros(K, D, device=dev)
            assert lib.lipvq_code_sort_i64(_ptr(idx), N, D, K, _ptr(ws), ws.numel(), stream) == 0, A.last_error()
            assert lib.lipvq_ema_stats_sorted_f32(_ptr(z), N, D, K, _ptr(ws), ws.numel(), _ptr(c2), _ptr(s2), stream) == 0, A.last_error()
            torch.cuda.synchronize(dev)
            assert torch.equal(counts.cpu(), c2.cpu()) and torch.equal(sums.cpu(), s2.cpu()), kind
        # beyond the built-in scratch: rejected with a clear error, nothing launched
        N, D, K = 1 << 20, 128, 8192
        assert int(lib.lipvq_code_sort_workspace_bytes(N, D, K)) > 8 << 20
        dummy = torch.zeros(16, device=dev)
        rc = lib.lipvq_ema_stats_static_f32(_ptr(dummy), _ptr(dummy), N, D, K, _ptr(dummy), _ptr(dummy), stream)
        assert rc != 0
        assert "scratch" in A.last_error() and "8388608" in A.last_error()
