"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/lipvq.h
declares (no compute calls -- there is no GPU here), the ctypes table is complete, and the Python modules keep the
reference's constructor/state_dict surface and refuse CPU tensors (no fallback)."""
import os
import re
import subprocess

import pytest
import torch

from conftest import ROOT, load_golden, sub

HEADER = os.path.join(ROOT, "include", "lipvq.h")


def _declared():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"LIPVQ_API\s+[\w\s\*]+?\b(lipvq_\w+)\s*\(", src)))


def test_header_declares_entry_points():
    names = _declared()
    assert len(names) >= 20 and "lipvq_nn_search_f32" in names and "lipvq_ema_stats_f32" in names


def test_library_exports_every_declared_symbol():
    import lipvq._cabi as A
    out = subprocess.run(["nm", "-D", "--defined-only", A.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\sT\s+(lipvq_\w+)", out))
    declared = set(_declared())
    assert declared <= exported, f"missing exports: {sorted(declared - exported)}"
    assert exported <= declared, f"undeclared exports: {sorted(exported - declared)}"
    assert set(A.SIGNATURES) == declared, "ctypes signature table out of sync with include/lipvq.h"


def test_abi_version_and_host_only_calls():
    import lipvq._cabi as A
    assert A.lib.lipvq_abi_version() == A.ABI_VERSION
    # workspace sizing and algorithm choice are host-side arithmetic: callable without a device
    n1 = A.lib.lipvq_nn_search_workspace_bytes(4096, 64, 512, A.METRIC_L2, A.ALGO_SIMT_F32)
    n2 = A.lib.lipvq_nn_search_workspace_bytes(65536, 64, 512, A.METRIC_L2, A.ALGO_SIMT_F32)
    assert 0 < n1 < n2
    assert A.lib.lipvq_nn_search_pick_algo(80, 208, 1024, A.METRIC_L2) in (A.ALGO_SIMT_F32, A.ALGO_TCGEN05)
    assert A.lib.lipvq_nn_search_workspace_bytes(-1, 64, 512, 0, 0) == 0


def test_error_path_reports_message():
    import lipvq._cabi as A
    # invalid shape is rejected before any CUDA call
    rc = A.lib.lipvq_nn_search_f32(None, 4, 0, None, 8, 0, 0, None, None, None, None, 0, None)
    assert rc == A.ERR_INVALID and "bad shape" in A.last_error()
    with pytest.raises(RuntimeError, match="bad shape"):
        A.check(rc, "nn_search")


def test_modules_keep_reference_surface():
    import lipvq
    for name, (F_, D, K) in {"t1_c1": (7, 64, 512), "t1_real": (12, 208, 128)}.items():
        g = load_golden(name)
        sd = sub(g, "sd")
        torch.manual_seed(0)
        m = lipvq.LLFQVAE_V4(F_, D, K)
        assert list(m.state_dict().keys()) == list(sd.keys())
        for k in sd:  # same RNG draw order as the reference constructor
            assert torch.equal(m.state_dict()[k], sd[k]), k
        m.load_state_dict(sd, strict=True)
    g = load_golden("quantizers")
    q = lipvq.LFQQuantizerEMA(64, 32)
    q.load_state_dict(sub(g, "pre"), strict=True)
    assert q.decay == 0.99 and q.epsilon == 1e-5


def test_cpu_tensors_are_refused():
    import lipvq
    m = lipvq.LLFQVAE_V4(7, 16, 32)
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(4, 7))
    with pytest.raises(RuntimeError, match="CUDA"):
        lipvq.ops.nn_search(torch.zeros(4, 16), torch.zeros(32, 16))


def test_hvq_trainer_has_no_cpu_path():
    import lipvq
    m = lipvq.HierarchicalVQVAE(7, embed_dim=32, num_subclusters=8, num_clusters=4, num_stages=1, num_layers_per_stage=1,
                                kmeans_init=False)
    assert not getattr(m, "_graph_safe", False)      # the module's own forward keeps the reference's host-driven behaviour
    with pytest.raises(RuntimeError, match="CUDA"):
        lipvq.HvqTrainer(m)


def test_missing_library_fails_loudly(tmp_path):
    env = dict(os.environ, LIPVQ_LIB=str(tmp_path / "nope.so"),
               PYTHONPATH=os.path.join(ROOT, "robot-manipulation-vq-vae_b200"))
    r = subprocess.run(["python", "-c", "import lipvq"], env=env, capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU fallback" in r.stderr


def test_host_side_planning_of_the_round2_entry_points():
    """No GPU needed: which shapes take the prepared search / the fused encoder, and that the workspace grows with the
    prepared state (the functions only evaluate geometry)."""
    import ctypes as C
    from lipvq import _cabi as A
    lib = A.lib
    # the tokenizer's own codebook stays resident in shared memory; big codebooks stream and have no prepared form
    assert lib.lipvq_nn_search_prepared_supported(65536, 64, 512, A.METRIC_L2) == 1
    assert lib.lipvq_nn_search_prepared_supported(65536, 64, 65536, A.METRIC_L2) == 0
    assert lib.lipvq_nn_search_prepared_supported(64, 64, 512, A.METRIC_L2) == 0        # tiny batch: FFMA path
    assert lib.lipvq_nn_search_prepared_supported(65536, 256, 4096, A.METRIC_L2) == 0
    ws_tc = lib.lipvq_nn_search_workspace_bytes(65536, 64, 512, A.METRIC_L2, A.ALGO_TCGEN05)
    img = 512 * 64 * 2 + 512 * 32
    assert ws_tc >= 2 * img + 3 * 65536 * 4, ws_tc   # two code images (self-contained call + prepared state) + the queues
    # fused three-layer forward: the encoder's and the decoder's widths fit, a 208-wide output does not (TMEM columns)
    assert lib.lipvq_mlp3_fwd_supported(65536, 12, 64, 128, 64) == 1
    assert lib.lipvq_mlp3_fwd_supported(65536, 64, 64, 128, 12) == 1
    assert lib.lipvq_mlp3_fwd_supported(65536, 12, 64, 128, 208) == 0
    assert lib.lipvq_mlp3_fwd_supported(1024, 12, 64, 128, 64) == 0                       # small batches: FFMA tiles
    # error path without a device: a streamed codebook is refused before any launch
    rc = lib.lipvq_nn_search_prepare_f32(C.c_void_p(16), 65536, 64, A.METRIC_L2, None, 0, C.c_void_p(16), 1 << 30, None)
    assert rc == A.ERR_UNSUPPORTED and b"resident" in lib.lipvq_last_error()
