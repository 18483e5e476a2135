"""CPU: oracle restatements of the caller-side pieces against the reference's own runs
(tests/golden/dataset_windows.npz, icrt_policy.npz bin_* entries; oracle/gen_golden.py)."""
import numpy as np
import torch

import lipvq_oracle as O
from conftest import load_golden, sub


def test_sequence_windows_vs_reference_get_item():
    g = load_golden("dataset_windows")
    demos_all = [g[f"demo.demo_{i}"].numpy() for i in range(len(g["lens"]))]
    for ci in range(4):
        T, fs, pad_seq, pad_fs, norm = (int(v) for v in g[f"c{ci}.cfg"])
        demos = [demos_all[int(d)] for d in g[f"c{ci}.demos"]]
        off, sc = (g["stats_offset"].numpy(), g["stats_scale"].numpy()) if norm else (None, None)
        n = g[f"c{ci}.actions"].shape[0]
        acts, masks = O.sequence_windows(demos, range(n), T, fs, bool(pad_seq), bool(pad_fs), off, sc)
        assert acts.shape == tuple(g[f"c{ci}.actions"].shape) == (n, fs - 1 + T, 12)
        assert np.array_equal(acts, g[f"c{ci}.actions"].numpy())
        assert np.array_equal(masks, g[f"c{ci}.pad_mask"].numpy())


def test_linspace_restatement_is_torch_linspace():
    rng = np.random.RandomState(1)
    for _ in range(200):
        lo, hi = np.float32(rng.randn()), np.float32(rng.randn() + 2.5)
        for steps in (21, 6, 2):
            assert np.array_equal(O.linspace_f32(lo, hi, steps), torch.linspace(float(lo), float(hi), steps).numpy())


def test_bin_embedding_vs_reference():
    g = load_golden("icrt_policy")
    p, a = sub(g, "bin_sd"), g["ctx_actions"].reshape(-1, 12)
    rmin, rmax = p.pop("running_min"), p.pop("running_max")
    out1, idx1, rmin, rmax = O.bin_embedding_forward(p, a, rmin, rmax)
    assert torch.allclose(out1, g["bin_out1"], rtol=1e-6, atol=1e-7)
    out2, _, rmin, rmax = O.bin_embedding_forward(p, a * 0.5, rmin, rmax)
    assert torch.allclose(out2, g["bin_out2"], rtol=1e-6, atol=1e-7)
    post = sub(g, "bin_post")
    assert torch.equal(rmin, post["running_min"]) and torch.equal(rmax, post["running_max"])
    assert torch.equal(O.bin_discretize(a, rmin, rmax, 20), g["bin_idx"])
