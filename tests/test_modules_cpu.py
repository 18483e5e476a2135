"""Host-side checks of the drop-in module surface (no GPU): constructor signatures, state_dict keys/shapes and RNG draw
order of every tokenizer class against the reference's own classes when /root/reference is present (this container),
else against the committed fixtures.  No compute call is made -- CPU tensors are refused by design."""
import importlib.util
import inspect
import os

import pytest
import torch

from conftest import load_golden, sub

REF = "/root/reference"


def _ref(name, rel):
    path = os.path.join(REF, rel)
    if not os.path.exists(path):
        pytest.skip("reference checkout not present on this machine")
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _same_state(a, b):
    sa, sb = a.state_dict(), b.state_dict()
    assert list(sa.keys()) == list(sb.keys())
    for k in sa:
        assert sa[k].shape == sb[k].shape and sa[k].dtype == sb[k].dtype, k
        assert torch.equal(sa[k], sb[k]), f"{k}: initialisation differs (RNG draw order)"


def _same_signature(ours, theirs):
    so, st = inspect.signature(ours.__init__), inspect.signature(theirs.__init__)
    assert [(p.name, p.default) for p in so.parameters.values()] == [(p.name, p.default) for p in st.parameters.values()]


def test_classic_vqvae_matches_reference_ctor():
    import lipvq
    bb = _ref("ref_bb", "robomimic/models/vq_vae/backbone.py")
    _same_signature(lipvq.VQVAE, bb.VQVAE)
    torch.manual_seed(7)
    a = lipvq.VQVAE(7, 32, num_embeddings=128)
    torch.manual_seed(7)
    b = bb.VQVAE(7, 32, num_embeddings=128)
    _same_state(a, b)


def test_hierarchical_vqvae_matches_reference_ctor():
    import lipvq
    hv = _ref("ref_hv", "robomimic/algo/heirarchical_vqvae.py")
    _same_signature(lipvq.HierarchicalVQVAE, hv.HierarchicalVQVAE)
    kw = dict(embed_dim=64, num_subclusters=48, num_clusters=24, num_stages=2, num_layers_per_stage=3)
    torch.manual_seed(7)
    a = lipvq.HierarchicalVQVAE(7, **kw)
    torch.manual_seed(7)
    b = hv.HierarchicalVQVAE(7, **kw)
    _same_state(a, b)
    for attr in ("num_subclusters", "num_clusters", "embed_dim", "beta"):  # read by icl.py:1384-1389 serialize()
        assert getattr(a, attr) == getattr(b, attr)


def test_hq_modules_match_reference_ctor():
    import lipvq
    hq = _ref("ref_hq", "robomimic/models/vq_vae/hq_vae.py")
    _same_signature(lipvq.LFQQuantizerEMA_KMeans, hq.LFQQuantizerEMA_KMeans)
    _same_signature(lipvq.HierarchicalLFQHVQVAE, hq.HierarchicalLFQHVQVAE)
    torch.manual_seed(3)
    a = lipvq.LFQQuantizerEMA_KMeans(64, 32)
    torch.manual_seed(3)
    b = hq.LFQQuantizerEMA_KMeans(64, 32)
    _same_state(a, b)
    torch.manual_seed(3)
    a = lipvq.HierarchicalLFQHVQVAE(7, 32, 16, num_z_codes=64, num_q_codes=32)
    torch.manual_seed(3)
    b = hq.HierarchicalLFQHVQVAE(7, 32, 16, num_z_codes=64, num_q_codes=32)
    _same_state(a, b)


def test_fixture_state_dicts_load_strict():
    import lipvq
    g = load_golden("t2_hvq")
    m = lipvq.HierarchicalVQVAE(7, embed_dim=64, num_subclusters=64, num_clusters=32, num_stages=2, num_layers_per_stage=3)
    m.load_state_dict(sub(g, "sd"), strict=True)
    g = load_golden("t4_classic")
    lipvq.VQVAE(7, 32, num_embeddings=128).load_state_dict(sub(g, "sd"), strict=True)
    g = load_golden("quantizers")
    lipvq.LFQQuantizerEMA_KMeans(64, 32).load_state_dict(sub(g, "lap_pre"), strict=True)


def test_new_modules_refuse_cpu_tensors():
    import lipvq
    with pytest.raises(RuntimeError):
        lipvq.VQVAE(7, 32)(torch.zeros(4, 7))
    with pytest.raises(RuntimeError):
        lipvq.HierarchicalVQVAE(7, embed_dim=64, num_stages=1, num_layers_per_stage=2)(torch.zeros(2, 8, 7))
    q = lipvq.LFQQuantizerEMA_KMeans(8, 4)
    q.initialized = True
    with pytest.raises(RuntimeError):
        q(torch.zeros(4, 4))
