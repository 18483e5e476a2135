"""Data-parallel reduction over NVLink peer memory (csrc/allreduce.cu) against the NCCL all-reduce: needs two GPUs."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_peer_allreduce_matches_nccl_two_ranks():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "tests", "dp_peer_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "PEER_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_module_on_second_device_with_current_device_zero():
    """The reference's nn.Modules work on any device after .to(): so must these -- kernels, streams and the per-device
    caches (SM count, shared-memory opt-in) follow the TENSORS' device, not torch's current device."""
    import lipvq as L
    import lipvq_oracle as O
    torch.cuda.set_device(0)
    torch.manual_seed(0)
    m = L.LLFQVAE_V4(12, 64, 512)
    p = {k: v.clone() for k, v in m.state_dict().items()}
    x = torch.rand(8192, 12, generator=torch.Generator().manual_seed(1)) * 2 - 1
    m.to("cuda:1")
    z, loss = m(x.to("cuda:1"))
    loss.backward()
    assert z.device == torch.device("cuda:1") and torch.cuda.current_device() == 0
    ref = O.t1_forward(p, x, row_chunk=1024)
    assert abs(loss.item() - ref["loss"].item()) <= 1e-4 * abs(ref["loss"].item())
    assert (m.last_indices.cpu() != ref["indices"]).sum().item() <= 4
    with pytest.raises(RuntimeError, match="one device"):
        L.ops.nn_search(torch.rand(64, 8, device="cuda:0"), torch.rand(16, 8, device="cuda:1"))
