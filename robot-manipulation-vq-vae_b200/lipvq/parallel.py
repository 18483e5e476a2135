"""Data-parallel plumbing of the tokenizer step: batch sharding and the single reduction slab.

The reference is single-process (no torch.distributed anywhere).  The path shards
over tokens: every rank encodes/searches/decodes its own batch shard against replicated weights,
and the only cross-rank quantities are sums -- parameter gradients, EMA counts[K] / sums[K,D]
(robomimic/models/vq_vae/hq_vae.py:264-275) and the loss scalars.  They live in ONE flat fp32 slab
so a step issues exactly one all_reduce(sum); afterwards each rank applies the same EMA + AdamW
update, which keeps replicas bit-identical without a broadcast.

Pure host logic (no CUDA needed): exercised with gloo world_size=2 in tests/test_parallel_cpu.py.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced [begin, end) of rank's share of n_items (first n%world ranks get one more)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(n_items, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


@dataclass
class SlabLayout:
    """[grads (P, padded to 4) | counts (K) | sums (K*D) | losses (4)] -- offsets in floats."""
    names: List[str]
    shapes: List[Tuple[int, ...]]
    K: int
    D: int
    ema: bool

    def __post_init__(self):
        self.sizes = [int(torch.Size(s).numel()) for s in self.shapes]
        self.P = sum(self.sizes)
        self.P4 = (self.P + 3) // 4 * 4  # counts/sums stay 16-byte aligned behind the gradients (float4 atomics)
        self.off_counts = self.P4
        self.off_sums = self.P4 + (self.K if self.ema else 0)
        self.off_losses = self.off_sums + (self.K * self.D if self.ema else 0)
        self.length = self.off_losses + 4
        self.param_offsets: Dict[str, int] = {}
        o = 0
        for n, s in zip(self.names, self.sizes):
            self.param_offsets[n] = o
            o += s

    def grad_view(self, slab: torch.Tensor, name: str) -> torch.Tensor:
        i = self.names.index(name)
        o = self.param_offsets[name]
        return slab[o:o + self.sizes[i]].view(self.shapes[i])

    def counts(self, slab: torch.Tensor) -> torch.Tensor:
        return slab[self.off_counts:self.off_counts + self.K]

    def sums(self, slab: torch.Tensor) -> torch.Tensor:
        return slab[self.off_sums:self.off_sums + self.K * self.D].view(self.K, self.D)

    def losses(self, slab: torch.Tensor) -> torch.Tensor:
        return slab[self.off_losses:self.off_losses + 4]


def all_reduce_slab(slab: torch.Tensor, group: Optional[dist.ProcessGroup] = None) -> int:
    """One sum all-reduce of the whole slab; returns the world size (gradient/loss scale is 1/world,
    EMA counts and sums are true sums)."""
    if not dist.is_initialized():
        return 1
    world = dist.get_world_size(group)
    if world > 1:
        dist.all_reduce(slab, op=dist.ReduceOp.SUM, group=group)
    return world


class PeerSlab:
    """Reduction slab in symmetric (peer-mapped) memory + the one-shot NVLink all-reduce kernel (csrc/allreduce.cu).

    ``slab``  -- this rank's [length] fp32 input (phase A accumulates into it; every peer can read it)
    ``out``   -- [length] fp32, local: sum over ranks after ``all_reduce()`` (same bits on every rank)
    The allocation / handle exchange is torch.distributed._symmetric_memory; the data path is one kernel of ours,
    capturable in a CUDA graph.  Raises if symmetric memory cannot be set up (caller falls back to NCCL)."""

    def __init__(self, length: int, device: torch.device, group: Optional[dist.ProcessGroup] = None):
        import torch.distributed._symmetric_memory as symm_mem

        from ._cabi import lib
        group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.n = (int(length) + 3) // 4 * 4
        ctas = int(lib.lipvq_allreduce_oneshot_ctas())
        flag_floats = int(lib.lipvq_allreduce_oneshot_flag_bytes(self.world)) // 4
        self.buf = symm_mem.empty(self.n + flag_floats, dtype=torch.float32, device=device)
        self.buf.zero_()
        self.hdl = symm_mem.rendezvous(self.buf, group.group_name)
        torch.cuda.synchronize(device)
        dist.barrier(group)  # every rank's flags are zero before anyone raises one
        self.slab = self.buf[:length]
        self._out = torch.zeros(self.n, dtype=torch.float32, device=device)
        self.out = self._out[:length]
        self.epoch = torch.zeros(ctas, dtype=torch.int32, device=device)
        self.peer_ptrs_dev = int(self.hdl.buffer_ptrs_dev)

        # Default: both flag exchanges inside the all-reduce kernel and a plain memset of the slab.  LIPVQ_PEER_DEFER=1:
        # the "I have read your slab" wait moves in front of the next step's clear (clear()).  Measured at 2 GPUs: no
        # difference (0.4803 vs 0.4806 ms per step) -- by the time a rank has summed its share the peers' signals are
        # there; what did pay was launching the kernel as a programmatic dependent (0.4855 -> 0.4805 ms).
        import os
        self.deferred = os.environ.get("LIPVQ_PEER_DEFER", "0") == "1"

    def all_reduce(self) -> None:
        from . import _cabi
        dev = self._out.device
        fn = _cabi.lib.lipvq_allreduce_oneshot_deferred_f32 if self.deferred else _cabi.lib.lipvq_allreduce_oneshot_f32
        with torch.cuda.device(dev):
            _cabi.check(fn(self.peer_ptrs_dev, self.rank, self.world, self.n, self.n * 4, self._out.data_ptr(),
                           self.epoch.data_ptr(), torch.cuda.current_stream(dev).cuda_stream), "allreduce_oneshot")

    def clear(self) -> None:
        """Zero the slab for the next step.  With the deferred exchange this is where a rank makes sure that no peer is
        still reading the slab of the last reduction (the only legal way to write the slab again)."""
        if not self.deferred:
            self.slab.zero_()
            return
        from . import _cabi
        dev = self._out.device
        with torch.cuda.device(dev):
            _cabi.check(_cabi.lib.lipvq_allreduce_oneshot_clear_f32(self.peer_ptrs_dev, self.rank, self.world, self.n, self.n * 4,
                                                                    self.epoch.data_ptr(),
                                                                    torch.cuda.current_stream(dev).cuda_stream), "allreduce_clear")

