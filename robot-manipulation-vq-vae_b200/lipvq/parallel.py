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
