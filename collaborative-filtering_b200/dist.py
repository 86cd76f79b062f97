"""Row sharding of the two ALS half-steps across the GPUs of one box (§8e).

One process per GPU (torchrun).  Each half-step's rows are independent
(als.py:414-433; 436-466 without the graph), so rank r solves a contiguous
slab of users, then of items; the freshly solved slab of factors and biases is
exchanged with one NCCL all-gather over NVLink, in place inside the replica
every rank keeps (gathers touch arbitrary rows, so U/V/b_u/b_i are replicated).
Slabs have equal row counts (the replica is padded to a multiple of the world
size) so a single `all_gather_into_tensor` moves them.  Residual sums are
combined with one float64 all-reduce per iteration.

`torch.distributed` is plumbing only; with a `gloo` group on CPU tensors the
same code paths run in the CPU test-suite.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as td


@dataclass
class Shard:
    rank: int = 0
    world: int = 1
    group: object = None

    @property
    def active(self) -> bool:
        return self.world > 1

    @staticmethod
    def from_env():
        if td.is_available() and td.is_initialized():
            return Shard(td.get_rank(), td.get_world_size(), None)
        return Shard()

    # ---- partition ---------------------------------------------------------
    def slab_rows(self, n_rows: int) -> int:
        """Rows per rank (replicas are padded to world * slab_rows)."""
        return (n_rows + self.world - 1) // self.world

    def padded_rows(self, n_rows: int) -> int:
        return self.slab_rows(n_rows) * self.world

    def my_range(self, n_rows: int):
        """[lo, hi) of real rows owned by this rank (hi clipped to n_rows)."""
        s = self.slab_rows(n_rows)
        lo = min(self.rank * s, n_rows)
        return lo, min(lo + s, n_rows)

    # ---- collectives -------------------------------------------------------
    def all_gather_rows(self, full: torch.Tensor, n_rows: int) -> None:
        """In-place all-gather of each rank's slab of `full` (rows padded)."""
        if not self.active:
            return
        s = self.slab_rows(n_rows)
        assert full.shape[0] == s * self.world, "replica must be padded to world * slab_rows"
        mine = full[self.rank * s:(self.rank + 1) * s]
        if full.is_cuda:
            td.all_gather_into_tensor(full, mine, group=self.group)
        else:  # gloo: list form
            parts = [full[r * s:(r + 1) * s] for r in range(self.world)]
            gathered = [torch.empty_like(mine) for _ in range(self.world)]
            td.all_gather(gathered, mine.contiguous(), group=self.group)
            for dst, src in zip(parts, gathered):
                dst.copy_(src)

    def all_reduce_sum(self, t: torch.Tensor) -> None:
        if self.active:
            td.all_reduce(t, op=td.ReduceOp.SUM, group=self.group)

    def barrier(self) -> None:
        if self.active:
            td.barrier(group=self.group)
