"""Output-row tensor parallelism for the Vec-LUT MUL_MAT path: one process per GPU, torch.distributed for plumbing.

The reference splits the output rows of one MUL_MAT over CPU threads (ggml/src/ggml-cpu/ggml-cpu.c:1413-1418: thread
``ith`` computes rows ``[ith*N/nth, (ith+1)*N/nth)``).  The same split generalises to GPUs: rank g holds the rows
``plan.rows(g)`` of every weight (a strided column slice of the N-fastest GGUF packing, cut on the host once at load by
``vlut_shard_rows_host``), quantises the replicated activations, computes its f32 slice ``[T][rows]`` and the slices
are all-gathered into ``[T][N]`` on every rank -- the one exchange step of the path.

Two gather implementations:
  * ``gather="nccl"``   torch.distributed.all_gather_into_tensor (NCCL over NVLink5/NVSwitch; gloo on CPU in tests) + a
                        permute copy from ``[G][T][rows]`` to ``[T][N]``                      -- the baseline
  * ``gather="fused"``  the GEMM epilogue stores every output tile straight into the ``[T][N]`` buffer of every rank:
                        through the NVSwitch multicast address of a symmetric-memory allocation (one ``multimem.st``
                        per element, replicated by the switch) or, without multicast, through peer pointers; followed
                        by one symmetric-memory barrier.  No separate collective kernel, no intermediate buffer.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import os

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class RowPlan:
    """Rows of an [N x K] weight owned by each of ``world`` ranks.  Shards are multiples of ``align`` rows (the kernels'
    128-row strip) except possibly the last, and differ by at most one strip."""
    N: int
    world: int
    align: int = 128

    def bounds(self) -> list[int]:
        strips = -(-self.N // self.align)
        base, extra = divmod(strips, self.world)
        b = [0]
        for g in range(self.world):
            b.append(min(self.N, b[-1] + (base + (1 if g < extra else 0)) * self.align))
        b[-1] = self.N
        return b

    def rows(self, rank: int) -> tuple[int, int]:
        b = self.bounds()
        return b[rank], b[rank + 1] - b[rank]

    def max_rows(self) -> int:
        b = self.bounds()
        return max(b[g + 1] - b[g] for g in range(self.world))


def shard_gguf(type_: int, K: int, N: int, plan: RowPlan, rank: int, gguf: np.ndarray) -> np.ndarray:
    """This rank's rows of a GGUF-packed matrix, still GGUF-packed (host side)."""
    from . import api
    row0, rows = plan.rows(rank)
    if rows == 0:
        return np.empty(0, dtype=np.uint8)
    return api.shard_rows(type_, K, N, row0, rows, gguf)


def all_gather_rows(local: torch.Tensor, plan: RowPlan, group=None) -> torch.Tensor:
    """``local`` f32 [T][rows(rank)] on every rank -> f32 [T][N] on every rank (any backend: nccl or gloo)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    T = local.shape[0]
    assert local.shape[1] == plan.rows(rank)[1], (local.shape, plan.rows(rank))
    mr = plan.max_rows()
    if local.shape[1] != mr:  # ragged last shard: pad to the common width
        padded = local.new_zeros((T, mr))
        padded[:, : local.shape[1]] = local
        local = padded
    flat = local.new_empty((world * T, mr))  # concatenated form: accepted by both NCCL and gloo
    dist.all_gather_into_tensor(flat, local.contiguous(), group=group)
    gathered = flat.view(world, T, mr)
    out = local.new_empty((T, plan.N))
    b = plan.bounds()
    for g in range(world):
        out[:, b[g]:b[g + 1]] = gathered[g, :, : b[g + 1] - b[g]]
    return out


class RowShardedWeight:
    """A vlut weight row-sharded over the ranks of a process group (one CUDA device per rank)."""

    def __init__(self, ctx, type_: int, K: int, N: int, gguf: np.ndarray | None = None, group=None,
                 local_gguf=None):
        self.ctx, self.type, self.K, self.N, self.group = ctx, type_, K, N, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.plan = RowPlan(N, self.world)
        self.row0, self.rows = self.plan.rows(self.rank)
        if local_gguf is None:
            local_gguf = shard_gguf(type_, K, N, self.plan, self.rank, gguf)
        self.local = ctx.upload(type_, K, self.rows, local_gguf)
        self._symm = {}
        self._turn = {}

    # ---- baseline: NCCL all-gather ----------------------------------------------------------------------
    def mul_mat(self, x: torch.Tensor, w_scale: float = 1.0) -> torch.Tensor:
        local = self.ctx.mul_mat(self.local, x, w_scale=w_scale)
        if self.world == 1:
            return local
        return all_gather_rows(local, self.plan, self.group)

    # ---- fused: epilogue stores into every rank's output through symmetric memory ---------------------------
    N_BUF = 2  # result buffers per token count, used in rotation

    def symm_out(self, T: int, slot: int = 0):
        """[T][N] f32 symmetric-memory output buffer (same virtual layout on every rank) + its rendezvous handle."""
        key = (T, slot)
        if key not in self._symm:
            import torch.distributed._symmetric_memory as symm_mem
            buf = symm_mem.empty((T, self.N), dtype=torch.float32, device=self.ctx.device)
            hdl = symm_mem.rendezvous(buf, self.group if self.group is not None else dist.group.WORLD)
            self._symm[key] = (buf, hdl)
        return self._symm[key]

    @staticmethod
    def flag_ptrs(hdl) -> list[int]:
        """Per-rank flag arrays for vlut_barrier: the tail of the allocation's signal pad (zero-initialised by torch,
        far from the words torch's own barrier uses)."""
        size = int(hdl.signal_pad_size)
        return [int(p) + size - 256 for p in hdl.signal_pad_ptrs]

    def mul_mat_fused(self, x: torch.Tensor, w_scale: float = 1.0) -> torch.Tensor:
        """Row-sharded MUL_MAT with the all-gather fused into the GEMM epilogue.  The result buffers of one token count
        rotate (N_BUF = 2) and every call ends with the cross-rank barrier that publishes its stores, so a fast rank can
        never overwrite a buffer a slower peer still reads: rank A reaches the call that rewrites buffer b only after the
        barrier of the call in between, which every peer joins after (in stream order) all its reads of b's previous
        contents.  A RESULT IS VALID UNTIL THE SECOND NEXT CALL WITH THE SAME TOKEN COUNT."""
        T = x.shape[0]
        if self.world == 1:
            return self.ctx.mul_mat(self.local, x, w_scale=w_scale)
        slot = self._turn.get(T, 0)
        self._turn[T] = (slot + 1) % self.N_BUF
        buf, hdl = self.symm_out(T, slot)
        # peer pointers by default (the CTA-pair kernel scatters staged tiles with bulk async copies);
        # VLUT_SHARD_MULTICAST=1: NVSwitch multicast stores (multimem.st) instead
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0) if os.environ.get("VLUT_SHARD_MULTICAST", "0") == "1" else 0
        peers = [int(p) for p in hdl.buffer_ptrs]
        if os.environ.get("VLUT_SHARD_PUSH", "0") == "1" and not mc:
            # local GEMM + the concurrent copy kernel (push.cu); falls back to the fused scatter for T <= 64
            self.ctx.mul_mat_scatter_push(self.local, x, peers=peers, self_index=self.rank, ldo=self.N, col0=self.row0,
                                          w_scale=w_scale)
        else:
            self.ctx.mul_mat_scatter(self.local, x, peers=peers, multicast_ptr=mc, ldo=self.N, col0=self.row0,
                                     w_scale=w_scale)
        self.ctx.barrier(self.flag_ptrs(hdl), self.rank)  # every rank's stores have landed everywhere
        return buf
