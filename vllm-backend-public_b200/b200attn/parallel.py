"""KV-head sharding across the GPUs of one box.

`tensor_parallel_size` comes from the reference's launch contract (reference benchmark-vllm/models.py:11 ->
`--tensor-parallel-size`, benchmark-vllm/vllm_templates.py:55).  One process per GPU (torchrun), rank r of tp owns

    kv heads [r * Hkv/tp, (r+1) * Hkv/tp)        and the q heads that read them [r * Hq/tp, (r+1) * Hq/tp)

Its KV cache is [num_blocks, Hkv/tp, bs, d]; block_tables, seq_lens and slot_mapping are REPLICATED (the same block
ids on every rank), so attention needs **no collective**: the following o_proj is row-parallel and consumes
head-sharded input.  `gather_heads` (an all-gather along the head axis; NCCL on GPUs, gloo in the CPU tests) exists
only for consumers that need unsharded heads, e.g. checking the sharded result against the 1-GPU one.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class HeadShard:
    tp: int
    rank: int
    Hq: int          # global q heads
    Hkv: int         # global kv heads

    def __post_init__(self):
        if self.tp < 1 or not 0 <= self.rank < self.tp:
            raise ValueError(f"rank {self.rank} outside tp {self.tp}")
        if self.Hkv % self.tp or self.Hq % self.Hkv:
            raise ValueError(f"Hkv={self.Hkv} must be divisible by tp={self.tp} and Hq={self.Hq} by Hkv")

    @property
    def Hq_local(self) -> int:
        return self.Hq // self.tp

    @property
    def Hkv_local(self) -> int:
        return self.Hkv // self.tp

    @property
    def q_heads(self) -> slice:
        return slice(self.rank * self.Hq_local, (self.rank + 1) * self.Hq_local)

    @property
    def kv_heads(self) -> slice:
        return slice(self.rank * self.Hkv_local, (self.rank + 1) * self.Hkv_local)

    # ---- views of unsharded tensors (no copies unless the consumer needs contiguity) ----
    def q(self, q_full: torch.Tensor) -> torch.Tensor:
        """[B, Hq, d] -> this rank's [B, Hq/tp, d] (a strided view; b2_paged_decode takes the row stride)."""
        return q_full[:, self.q_heads]

    def kv(self, kv_full: torch.Tensor) -> torch.Tensor:
        """[T, Hkv, d] -> this rank's [T, Hkv/tp, d] (strided view; b2_reshape_and_cache takes the token stride)."""
        return kv_full[:, self.kv_heads]

    def cache_shape(self, num_blocks: int, bs: int, d: int):
        return (num_blocks, self.Hkv_local, bs, d)

    def cache_bytes_per_block(self, bs: int, d: int, kv_bytes: int) -> int:
        """K + V bytes one block costs on this rank: sizes the pool from gpu_memory_utilization (models.py:8)."""
        return 2 * self.Hkv_local * bs * d * kv_bytes


def from_env(Hq: int, Hkv: int) -> HeadShard:
    """Shard of the calling torchrun rank (WORLD_SIZE / RANK), or the trivial shard outside torchrun."""
    if dist.is_available() and dist.is_initialized():
        return HeadShard(dist.get_world_size(), dist.get_rank(), Hq, Hkv)
    return HeadShard(1, 0, Hq, Hkv)


def gather_heads(out_local: torch.Tensor, shard: HeadShard, group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """All-gather [B, Hq/tp, d] over the ranks -> [B, Hq, d] (head axis in rank order).  4 MiB in total at
    BASELINE configs[3]; microseconds over NVLink 5 / NVSwitch, and off the attention path."""
    if shard.tp == 1:
        return out_local
    B, Hl, d = out_local.shape
    if Hl != shard.Hq_local:
        raise ValueError(f"expected {shard.Hq_local} local heads, got {Hl}")
    buf = torch.empty(shard.tp * B, Hl, d, dtype=out_local.dtype, device=out_local.device)   # rank-major rows
    dist.all_gather_into_tensor(buf, out_local.contiguous(), group=group)
    return buf.view(shard.tp, B, Hl, d).permute(1, 0, 2, 3).reshape(B, shard.Hq, d)


def shard_weight(kind: str, full: torch.Tensor, tp: int, rank: int, Hq: int, Hkv: int, d: int, inter: int,
                 vocab_parallel: bool = True) -> torch.Tensor:
    """Rank `rank`'s shard of one decoder weight for tensor-parallel serving (b200attn.model / b200attn.tp).

    kind (the last component of the model's weight names):
      wqkv / bqkv  [(Hq + 2 Hkv) d, hidden] (or its bias)   column-parallel: this rank's q heads | its kv heads | its v heads
      wo           [hidden, Hq d]                           row-parallel: the input columns of this rank's q heads
      wgu          [2 inter, hidden] = gate | up            column-parallel: this rank's slice of gate | of up
      wdown        [hidden, inter]                          row-parallel: the input columns of that slice
      lm_head      [vocab, hidden]                          vocabulary-parallel rows (when vocab_parallel)
    anything else (embedding, norms) is replicated.  Row-parallel outputs are PARTIAL sums: the caller adds them over
    the ranks (csrc/comm.cu or an all-reduce).  Shards are cut from the full tensor, so sum / concatenation over the
    ranks reproduces the tp = 1 model exactly."""
    if tp == 1:
        return full
    Hq_l, Hkv_l, inter_l = Hq // tp, Hkv // tp, inter // tp
    r = rank
    if kind in ("wqkv", "bqkv"):
        q, k, v = full.split([Hq * d, Hkv * d, Hkv * d], dim=0)
        return torch.cat([q[r * Hq_l * d:(r + 1) * Hq_l * d], k[r * Hkv_l * d:(r + 1) * Hkv_l * d],
                          v[r * Hkv_l * d:(r + 1) * Hkv_l * d]])
    if kind == "wo":
        return full[:, r * Hq_l * d:(r + 1) * Hq_l * d].contiguous()
    if kind == "wgu":
        g, u = full.split([inter, inter], dim=0)
        return torch.cat([g[r * inter_l:(r + 1) * inter_l], u[r * inter_l:(r + 1) * inter_l]])
    if kind == "wdown":
        return full[:, r * inter_l:(r + 1) * inter_l].contiguous()
    if kind == "lm_head" and vocab_parallel:
        vs = full.shape[0] // tp
        return full[r * vs:(r + 1) * vs].clone()          # a copy: the full tensor can then be freed
    return full
