"""Tensor-parallel layout of the attention path: shard by heads, no data-path collective.

The reference gives every rank Hq/tp query heads and Hkv/tp KV heads (models/qwen3.py:28-37,
layers/linear.py:92-123) and a KV cache holding only its own KV heads (engine/model_runner.py:123,
137-144); block tables, slot mapping and context lens are replicated. GQA groups never straddle
ranks because both head counts are divided by tp. The only collective near the path is the
existing all_reduce after o_proj (layers/linear.py:146), which stays NCCL and is outside it.
So the N>1 story here is: each rank runs the same kernels on its head slice; the only
cross-rank operation in our harnesses is the max-over-ranks of the measured time.
"""
from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class HeadShard:
    num_q_heads: int     # rank-local Hq
    num_kv_heads: int    # rank-local Hkv
    q_head_start: int    # first global query head of this rank
    kv_head_start: int   # first global KV head of this rank

    @property
    def group(self) -> int:
        return self.num_q_heads // self.num_kv_heads


def shard_heads(total_q_heads: int, total_kv_heads: int, world_size: int, rank: int) -> HeadShard:
    """Rank-local head counts and offsets; same divisibility asserts as qwen3.py:30,33."""
    assert 0 <= rank < world_size
    assert total_q_heads % world_size == 0
    assert total_kv_heads % world_size == 0
    hq = total_q_heads // world_size
    hkv = total_kv_heads // world_size
    return HeadShard(hq, hkv, rank * hq, rank * hkv)


def max_over_ranks(value: float) -> float:
    """MAX-reduce a host float over the default process group (identity when not initialised).
    Works for both NCCL (value travels through a device tensor) and gloo (CPU tensor)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
