"""Host-side builders of the per-step attention metadata, as ModelRunner forms it.

Restates reference nanovllm/engine/model_runner.py:152-158 (_prepare_block_tables), :160-204
(_prepare_prefill) and :206-226 (_prepare_decode) for the fields the attention path consumes,
on plain (length, num_cached_tokens, block_table) triples instead of `Sequence` objects
(engine/sequence.py:55-69 for num_blocks / last_block_num_tokens / num_cached_blocks). Used by
the standalone harnesses (tests, bench, smoke) to drive the op exactly as the engine would; under
a real nanovllm the engine's own builders run and this module is not involved. It is HARNESS code
(it lives in tools/, not in the product package): the device-side builders of
byox_vllm_b200/fast_meta.py are checked against it on the CPU and against the reference's own
ModelRunner builders on the GPU box (tests/test_engine_parity.py).

All tensors are built on the CPU (int32, like the reference) and optionally pinned; the caller
moves them with .cuda(non_blocking=True) as the reference does.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch


@dataclass
class SeqState:
    """The part of engine/sequence.py:Sequence the metadata builders read."""
    num_tokens: int
    block_table: list[int] = field(default_factory=list)
    num_cached_tokens: int = 0

    def num_blocks(self, block_size: int) -> int:  # sequence.py:62-64
        return (self.num_tokens + block_size - 1) // block_size

    def num_cached_blocks(self, block_size: int) -> int:  # sequence.py:58-60
        return self.num_cached_tokens // block_size

    def last_block_num_tokens(self, block_size: int) -> int:  # sequence.py:66-68
        return self.num_tokens - (self.num_blocks(block_size) - 1) * block_size


def _tensor(data, dtype, pin: bool) -> torch.Tensor:
    return torch.tensor(data, dtype=dtype, pin_memory=pin)


def prepare_block_tables(seqs: list[SeqState], pin: bool = False) -> torch.Tensor:
    """[num_seqs, max blocks] int32, right-padded with -1 (model_runner.py:152-158)."""
    max_len = max(len(s.block_table) for s in seqs)
    rows = [s.block_table + [-1] * (max_len - len(s.block_table)) for s in seqs]
    return _tensor(rows, torch.int32, pin)


def prepare_prefill(seqs: list[SeqState], block_size: int, pin: bool = False) -> dict:
    """Fields of set_context(True, ...) plus `positions` (model_runner.py:160-204)."""
    positions: list[int] = []
    cu_q, cu_k = [0], [0]
    max_q = max_k = 0
    slot_mapping: list[int] = []
    for s in seqs:
        seqlen = s.num_tokens
        positions.extend(range(s.num_cached_tokens, seqlen))
        seqlen_q = seqlen - s.num_cached_tokens
        cu_q.append(cu_q[-1] + seqlen_q)
        cu_k.append(cu_k[-1] + seqlen)
        max_q = max(max_q, seqlen_q)
        max_k = max(max_k, seqlen)
        if not s.block_table:  # warm-up sequences have no blocks (model_runner.py:179-180)
            continue
        nb = s.num_blocks(block_size)
        for i in range(s.num_cached_blocks(block_size), nb):
            start = s.block_table[i] * block_size
            end = start + (block_size if i != nb - 1 else s.last_block_num_tokens(block_size))
            slot_mapping.extend(range(start, end))
    block_tables = None
    if cu_k[-1] > cu_q[-1]:  # some sequence has a cached prefix
        block_tables = prepare_block_tables(seqs, pin)
    return dict(
        is_prefill=True,
        cu_seqlens_q=_tensor(cu_q, torch.int32, pin),
        cu_seqlens_k=_tensor(cu_k, torch.int32, pin),
        max_seqlen_q=max_q,
        max_seqlen_k=max_k,
        slot_mapping=_tensor(slot_mapping, torch.int32, pin),
        context_lens=None,
        block_tables=block_tables,
        positions=_tensor(positions, torch.int64, pin),
    )


def prepare_decode(seqs: list[SeqState], block_size: int, pin: bool = False) -> dict:
    """Fields of set_context(False, ...) plus `positions` (model_runner.py:206-226)."""
    positions, slot_mapping, context_lens = [], [], []
    for s in seqs:
        positions.append(s.num_tokens - 1)
        context_lens.append(s.num_tokens)  # includes the token being decoded
        slot_mapping.append(s.block_table[-1] * block_size + s.last_block_num_tokens(block_size) - 1)
    return dict(
        is_prefill=False,
        cu_seqlens_q=None,
        cu_seqlens_k=None,
        max_seqlen_q=0,
        max_seqlen_k=0,
        slot_mapping=_tensor(slot_mapping, torch.int32, pin),
        context_lens=_tensor(context_lens, torch.int32, pin),
        block_tables=prepare_block_tables(seqs, pin),
        positions=_tensor(positions, torch.int64, pin),
    )


def context_kwargs(meta: dict, device=None) -> dict:
    """The keyword arguments of set_context() from a prepare_* result, moved to `device`."""
    keys = ("is_prefill", "cu_seqlens_q", "cu_seqlens_k", "max_seqlen_q", "max_seqlen_k",
            "slot_mapping", "context_lens", "block_tables")
    out = {}
    for k in keys:
        v = meta[k]
        if device is not None and isinstance(v, torch.Tensor):
            v = v.to(device, non_blocking=True)
        out[k] = v
    return out
