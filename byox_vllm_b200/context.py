"""Per-step attention metadata: the `get_context()` contract of the reference, field for field.

Mirror of reference nanovllm/utils/context.py:6-50 (Context / get_context / set_context /
reset_context). When the hot path is installed under nanovllm (byox_vllm_b200/install.py) the
reference's own module is used instead, so ModelRunner (model_runner.py:201-203,225,275-280) and
ParallelLMHead (embed_head.py:55-58) keep talking to the object they already know; this module is
what the Attention op reads when it is used standalone (tests, bench, smoke).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch


@dataclass
class Context:
    is_prefill: bool = False
    cu_seqlens_q: torch.Tensor | None = None
    cu_seqlens_k: torch.Tensor | None = None
    max_seqlen_q: int = 0
    max_seqlen_k: int = 0
    slot_mapping: torch.Tensor | None = None
    context_lens: torch.Tensor | None = None
    block_tables: torch.Tensor | None = None


_CONTEXT = Context()


def get_context() -> Context:
    return _CONTEXT


def set_context(is_prefill, cu_seqlens_q=None, cu_seqlens_k=None, max_seqlen_q=0, max_seqlen_k=0,
                slot_mapping=None, context_lens=None, block_tables=None) -> None:
    global _CONTEXT
    _CONTEXT = Context(is_prefill, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                       slot_mapping, context_lens, block_tables)


def reset_context() -> None:
    global _CONTEXT
    _CONTEXT = Context()
