"""the steps between qkv_proj and o_proj of reference
nanovllm/models/qwen3.py:75-84 as fused library calls.

    q, k, v = qkv.split(...); q = q_norm(q); k = k_norm(k)      qwen3.py:75-82
    q, k = rotary_emb(positions, q, k)                          qwen3.py:83
    o = attn(q, k, v)                                           qwen3.py:84 (store + attention)

  qk_norm_rope_store(...)   one pass over the QKV rows: q/k RMSNorm + RoPE + KV-cache scatter
  qkv_decode(...)           the whole decode step of a layer in ONE launch
  qkv_attention(attn, qkv, positions, ...)   dispatch on the Context exactly as Attention.forward
  qwen3_attention_forward   drop-in for Qwen3Attention.forward (install.py binds it)

Same rules as attention.py: raw device pointers + the current torch stream go to libb200attn.so,
outputs / workspaces come from torch's caching allocator, no CPU or eager fallback.
"""
from __future__ import annotations

import torch

from . import _lib
from . import attention as _attn
from .attention import _ptr, _require_cuda, _stream


def _check_norm_rope(qkv, positions, q_weight, k_weight, cos_sin_cache, Hq, Hkv, D):
    _require_cuda(qkv, positions, q_weight, k_weight, cos_sin_cache)
    if qkv.dtype != torch.bfloat16:
        raise RuntimeError(f"qkv: bf16 only, got {qkv.dtype}")
    assert qkv.dim() == 2 and qkv.stride(1) == 1 and qkv.size(1) == (Hq + 2 * Hkv) * D
    assert positions.dtype == torch.int64 and positions.is_contiguous()
    assert positions.numel() == qkv.size(0)
    assert (q_weight is None) == (k_weight is None)
    if q_weight is not None:
        assert q_weight.dtype == torch.bfloat16 and k_weight.dtype == torch.bfloat16
        assert q_weight.numel() == D and k_weight.numel() == D
        assert q_weight.is_contiguous() and k_weight.is_contiguous()
    # RotaryEmbedding.cos_sin_cache: [max_position, 1, D] fp32 (rotary_embedding.py:27-35)
    assert cos_sin_cache.dtype == torch.float32 and cos_sin_cache.is_contiguous()
    assert cos_sin_cache.numel() == cos_sin_cache.size(0) * D


def qk_norm_rope_store(qkv: torch.Tensor, positions: torch.Tensor, q_weight, k_weight,
                       eps: float, cos_sin_cache: torch.Tensor, num_q_heads: int,
                       num_kv_heads: int, head_dim: int = 128, *,
                       k_cache: torch.Tensor | None = None, v_cache: torch.Tensor | None = None,
                       slot_mapping: torch.Tensor | None = None, write_k: bool = True):
    """qkv [N, (Hq + 2 Hkv) * D] (raw qkv_proj output) -> (q, k, v).

    q is a new contiguous [N, Hq, D] tensor (normalised + rotated); k and v are the [N, Hkv, D]
    strided views of qkv the reference would pass on (qwen3.py:77-79), k rewritten in place with
    the normalised + rotated rows when write_k. With k_cache / v_cache / slot_mapping the K rows
    (rotated) and V rows are also scattered into the paged cache, as store_kvcache would.
    """
    Hq, Hkv, D = num_q_heads, num_kv_heads, head_dim
    _check_norm_rope(qkv, positions, q_weight, k_weight, cos_sin_cache, Hq, Hkv, D)
    N = qkv.size(0)
    q = torch.empty(N, Hq, D, dtype=qkv.dtype, device=qkv.device)
    k = qkv[:, Hq * D:(Hq + Hkv) * D].view(N, Hkv, D)
    v = qkv[:, (Hq + Hkv) * D:].view(N, Hkv, D)
    num_slots = 0
    if slot_mapping is not None:
        _require_cuda(k_cache, v_cache, slot_mapping)
        assert k_cache.is_contiguous() and v_cache.is_contiguous()
        assert k_cache.dtype == qkv.dtype and v_cache.dtype == qkv.dtype
        assert slot_mapping.dtype == torch.int32 and slot_mapping.is_contiguous()
        assert slot_mapping.numel() == N
        num_slots = k_cache.numel() // (Hkv * D)
    if N == 0:
        return q, k, v
    lib = _lib.load()
    rc = lib.b200attn_qk_norm_rope_store(
        qkv.data_ptr(), qkv.stride(0), positions.data_ptr(), _ptr(q_weight), _ptr(k_weight),
        float(eps), cos_sin_cache.data_ptr(), cos_sin_cache.size(0), q.data_ptr(),
        k.data_ptr() if write_k else None, qkv.stride(0),
        None if slot_mapping is None else k_cache.data_ptr(),
        None if slot_mapping is None else v_cache.data_ptr(), _ptr(slot_mapping), N, Hq, Hkv, D,
        num_slots, _lib.DTYPE_BF16, _stream(qkv.device))
    _lib.check("b200attn_qk_norm_rope_store", rc)
    return q, k, v


def qkv_decode(qkv: torch.Tensor, positions: torch.Tensor, q_weight, k_weight, eps: float,
               cos_sin_cache: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
               context_lens: torch.Tensor, block_tables: torch.Tensor,
               slot_mapping: torch.Tensor, scale: float, num_q_heads: int, num_kv_heads: int,
               *, num_splits: int | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
    """One launch per layer per decode step: q/k RMSNorm + RoPE, scatter of the new K/V rows and
    paged attention over the updated cache, reading q/k/v straight from qkv [B, (Hq+2Hkv)*D].
    Returns [B, Hq, D]."""
    Hq, Hkv = num_q_heads, num_kv_heads
    num_blocks, block_size, Hkv_c, D = k_cache.shape
    assert Hkv_c == Hkv and v_cache.shape == k_cache.shape
    _check_norm_rope(qkv, positions, q_weight, k_weight, cos_sin_cache, Hq, Hkv, D)
    _require_cuda(k_cache, v_cache, context_lens, block_tables, slot_mapping)
    B = qkv.size(0)
    assert k_cache.dtype == qkv.dtype and v_cache.dtype == qkv.dtype
    assert k_cache.is_contiguous() and v_cache.is_contiguous()
    assert context_lens.dtype == torch.int32 and context_lens.numel() == B
    assert context_lens.is_contiguous()
    assert block_tables.dtype == torch.int32 and block_tables.dim() == 2
    assert block_tables.size(0) == B and block_tables.stride(1) == 1
    assert slot_mapping.dtype == torch.int32 and slot_mapping.numel() == B
    assert slot_mapping.is_contiguous()
    if out is None:
        out = torch.empty(B, Hq, D, dtype=qkv.dtype, device=qkv.device)
    else:
        assert out.shape == (B, Hq, D) and out.dtype == qkv.dtype and out.is_contiguous()
    if B == 0:
        return out
    W = block_tables.size(1)
    if num_splits is None:
        num_splits = _attn.decode_num_splits(B, Hkv, W, block_size)
    lib = _lib.load()
    ws, ws_bytes = None, 0
    if num_splits > 1:
        _attn.ensure_split_counters(qkv.device, B * Hkv + 2)
        ws_bytes = lib.b200attn_decode_workspace_bytes(B, Hq, D, num_splits, 0)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=qkv.device)
    rc = lib.b200attn_qkv_decode(
        qkv.data_ptr(), qkv.stride(0), positions.data_ptr(), _ptr(q_weight), _ptr(k_weight),
        float(eps), cos_sin_cache.data_ptr(), cos_sin_cache.size(0), slot_mapping.data_ptr(),
        k_cache.data_ptr(), v_cache.data_ptr(), block_tables.data_ptr(), block_tables.stride(0),
        W, context_lens.data_ptr(), out.data_ptr(), _ptr(ws), ws_bytes, B, Hq, Hkv, D,
        block_size, num_blocks, float(scale), num_splits, _lib.DTYPE_BF16, _stream(qkv.device))
    _lib.check("b200attn_qkv_decode", rc)
    return out


def qkv_attention(attn, qkv: torch.Tensor, positions: torch.Tensor, q_weight, k_weight,
                  eps: float, cos_sin_cache: torch.Tensor) -> torch.Tensor:
    """qwen3.py:75-84 for an `Attention` module `attn` (its caches, head counts and scale), with
    the same dispatch on the Context as Attention.forward (attention.py:67-97). Returns
    [N, Hq, D]."""
    context = _attn.get_context()
    Hq, Hkv, D = attn.num_heads, attn.num_kv_heads, attn.head_dim
    k_cache, v_cache = attn.k_cache, attn.v_cache
    have_cache = bool(k_cache.numel() and v_cache.numel())
    if qkv.stride(1) != 1:
        qkv = qkv.contiguous()
    cache_args = dict(k_cache=k_cache, v_cache=v_cache,
                      slot_mapping=context.slot_mapping) if have_cache else {}
    if context.is_prefill:
        dense = context.block_tables is None
        q, k, v = qk_norm_rope_store(qkv, positions, q_weight, k_weight, eps, cos_sin_cache, Hq,
                                     Hkv, D, write_k=dense, **cache_args)
        if not dense:  # prefix cache: keys come from the cache the call above just filled
            k, v = k_cache, v_cache
        return _attn.prefill_varlen(q, k, v, context.cu_seqlens_q, context.cu_seqlens_k,
                                    context.max_seqlen_q, context.max_seqlen_k, attn.scale,
                                    block_tables=context.block_tables)
    if not have_cache:
        # Attention.forward would be called with the empty placeholder caches of __init__
        # (attention.py:62): a decode step has nothing to attend to before ModelRunner injected
        # the cache views (model_runner.py:145-150)
        raise RuntimeError("byox_vllm_b200: decode step before the KV cache was injected into "
                           "the Attention module (k_cache / v_cache are still empty)")
    plan, chunks = _attn._step_plan(context, qkv.size(0), Hq, Hkv, k_cache)
    if plan is not None:
        # shared-prefix cascade: its prefix pass reads q through a TMA map, so q is materialised
        q, _, _ = qk_norm_rope_store(qkv, positions, q_weight, k_weight, eps, cos_sin_cache, Hq,
                                     Hkv, D, write_k=False, **cache_args)
        return _attn.paged_decode(q, k_cache, v_cache, context.context_lens,
                                  context.block_tables, attn.scale, plan=plan,
                                  prefix_chunks=chunks)
    return qkv_decode(qkv, positions, q_weight, k_weight, eps, cos_sin_cache, k_cache, v_cache,
                      context.context_lens, context.block_tables, context.slot_mapping,
                      attn.scale, Hq, Hkv)


def qwen3_attention_forward(self, positions: torch.Tensor,
                            hidden_states: torch.Tensor) -> torch.Tensor:
    """Drop-in for Qwen3Attention.forward (models/qwen3.py:70-86): qkv_proj and o_proj stay the
    reference's own layers (and o_proj's all_reduce stays NCCL); everything between them is the
    fused path."""
    if hidden_states.dtype != torch.bfloat16:
        # The fused norm / RoPE kernels restate the reference's bf16 rounding points and exist for
        # bf16 only; an fp16 checkpoint (model_runner.py:30 takes the dtype from the config) keeps
        # the reference's own q_norm / k_norm / rotary_emb and enters at Attention.forward, whose
        # kernels (store, decode, cascade, prefill) do take fp16.
        from . import install as _install

        if _install._ORIG_QWEN3_FORWARD is None:
            raise RuntimeError("byox_vllm_b200: fused QKV path is bf16 only and the reference's "
                               "Qwen3Attention.forward is not available to fall back to")
        return _install._ORIG_QWEN3_FORWARD(self, positions, hidden_states)
    qkv = self.qkv_proj(hidden_states)
    has_norm = not self.qkv_bias
    o = qkv_attention(self.attn, qkv, positions,
                      self.q_norm.weight if has_norm else None,
                      self.k_norm.weight if has_norm else None,
                      self.q_norm.eps if has_norm else 0.0, self.rotary_emb.cos_sin_cache)
    return self.o_proj(o.flatten(1, -1))
