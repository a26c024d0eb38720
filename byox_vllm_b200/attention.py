"""Drop-in for reference nanovllm/layers/attention.py — same names, arguments and return shapes.

  store_kvcache(key, value, k_cache, v_cache, slot_mapping)   reference attention.py:34-49
  Attention(num_heads, head_dim, scale, num_kv_heads).forward(q, k, v)   attention.py:52-97

The arithmetic is done by libb200attn.so (hand-written sm_100a CUDA behind the C-ABI of
include/b200attn.h); this file only validates tensors, hands raw device pointers + the current
torch stream to the library, and allocates outputs / split workspaces from torch's caching
allocator so that everything stays CUDA-graph capturable (model_runner.py:256-297). There is no
Triton, no flash-attn and no CPU or eager fallback: CPU tensors raise.
"""
from __future__ import annotations

import torch
from torch import nn

from . import _lib
from . import context as _own_context

# `get_context` is a module attribute on purpose: install.py rebinds it to the reference's
# nanovllm.utils.context.get_context so ModelRunner's set_context() is what forward() sees.
get_context = _own_context.get_context


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _stream(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "byox_vllm_b200 attention runs only on CUDA tensors (sm_100a); there is no CPU path")


def _kv_rows(x: torch.Tensor, head_dim: int) -> torch.Tensor:
    """[N, H, D] with unit inner stride and stride(1) == D (row stride free), else a copy."""
    if x.stride(-1) != 1 or x.stride(1) != head_dim:
        x = x.contiguous()
    return x


def store_kvcache(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor,
                  v_cache: torch.Tensor, slot_mapping: torch.Tensor) -> None:
    """k_cache.view(-1, H*D)[slot_mapping[i]] = key[i]; same for value; negative slots skipped.

    Same preconditions as the reference launcher (attention.py:41-46); bit-exact 16-bit copy.
    """
    N, num_heads, head_dim = key.shape
    D = num_heads * head_dim
    assert key.stride(-1) == 1 and value.stride(-1) == 1
    assert key.stride(1) == head_dim and value.stride(1) == head_dim
    assert k_cache.stride(1) == D and v_cache.stride(1) == D
    assert slot_mapping.numel() == N
    _require_cuda(key, value, k_cache, v_cache, slot_mapping)
    if key.dtype != torch.bfloat16 and key.dtype != torch.float16:
        raise RuntimeError(f"store_kvcache: 16-bit dtypes only, got {key.dtype}")
    assert value.dtype == key.dtype and k_cache.dtype == key.dtype and v_cache.dtype == key.dtype
    assert slot_mapping.dtype == torch.int32 and slot_mapping.is_contiguous()
    assert k_cache.is_contiguous() and v_cache.is_contiguous()
    if N == 0:
        return
    num_slots = k_cache.numel() // D
    lib = _lib.load()
    rc = lib.b200attn_store_kvcache(
        key.data_ptr(), key.stride(0), value.data_ptr(), value.stride(0), k_cache.data_ptr(),
        v_cache.data_ptr(), slot_mapping.data_ptr(), N, D, num_slots, _lib.dtype_code(key.dtype),
        _stream(key.device))
    _lib.check("b200attn_store_kvcache", rc)


_SPLIT_COUNTERS: dict[int, list[torch.Tensor]] = {}  # per device, newest (registered) last
_SPLIT_COUNTER_SLOTS = 1 << 16  # (sequence, KV head) pairs; 256 KiB per device


def ensure_split_counters(device: torch.device, need: int = 0) -> None:
    """Register the zeroed arrival counters that let split-KV decode merge inside its own launch
    (b200attn_set_split_counters). Allocated once per device, outside any graph capture (the
    reference warms every batch size up eagerly before capturing it, model_runner.py:281), kept
    alive for the life of the process; every launch leaves them zero.

    `need` = counters the coming call uses (rows x KV heads + 2). The default buffer covers every
    batch the engine issues; an eager call beyond it (>= 65535 (sequence, KV head) pairs with an
    explicit num_splits > 1) registers a larger buffer. Buffers already handed to the library stay
    allocated: graphs captured earlier keep launching on them."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    bufs = _SPLIT_COUNTERS.setdefault(idx, [])
    if bufs and bufs[-1].numel() >= need:
        return
    if torch.cuda.is_current_stream_capturing():
        raise RuntimeError(
            "byox_vllm_b200: the decode arrival counters must be registered before CUDA-graph "
            "capture; run one eager decode step (or attention.ensure_split_counters(device)) "
            "first, as ModelRunner's warm-up before capture does (model_runner.py:281)")
    slots = max(_SPLIT_COUNTER_SLOTS, 1 << max(int(need) - 1, 1).bit_length())
    with torch.cuda.device(idx):
        buf = torch.zeros(slots, dtype=torch.int32, device=device)
        torch.cuda.current_stream(device).synchronize()
        rc = _lib.load().b200attn_set_split_counters(buf.data_ptr(), buf.numel())
    _lib.check("b200attn_set_split_counters", rc)
    bufs.append(buf)


def set_decode_overlap(on: bool) -> bool:
    """Overlap mode of the decode launches (b200attn_set_decode_overlap): each launch becomes a
    programmatic dependent of the kernel before it in the stream and runs its prologue (scan of
    context_lens, first work item, block-table window) while that kernel drains. Only valid when
    context_lens / block_tables are not written by the kernel right before a decode call - true
    inside the engine, where they are set once per step (model_runner.py:206-226); install()
    switches it on. Returns the previous setting."""
    return bool(_lib.load().b200attn_set_decode_overlap(1 if on else 0))


def decode_num_splits(batch: int, num_kv_heads: int, table_width: int, block_size: int) -> int:
    """Host-only choice of the number of persistent CTAs the batch's KV tile stream is cut into
    (a function of host-visible sizes: graph-capture rule; the cut itself happens on the device
    from context_lens)."""
    lib = _lib.load()
    return lib.b200attn_decode_pick_splits(batch, num_kv_heads, table_width, block_size,
                                           lib.b200attn_sm_count())


def cascade_chunks(batch: int, num_q_heads: int, num_kv_heads: int, table_width: int,
                   block_size: int) -> int:
    """Host-only: split-KV chunks of the shared-prefix pass, 0 = cascade not armed for this shape
    (B200ATTN_CASCADE=0|1 forces it off / on)."""
    lib = _lib.load()
    return lib.b200attn_cascade_pick_chunks(batch, num_q_heads, num_kv_heads, table_width,
                                            block_size, lib.b200attn_sm_count())


def cascade_plan(context_lens: torch.Tensor, block_tables: torch.Tensor,
                 block_size: int) -> torch.Tensor:
    """Device int32[4] {shared pages P, first active row, active rows, 0} for this step's metadata
    (one tiny launch; depends only on block_tables / context_lens, so all layers share it)."""
    _require_cuda(context_lens, block_tables)
    assert context_lens.dtype == torch.int32 and context_lens.is_contiguous()
    assert block_tables.dtype == torch.int32 and block_tables.dim() == 2
    assert block_tables.stride(1) == 1 and block_tables.size(0) == context_lens.numel()
    plan = torch.empty(4, dtype=torch.int32, device=block_tables.device)
    lib = _lib.load()
    rc = lib.b200attn_cascade_plan(
        block_tables.data_ptr(), block_tables.stride(0), block_tables.size(1),
        context_lens.data_ptr(), context_lens.numel(), block_size, plan.data_ptr(),
        _stream(block_tables.device))
    _lib.check("b200attn_cascade_plan", rc)
    return plan


def paged_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                 context_lens: torch.Tensor, block_tables: torch.Tensor, scale: float,
                 *, key: torch.Tensor | None = None, value: torch.Tensor | None = None,
                 slot_mapping: torch.Tensor | None = None, num_splits: int | None = None,
                 out: torch.Tensor | None = None, plan: torch.Tensor | None = None,
                 prefix_chunks: int = 0) -> torch.Tensor:
    """Paged single-query attention (the decode branch, reference attention.py:87-96).

    q [B, Hq, D]; caches [num_blocks, block_size, Hkv, D]; context_lens [B] int32 (includes the
    current token); block_tables [B, W] int32. If key/value/slot_mapping are given, this step's
    K/V rows are scattered into the cache first (same stream), as Attention.forward does.
    `plan` (from cascade_plan) + `prefix_chunks` (from cascade_chunks, > 0) arm the shared-prefix
    cascade: same result, shared leading pages attended once per KV head on the tensor cores.
    Returns [B, Hq, D].
    """
    _require_cuda(q, k_cache, v_cache, context_lens, block_tables, key, value, slot_mapping)
    B, Hq, D = q.shape
    num_blocks, block_size, Hkv, Dc = k_cache.shape
    assert Dc == D and v_cache.shape == k_cache.shape
    dcode = _lib.dtype_code(q.dtype)  # bf16 or fp16
    assert k_cache.dtype == q.dtype and v_cache.dtype == q.dtype
    assert k_cache.is_contiguous() and v_cache.is_contiguous()
    assert context_lens.dtype == torch.int32 and context_lens.numel() == B
    assert context_lens.is_contiguous()
    assert block_tables.dtype == torch.int32 and block_tables.dim() == 2
    assert block_tables.size(0) == B and block_tables.stride(1) == 1
    if not q.is_contiguous():
        q = q.contiguous()
    if out is None:
        out = torch.empty_like(q)
    else:
        assert out.shape == q.shape and out.dtype == q.dtype and out.is_contiguous()
    if B == 0:
        return out
    W = block_tables.size(1)
    if num_splits is None:
        num_splits = decode_num_splits(B, Hkv, W, block_size)
    lib = _lib.load()
    if plan is None or prefix_chunks <= 0:
        plan, prefix_chunks = None, 0
    else:
        assert plan.dtype == torch.int32 and plan.numel() >= 4 and plan.is_cuda
    ws, ws_bytes = None, 0
    if num_splits > 1:
        ensure_split_counters(q.device, B * Hkv + 2)
    if num_splits > 1 or prefix_chunks > 0:
        ws_bytes = lib.b200attn_decode_workspace_bytes(B, Hq, D, num_splits, prefix_chunks)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=q.device)
    stream = _stream(q.device)
    if key is not None:
        assert value is not None and slot_mapping is not None
        assert key.shape == (B, Hkv, D) and value.shape == (B, Hkv, D)
        assert key.stride(-1) == 1 and value.stride(-1) == 1
        assert key.stride(1) == D and value.stride(1) == D
        assert slot_mapping.dtype == torch.int32 and slot_mapping.numel() == B
        assert slot_mapping.is_contiguous()
    if plan is not None:
        rc = lib.b200attn_store_decode_cascade(
            q.data_ptr(), _ptr(key), 0 if key is None else key.stride(0), _ptr(value),
            0 if key is None else value.stride(0), None if key is None else _ptr(slot_mapping),
            k_cache.data_ptr(), v_cache.data_ptr(), block_tables.data_ptr(),
            block_tables.stride(0), W, context_lens.data_ptr(), out.data_ptr(), _ptr(ws),
            ws_bytes, B, Hq, Hkv, D, block_size, num_blocks, float(scale), num_splits,
            plan.data_ptr(), prefix_chunks, dcode, stream)
        _lib.check("b200attn_store_decode_cascade", rc)
    elif key is not None:
        rc = lib.b200attn_store_decode(
            q.data_ptr(), key.data_ptr(), key.stride(0), value.data_ptr(), value.stride(0),
            slot_mapping.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
            block_tables.data_ptr(), block_tables.stride(0), W, context_lens.data_ptr(),
            out.data_ptr(), _ptr(ws), ws_bytes, B, Hq, Hkv, D, block_size, num_blocks,
            float(scale), num_splits, dcode, stream)
        _lib.check("b200attn_store_decode", rc)
    else:
        rc = lib.b200attn_decode(
            q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), block_tables.data_ptr(),
            block_tables.stride(0), W, context_lens.data_ptr(), out.data_ptr(), _ptr(ws),
            ws_bytes, B, Hq, Hkv, D, block_size, num_blocks, float(scale), num_splits,
            dcode, stream)
        _lib.check("b200attn_decode", rc)
    return out


def prefill_varlen(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens_q: torch.Tensor,
                   cu_seqlens_k: torch.Tensor, max_seqlen_q: int, max_seqlen_k: int, scale: float,
                   block_tables: torch.Tensor | None = None,
                   out: torch.Tensor | None = None) -> torch.Tensor:
    """Varlen causal attention, bottom-right aligned (the prefill branch, attention.py:72-86).

    Dense (block_tables None): k, v are [total_k, Hkv, D] rows of this step (row stride free).
    Paged (block_tables [S, W]): k, v are the caches [num_blocks, block_size, Hkv, D].
    Returns [total_q, Hq, D].
    """
    _require_cuda(q, k, v, cu_seqlens_q, cu_seqlens_k, block_tables)
    total_q, Hq, D = q.shape
    dcode = _lib.dtype_code(q.dtype)  # bf16 or fp16
    assert k.dtype == q.dtype and v.dtype == q.dtype
    assert cu_seqlens_q.dtype == torch.int32 and cu_seqlens_k.dtype == torch.int32
    assert cu_seqlens_q.is_contiguous() and cu_seqlens_k.is_contiguous()
    num_seqs = cu_seqlens_q.numel() - 1
    assert cu_seqlens_k.numel() == num_seqs + 1
    if not q.is_contiguous():
        q = q.contiguous()
    if out is None:
        out = torch.empty_like(q)
    else:
        assert out.shape == q.shape and out.dtype == q.dtype and out.is_contiguous()
    if total_q == 0 or num_seqs <= 0:
        return out
    lib = _lib.load()
    stream = _stream(q.device)
    if block_tables is None:
        assert k.dim() == 3 and v.shape == k.shape and k.size(2) == D
        Hkv = k.size(1)
        k, v = _kv_rows(k, D), _kv_rows(v, D)
        rc = lib.b200attn_prefill_varlen(
            q.data_ptr(), k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0), None, None, None,
            0, 0, cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr(), num_seqs, total_q, k.size(0),
            int(max_seqlen_q), int(max_seqlen_k), out.data_ptr(), Hq, Hkv, D, 0, 0, float(scale),
            dcode, stream)
    else:
        assert k.dim() == 4 and v.shape == k.shape and k.size(3) == D
        assert k.is_contiguous() and v.is_contiguous()
        num_blocks, block_size, Hkv, _ = k.shape
        assert block_tables.dtype == torch.int32 and block_tables.dim() == 2
        assert block_tables.size(0) == num_seqs and block_tables.stride(1) == 1
        rc = lib.b200attn_prefill_varlen(
            q.data_ptr(), None, 0, None, 0, k.data_ptr(), v.data_ptr(), block_tables.data_ptr(),
            block_tables.stride(0), block_tables.size(1), cu_seqlens_q.data_ptr(),
            cu_seqlens_k.data_ptr(), num_seqs, total_q, 0, int(max_seqlen_q), int(max_seqlen_k),
            out.data_ptr(), Hq, Hkv, D, block_size, num_blocks, float(scale), dcode,
            stream)
    _lib.check("b200attn_prefill_varlen", rc)
    return out


class Attention(nn.Module):
    """Same constructor, attributes and forward contract as reference attention.py:52-97.

    `k_cache` / `v_cache` start as empty tensors and are replaced by ModelRunner with per-layer
    views of the KV cache (model_runner.py:145-150). Metadata comes only from get_context().
    """

    def __init__(self, num_heads, head_dim, scale, num_kv_heads):
        super().__init__()
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.scale = scale
        self.num_kv_heads = num_kv_heads
        self.k_cache = self.v_cache = torch.tensor([])

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor):
        context = get_context()
        k_cache, v_cache = self.k_cache, self.v_cache
        have_cache = bool(k_cache.numel() and v_cache.numel())
        if context.is_prefill:
            # The scatter must complete before attention reads the cache: a later sequence of the
            # same batch may hit blocks an earlier one is writing in this very pass
            # (block_manager.py:68-70), so store and prefill are never fused.
            if have_cache:
                store_kvcache(k, v, k_cache, v_cache, context.slot_mapping)
            if context.block_tables is not None:  # prefix cache
                k, v = k_cache, v_cache
            return prefill_varlen(q, k, v, context.cu_seqlens_q, context.cu_seqlens_k,
                                  context.max_seqlen_q, context.max_seqlen_k, self.scale,
                                  block_tables=context.block_tables)
        # decode: one fused library call = scatter of this step's rows + paged attention
        plan, chunks = _step_plan(context, q.size(0), self.num_heads, self.num_kv_heads, k_cache)
        o = paged_decode(q, k_cache, v_cache, context.context_lens, context.block_tables,
                         self.scale, key=k if have_cache else None,
                         value=v if have_cache else None,
                         slot_mapping=context.slot_mapping if have_cache else None,
                         plan=plan, prefix_chunks=chunks)
        return o.unsqueeze(1)  # [B, 1, Hq, D] as flash_attn_with_kvcache returns


def _step_plan(context, batch: int, num_heads: int, num_kv_heads: int, k_cache: torch.Tensor):
    """(plan, prefix_chunks) of the shared-prefix cascade for this decode step, or (None, 0).

    The plan depends only on this step's metadata, and the reference builds a fresh Context per
    step and per captured graph (utils/context.py:23-45, model_runner.py:225,275-280), so the first
    layer computes it and parks it on the Context object; the other layers (and, under capture, the
    other layers of the same graph) reuse it: one plan launch per step, not per layer."""
    bt = context.block_tables
    if bt is None or k_cache.dim() != 4:
        return None, 0
    block_size = k_cache.size(1)
    chunks = cascade_chunks(batch, num_heads, num_kv_heads, bt.size(1), block_size)
    if chunks <= 0:
        return None, 0
    # ModelRunner warms a graph batch size up and then captures it under the SAME Context
    # (model_runner.py:281-284): a plan launched eagerly must not be reused inside the capture,
    # or the graph would replay without its plan kernel.
    key = (bt.data_ptr(), context.context_lens.data_ptr(), batch,
           torch.cuda.is_current_stream_capturing())
    cached = getattr(context, "_b200_cascade_plan", None)
    if cached is not None and cached[0] == key:
        return cached[1], chunks
    plan = cascade_plan(context.context_lens, bt, block_size)
    context._b200_cascade_plan = (key, plan)
    return plan, chunks
