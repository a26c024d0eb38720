"""CPU oracle for nanovllm's attention hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may
import this module, and only as the checker or the timed CPU baseline. The product path
(byox_vllm_b200/) never imports it and has no CPU fallback.

What it restates (paths relative to the reference tree /root/reference):
  store_kvcache_ref      nanovllm/layers/attention.py:10-31 (Triton kernel) + :34-49 (launcher)
  paged_decode_ref       nanovllm/layers/attention.py:87-96 -> flash_attn_with_kvcache
  prefill_varlen_ref     nanovllm/layers/attention.py:72-86 -> flash_attn_varlen_func
  attention_forward_ref  nanovllm/layers/attention.py:67-97 (dispatch on the Context)
  qk_norm_rope_ref       nanovllm/models/qwen3.py:75-83 (split, q_norm/k_norm, rotary_emb) with
                         layers/layernorm.py:16-26 and layers/rotary_embedding.py:7-16,27-48
  qkv_attention_ref      qwen3.py:75-84: the steps above followed by attention_forward_ref

The attention arithmetic itself lives in a third-party dependency that is NOT in the reference
tree: flash-attn (unpinned in the reference's pyproject.toml:18; this image carries 2.8.3). Its
published semantics are restated here in fp32 torch:
  * GQA: query head h reads KV head h // (Hq / Hkv)       (flash_attn_interface.py:1389-1392)
  * causal mask aligned to the bottom-right corner; rows with no visible key give 0
                                                          (flash_attn_interface.py:1394-1404)
  * paged KV cache [num_blocks, block_size, Hkv, D] addressed by block_table [B, W] int32 and
    cache_seqlens [B] int32                               (flash_attn_interface.py:1528-1547)

Parity pins (see DESIGN.md "Oracle"):
  * scatter: bit-exact against the reference's own Triton kernel executed on CPU under
    TRITON_INTERPRET=1 — fixtures tests/golden/scatter_*.npz made by tools/make_golden_scatter.py
  * decode / prefill: against flash-attn 2.8.3 executed on a B200 through the reference's exact
    call sites — fixtures tests/golden/fa_*.npz made by tools/make_golden_flashattn.py
    (tolerance 2e-2 abs + 1e-2 rel, the bound north_star states for bf16).
  * q/k norm + RoPE: bit-exact against the reference's own RMSNorm / RotaryEmbedding modules run
    on CPU (eager: TORCHDYNAMO_DISABLE=1) — fixtures tests/golden/normrope_*.npz made by
    tools/make_golden_normrope.py.
"""
from __future__ import annotations

import math

import torch


def store_kvcache_ref(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor,
                      v_cache: torch.Tensor, slot_mapping: torch.Tensor) -> None:
    """In place: cache.view(-1, H*D)[slot[i]] = row i, for slot[i] >= 0 (reference skips -1,
    attention.py:23-24; other negative slots never occur, model_runner.py:181-187,215).
    Later tokens win on duplicate slots, as a sequential loop would."""
    N, H, D = key.shape
    row = H * D
    assert slot_mapping.numel() == N
    kc = k_cache.view(-1, row)
    vc = v_cache.view(-1, row)
    slots = slot_mapping.to(torch.int64)
    keep = slots >= 0
    idx = slots[keep]
    kc[idx] = key.reshape(N, row)[keep]
    vc[idx] = value.reshape(N, row)[keep]


def _gather_kv(cache: torch.Tensor, table_row: torch.Tensor, ctx: int) -> torch.Tensor:
    """[ctx, Hkv, D] fp32 rows of one sequence, pages taken in table order."""
    block_size = cache.shape[1]
    n_pages = (ctx + block_size - 1) // block_size
    pages = cache[table_row[:n_pages].to(torch.int64)]  # [n_pages, bs, Hkv, D]
    return pages.reshape(-1, cache.shape[2], cache.shape[3])[:ctx].float()


def _attend(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, scale: float,
            causal_offset: int | None) -> torch.Tensor:
    """q [Lq, Hq, D] fp32, k/v [Lk, Hkv, D] fp32 -> [Lq, Hq, D] fp32.
    causal_offset = Lk - Lq for bottom-right causal masking, None for no mask."""
    Lq, Hq, D = q.shape
    Lk, Hkv, _ = k.shape
    G = Hq // Hkv
    if Lk == 0:
        return torch.zeros_like(q)
    kk = k.repeat_interleave(G, dim=1)  # head h <- kv head h // G
    vv = v.repeat_interleave(G, dim=1)
    s = torch.einsum("qhd,khd->hqk", q, kk) * scale
    if causal_offset is not None:
        i = torch.arange(Lq).unsqueeze(1)
        j = torch.arange(Lk).unsqueeze(0)
        masked = j > i + causal_offset
        s = s.masked_fill(masked.unsqueeze(0), float("-inf"))
        dead = masked.all(dim=1)  # rows that see no key -> 0 output
    else:
        dead = torch.zeros(Lq, dtype=torch.bool)
    m = s.max(dim=-1, keepdim=True).values
    m = torch.where(torch.isinf(m), torch.zeros_like(m), m)
    p = torch.exp(s - m)
    denom = p.sum(dim=-1, keepdim=True)
    p = torch.where(denom > 0, p / denom, torch.zeros_like(p))
    o = torch.einsum("hqk,khd->qhd", p, vv)
    o[dead] = 0
    return o


def paged_decode_ref(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                     context_lens: torch.Tensor, block_tables: torch.Tensor,
                     scale: float | None = None) -> torch.Tensor:
    """q [B, Hq, D] -> fp32 [B, Hq, D]. Rows with context_lens <= 0 give zeros and touch no
    table entry (CUDA-graph padding rows, model_runner.py:248-251)."""
    B, Hq, D = q.shape
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    out = torch.zeros(B, Hq, D, dtype=torch.float32)
    lens = context_lens.tolist()
    for b in range(B):
        ctx = int(lens[b])
        if ctx <= 0:
            continue
        k = _gather_kv(k_cache, block_tables[b], ctx)
        v = _gather_kv(v_cache, block_tables[b], ctx)
        out[b] = _attend(q[b : b + 1].float(), k, v, scale, None)[0]
    return out


def shared_prefix_pages_ref(context_lens: torch.Tensor, block_tables: torch.Tensor,
                            block_size: int) -> tuple[int, int, int]:
    """(P, b0, active): checker for the cascade plan kernel (byox_vllm_b200/csrc/cascade.cu).
    P = leading table columns whose block id is equal in every row with context_lens > 0 and whose
    page ends before every such row's newest token (index ctx-1); 0 with fewer than two active
    rows. The sharing itself is what reference engine/block_manager.py:41-71 produces on prefix
    hits; the attention result does not depend on P (paged_decode_ref is the oracle for that)."""
    lens = [int(x) for x in context_lens.tolist()]
    rows = [b for b, c in enumerate(lens) if c > 0]
    if len(rows) < 2:
        return 0, 0, len(rows)
    pmax = min(min((lens[b] - 1) // block_size for b in rows), block_tables.shape[1])
    t0 = block_tables[rows[0]].tolist()
    P = 0
    while P < pmax and all(int(block_tables[b, P]) == t0[P] for b in rows):
        P += 1
    return P, rows[0], len(rows)


def prefill_varlen_ref(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                       cu_seqlens_q: torch.Tensor, cu_seqlens_k: torch.Tensor,
                       scale: float | None = None,
                       block_tables: torch.Tensor | None = None) -> torch.Tensor:
    """q [total_q, Hq, D] -> fp32 [total_q, Hq, D]. Dense: k, v [total_k, Hkv, D]; paged: k, v
    are the caches and rows come through block_tables (prefix-cache case, attention.py:73-74)."""
    total_q, Hq, D = q.shape
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    out = torch.zeros(total_q, Hq, D, dtype=torch.float32)
    cq = cu_seqlens_q.tolist()
    ck = cu_seqlens_k.tolist()
    for s in range(len(cq) - 1):
        q0, q1 = cq[s], cq[s + 1]
        k0, k1 = ck[s], ck[s + 1]
        Lq, Lk = q1 - q0, k1 - k0
        if Lq == 0:
            continue
        if block_tables is None:
            ks = k[k0:k1].float()
            vs = v[k0:k1].float()
        else:
            ks = _gather_kv(k, block_tables[s], Lk)
            vs = _gather_kv(v, block_tables[s], Lk)
        out[q0:q1] = _attend(q[q0:q1].float(), ks, vs, scale, Lk - Lq)
    return out


def attention_forward_ref(q, k, v, k_cache, v_cache, context, scale):
    """Attention.forward (attention.py:67-97) on CPU tensors: scatter (if the cache exists), then
    prefill or decode. Mutates the caches like the reference; returns fp32 in the reference's
    output shape (prefill [total_q, Hq, D], decode [B, 1, Hq, D])."""
    if k_cache.numel() and v_cache.numel():
        store_kvcache_ref(k, v, k_cache, v_cache, context.slot_mapping)
    if context.is_prefill:
        if context.block_tables is not None:
            k, v = k_cache, v_cache
        return prefill_varlen_ref(q, k, v, context.cu_seqlens_q, context.cu_seqlens_k, scale,
                                  context.block_tables)
    o = paged_decode_ref(q, k_cache, v_cache, context.context_lens, context.block_tables, scale)
    return o.unsqueeze(1)


def within_tolerance(got: torch.Tensor, want: torch.Tensor, atol: float = 2e-2,
                     rtol: float = 1e-2) -> tuple[bool, float]:
    """north_star bound: |got - want| <= atol + rtol * |want| elementwise. Returns (ok, max
    violation ratio) where ratio = max(|diff| / (atol + rtol |want|))."""
    got = got.float().cpu()
    want = want.float().cpu()
    if got.numel() == 0:
        return True, 0.0
    ratio = ((got - want).abs() / (atol + rtol * want.abs())).max().item()
    ok = bool(torch.isfinite(got).all()) and ratio <= 1.0
    return ok, ratio


def rope_cos_sin_cache(head_dim: int, max_position: int, base: float) -> torch.Tensor:
    """[max_position, head_dim] fp32 = cos(head_dim/2) | sin(head_dim/2), the table
    RotaryEmbedding.__init__ builds (rotary_embedding.py:27-35, without its unsqueeze)."""
    inv_freq = 1.0 / (base ** (torch.arange(0, head_dim, 2, dtype=torch.float) / head_dim))
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


def _rms_norm_ref(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    """layers/layernorm.py:16-26 (rms_forward), eager rounding points."""
    orig = x.dtype
    assert weight.dtype == orig, "the reference builds RMSNorm weights in the model dtype"
    x = x.float()
    var = x.pow(2).mean(dim=-1, keepdim=True)
    x = x * torch.rsqrt(var + eps)
    return x.to(orig) * weight


def _rope_ref(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """layers/rotary_embedding.py:7-16 (apply_rotary_emb)."""
    x1, x2 = torch.chunk(x.float(), 2, dim=-1)
    y1 = x1 * cos - x2 * sin
    y2 = x2 * cos + x1 * sin
    return torch.cat((y1, y2), dim=-1).to(x.dtype)


def qk_norm_rope_ref(qkv: torch.Tensor, positions: torch.Tensor, q_weight, k_weight, eps: float,
                     cos_sin_cache: torch.Tensor, num_q_heads: int, num_kv_heads: int,
                     head_dim: int = 128):
    """qwen3.py:75-83: split qkv rows into heads, per-head RMSNorm of q and k (skipped when the
    weights are None: qkv_bias models), neox RoPE at `positions`. Returns (q, k, v) with v a
    strided view of qkv as in the reference."""
    q_size, kv_size = num_q_heads * head_dim, num_kv_heads * head_dim
    q, k, v = qkv.split([q_size, kv_size, kv_size], dim=-1)
    q = q.reshape(-1, num_q_heads, head_dim)
    k = k.reshape(-1, num_kv_heads, head_dim)
    v = v.view(-1, num_kv_heads, head_dim)
    if q_weight is not None:
        q = _rms_norm_ref(q, q_weight, eps)
        k = _rms_norm_ref(k, k_weight, eps)
    cos_sin = cos_sin_cache[positions.long()].unsqueeze(1)
    cos, sin = cos_sin.chunk(2, dim=-1)
    return _rope_ref(q, cos, sin), _rope_ref(k, cos, sin), v


def qkv_attention_ref(qkv, positions, q_weight, k_weight, eps, cos_sin_cache, k_cache, v_cache,
                      context, scale, num_q_heads, num_kv_heads):
    """qwen3.py:75-84 up to (not including) o_proj; caches updated in place."""
    q, k, v = qk_norm_rope_ref(qkv, positions, q_weight, k_weight, eps, cos_sin_cache,
                               num_q_heads, num_kv_heads)
    return attention_forward_ref(q, k, v, k_cache, v_cache, context, scale)
