"""Seeded synthetic workloads for the attention hot path (tests, golden makers, bench, smoke).

Inputs follow SURVEY.md §8(d): bf16, q/k/v ~ N(0,1), KV cache pre-filled N(0,1), block tables a
random permutation of physical block ids (scattered pages), scale = head_dim ** -0.5. Metadata is
formed by byox_vllm_b200.metadata exactly as ModelRunner would form it.
Small cases are generated on the CPU (bit-identical on every machine of this image: fixtures
store only outputs + an input checksum); large bench cases are generated on the device.
"""
from __future__ import annotations

import hashlib
import random
from dataclasses import dataclass

import torch

from tools import metadata_ref as md


@dataclass
class DecodeCase:
    q: torch.Tensor            # [B, Hq, D]
    k: torch.Tensor            # [B, Hkv, D] this step's key rows (contiguous)
    v: torch.Tensor            # [B, Hkv, D] this step's value rows (strided view, like qkv.split)
    k_cache: torch.Tensor      # [num_blocks, block_size, Hkv, D]
    v_cache: torch.Tensor
    meta: dict                 # prepare_decode() result (CPU tensors)
    seqs: list
    scale: float


@dataclass
class PrefillCase:
    q: torch.Tensor            # [total_q, Hq, D]
    k: torch.Tensor            # [total_q, Hkv, D]
    v: torch.Tensor            # strided view
    k_cache: torch.Tensor
    v_cache: torch.Tensor
    meta: dict                 # prepare_prefill() result
    seqs: list
    scale: float


def _randn(shape, gen, device, dtype=torch.bfloat16):
    return torch.randn(shape, generator=gen, device=device, dtype=torch.float32).to(dtype)


def _qkv(n, Hq, Hkv, D, gen, device):
    """q, k contiguous; v a strided view of a packed [n, (Hq+2Hkv)*D] buffer, as qkv.split gives
    (reference models/qwen3.py:76-79: v keeps the row stride of the fused QKV output)."""
    q = _randn((n, Hq, D), gen, device)
    k = _randn((n, Hkv, D), gen, device)
    packed = _randn((n, (Hq + 2 * Hkv) * D), gen, device)
    v = packed[:, (Hq + Hkv) * D:].view(n, Hkv, D)
    return q, k, v


def _alloc_blocks(n_needed: int, num_blocks: int, rng: random.Random) -> list[int]:
    ids = list(range(num_blocks))
    rng.shuffle(ids)
    return ids[:n_needed]


def decode_case(batch: int, ctx, Hq: int, Hkv: int, D: int = 128, block_size: int = 256,
                seed: int = 0, device="cpu", spare_blocks: int = 3,
                fill_cache: bool = True) -> DecodeCase:
    """`ctx`: int (every row) or list of per-row context lengths (each includes the new token)."""
    lens = [ctx] * batch if isinstance(ctx, int) else list(ctx)
    assert len(lens) == batch and all(c >= 1 for c in lens)
    rng = random.Random(seed)
    gen = torch.Generator(device=device).manual_seed(seed)
    need = [(c + block_size - 1) // block_size for c in lens]
    num_blocks = sum(need) + spare_blocks
    ids = _alloc_blocks(sum(need), num_blocks, rng)
    seqs, at = [], 0
    for c, n in zip(lens, need):
        seqs.append(md.SeqState(num_tokens=c, block_table=ids[at:at + n]))
        at += n
    shape = (num_blocks, block_size, Hkv, D)
    if fill_cache:
        k_cache, v_cache = _randn(shape, gen, device), _randn(shape, gen, device)
    else:
        k_cache = torch.zeros(shape, dtype=torch.bfloat16, device=device)
        v_cache = torch.zeros(shape, dtype=torch.bfloat16, device=device)
    q, k, v = _qkv(batch, Hq, Hkv, D, gen, device)
    return DecodeCase(q, k, v, k_cache, v_cache, md.prepare_decode(seqs, block_size), seqs,
                      float(D) ** -0.5)


def shared_prefix_decode_case(ctx, shared_pages, Hq: int, Hkv: int, D: int = 128,
                              block_size: int = 256, seed: int = 0, device="cpu",
                              spare_blocks: int = 3) -> DecodeCase:
    """Decode batch whose rows start with the same physical pages, as BlockManager hands out for
    prefix-cache hits (reference engine/block_manager.py:41-71). `shared_pages`: int (every row
    shares that many leading pages with row 0) or a per-row list (rows with fewer shared pages
    diverge earlier; 0 = unrelated row). A row shares at most the pages it has."""
    lens = list(ctx)
    batch = len(lens)
    sp = [shared_pages] * batch if isinstance(shared_pages, int) else list(shared_pages)
    assert len(sp) == batch and all(c >= 1 for c in lens)
    rng = random.Random(seed)
    gen = torch.Generator(device=device).manual_seed(seed)
    need = [(c + block_size - 1) // block_size for c in lens]
    sp = [min(s, n) for s, n in zip(sp, need)]
    n_shared = max(sp) if sp else 0
    own = [n - s for n, s in zip(need, sp)]
    num_blocks = n_shared + sum(own) + spare_blocks
    ids = _alloc_blocks(n_shared + sum(own), num_blocks, rng)
    shared_ids, at = ids[:n_shared], n_shared
    seqs = []
    for c, s, o in zip(lens, sp, own):
        seqs.append(md.SeqState(num_tokens=c, block_table=shared_ids[:s] + ids[at:at + o]))
        at += o
    shape = (num_blocks, block_size, Hkv, D)
    k_cache, v_cache = _randn(shape, gen, device), _randn(shape, gen, device)
    q, k, v = _qkv(batch, Hq, Hkv, D, gen, device)
    return DecodeCase(q, k, v, k_cache, v_cache, md.prepare_decode(seqs, block_size), seqs,
                      float(D) ** -0.5)


def prefill_case(lens, Hq: int, Hkv: int, cached=None, D: int = 128, block_size: int = 256,
                 seed: int = 0, device="cpu", spare_blocks: int = 3) -> PrefillCase:
    """`lens`: full lengths; `cached`: per-sequence cached prefix tokens (multiples of block_size,
    < len) or None. The cached prefix rows already sit in the cache (random N(0,1))."""
    lens = list(lens)
    cached = [0] * len(lens) if cached is None else list(cached)
    assert all(c % block_size == 0 and c < n for c, n in zip(cached, lens))
    rng = random.Random(seed)
    gen = torch.Generator(device=device).manual_seed(seed)
    need = [(n + block_size - 1) // block_size for n in lens]
    num_blocks = sum(need) + spare_blocks
    ids = _alloc_blocks(sum(need), num_blocks, rng)
    seqs, at = [], 0
    for n, nb, c in zip(lens, need, cached):
        seqs.append(md.SeqState(num_tokens=n, block_table=ids[at:at + nb], num_cached_tokens=c))
        at += nb
    shape = (num_blocks, block_size, Hkv, D)
    k_cache, v_cache = _randn(shape, gen, device), _randn(shape, gen, device)
    total_q = sum(n - c for n, c in zip(lens, cached))
    q, k, v = _qkv(total_q, Hq, Hkv, D, gen, device)
    return PrefillCase(q, k, v, k_cache, v_cache, md.prepare_prefill(seqs, block_size), seqs,
                       float(D) ** -0.5)


def bench_shape_lengths(num_seqs: int = 256, lo: int = 100, hi: int = 1024, seed: int = 0):
    """Prompt / output lengths of the reference bench (bench.py:17-36: seed(0), randint(100, max)
    for each prompt then for each max_tokens)."""
    rng = random.Random(seed)
    prompts = []
    for _ in range(num_seqs):  # same RNG consumption order as the reference: length, then ids
        n = rng.randint(lo, hi)
        for _ in range(n):
            rng.randint(0, 10000)
        prompts.append(n)
    outputs = [rng.randint(lo, hi) for _ in range(num_seqs)]
    return prompts, outputs


def checksum(*tensors: torch.Tensor) -> str:
    """sha256 over the raw bytes of the given tensors (guards fixtures against generator drift)."""
    h = hashlib.sha256()
    for t in tensors:
        t = t.detach().cpu().contiguous()
        if t.dtype == torch.bfloat16:
            t = t.view(torch.int16)
        h.update(t.numpy().tobytes())
    return h.hexdigest()[:16]
