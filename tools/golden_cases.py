"""Case tables shared by the golden-vector makers (tools/make_golden_*.py) and the tests.

Every case is regenerated from its seed by tools/synth.py on the CPU; fixtures under
tests/golden/ store the reference's OUTPUT for it plus a checksum of the regenerated inputs.
"""
from __future__ import annotations

import random

import torch

from tools import synth

# ---- scatter: (name, n_tokens, Hkv, num_blocks, block_size, hole_every, seed) -------------------
SCATTER_CASES = [
    ("n1_h8", 1, 8, 4, 16, 0, 1),
    ("n7_h8_holes", 7, 8, 4, 16, 3, 2),
    ("n33_h2_holes", 33, 2, 6, 16, 4, 3),
    ("n64_h1", 64, 1, 8, 16, 0, 4),
    ("n40_h4_holes", 40, 4, 5, 16, 5, 5),
]


def scatter_inputs(n, Hkv, num_blocks, block_size, hole_every, seed, D=128, Hq=None):
    """key contiguous, value strided (row stride (Hq+2Hkv)*D), random distinct slots with -1 holes,
    cache pre-filled with a recognisable pattern so untouched rows are checked too."""
    Hq = Hq or 2 * Hkv
    gen = torch.Generator().manual_seed(seed)
    rng = random.Random(seed)
    q, k, v = synth._qkv(n, Hq, Hkv, D, gen, "cpu")
    slots = list(range(num_blocks * block_size))
    rng.shuffle(slots)
    slots = slots[:n]
    if hole_every:
        slots = [-1 if i % hole_every == 1 else s for i, s in enumerate(slots)]
    slot_mapping = torch.tensor(slots, dtype=torch.int32)
    shape = (num_blocks, block_size, Hkv, D)
    k_cache = synth._randn(shape, gen, "cpu")
    v_cache = synth._randn(shape, gen, "cpu")
    return k, v, k_cache, v_cache, slot_mapping


# ---- decode: (name, ctx list, Hq, Hkv, seed) -----------------------------------------------------
DECODE_CASES = [
    ("c1_b4_ctx512_16_8", [512] * 4, 16, 8, 11),            # BASELINE config 1
    ("ragged_b5_16_8", [1, 255, 256, 257, 700], 16, 8, 12),
    ("b3_32_8", [300, 1025, 64], 32, 8, 13),                 # G = 4 (Qwen3-8B)
    ("b4_8_1", [513, 100, 2048, 31], 8, 1, 14),              # G = 8, one KV head (32B TP=8)
    ("b2_4_1", [777, 1500], 4, 1, 15),                       # 8B TP=8
    ("b2_16_4", [900, 130], 16, 4, 16),                      # 8B TP=2
]

# ---- prefill: (name, lens, cached, Hq, Hkv, seed) ------------------------------------------------
PREFILL_CASES = [
    ("dense_3seq_4_2", [5, 100, 180], None, 4, 2, 21),
    ("dense_1seq_64_8_2", [64], None, 8, 2, 22),
    ("dense_ragged_4_1", [1, 63, 65, 129, 200], None, 4, 1, 23),
    ("dense_g8_8_1", [70, 130], None, 8, 1, 26),
    ("paged_prefix_8_4", [280, 600, 40], [256, 512, 0], 8, 4, 24),
    ("paged_prefix_4_1", [1000, 257], [768, 256], 4, 1, 25),
]


# ---- q/k RMSNorm + RoPE: (name, n_tokens, Hq, Hkv, has_norm, max_position, rope_base, seed) -----
NORMROPE_CASES = [
    ("n1_16_8", 1, 16, 8, True, 4096, 1000000.0, 11),
    ("n37_16_8", 37, 16, 8, True, 4096, 1000000.0, 12),
    ("n64_32_8", 64, 32, 8, True, 40960, 1000000.0, 13),
    ("n19_8_1", 19, 8, 1, True, 40960, 1000000.0, 14),
    ("n23_4_1_nonorm", 23, 4, 1, False, 4096, 10000.0, 15),
]
NORMROPE_EPS = 1e-6


def normrope_inputs(n, Hq, Hkv, has_norm, max_position, seed, D=128):
    """qkv [n, (Hq + 2 Hkv) * D] bf16 ~ N(0, 2^2) (so the norm has work to do), positions int64 in
    [0, max_position) including both ends, bf16 norm weights ~ 1 + 0.25 N(0,1) (None without
    norm: the qkv_bias models of qwen3.py:66-68)."""
    gen = torch.Generator().manual_seed(seed)
    qkv = (2.0 * torch.randn(n, (Hq + 2 * Hkv) * D, generator=gen)).to(torch.bfloat16)
    positions = torch.randint(0, max_position, (n,), generator=gen, dtype=torch.int64)
    positions[0] = max_position - 1
    if n > 1:
        positions[1] = 0
    q_w = k_w = None
    if has_norm:
        q_w = (1.0 + 0.25 * torch.randn(D, generator=gen)).to(torch.bfloat16)
        k_w = (1.0 + 0.25 * torch.randn(D, generator=gen)).to(torch.bfloat16)
    return qkv, positions, q_w, k_w
