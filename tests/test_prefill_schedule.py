"""The work items of the persistent tcgen05 prefill kernel and their order, checked WITHOUT a GPU:
b200attn_prefill_schedule_host runs the kernel's own item code (csrc/prefill_plan.cuh) on the host.

Checked against an independent restatement in Python of what flash_attn_varlen_func's bottom-right
causal mask implies for a 128-row sub-tile (reference nanovllm/layers/attention.py:72-86): which
query tokens a sub-tile holds and how many 128-token KV tiles can contain a visible key."""
import ctypes
import random

import numpy as np
import pytest


@pytest.fixture(scope="module")
def lib(built_lib):
    from byox_vllm_b200 import _lib

    return _lib.load()


def schedule(lib, lens_q, lens_k, Hq, Hkv, max_items=None):
    cu_q = np.concatenate([[0], np.cumsum(lens_q)]).astype(np.int32)
    cu_k = np.concatenate([[0], np.cumsum(lens_k)]).astype(np.int32)
    n = lib.b200attn_prefill_schedule_host(cu_q.ctypes.data, cu_k.ctypes.data, len(lens_q), Hq, Hkv,
                                           None, 0)
    assert n >= 0, n
    cap = n if max_items is None else max_items
    items = np.zeros((max(cap, 1), 8), dtype=np.int32)
    n2 = lib.b200attn_prefill_schedule_host(cu_q.ctypes.data, cu_k.ctypes.data, len(lens_q), Hq,
                                            Hkv, items.ctypes.data, cap)
    assert n2 == n
    return n, items[:min(n, cap)]


def expected_sub_tile(pair, i, Lq, Lk, G):
    """(first token, KV tiles) of sub-tile i of a pair, from the mask definition: query token t of
    a sequence sees keys 0 .. t + (Lk - Lq); a sub-tile holds 128 / G consecutive tokens."""
    toks = 128 // G
    row0 = (2 * pair + i) * toks
    rows = [t for t in range(row0, row0 + toks) if t < Lq]
    if not rows:
        return row0, 0
    last_visible = min(Lk - 1, rows[-1] + (Lk - Lq))
    return row0, 0 if last_visible < 0 else last_visible // 128 + 1


CASES = [
    ("bench_prompts", "rand", 24, 100, 1024, 16, 8),
    ("many_short", "rand", 300, 1, 90, 16, 8),
    ("gqa8", "rand", 17, 1, 3000, 8, 1),
    ("mha", "rand", 9, 1, 1500, 4, 4),
    ("gqa4", "rand", 40, 1, 700, 32, 8),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("cached", [False, True])
def test_items_cover_every_row_once_in_cost_order(lib, case, cached):
    name, _, S, lo, hi, Hq, Hkv = case
    G = Hq // Hkv
    rng = random.Random(len(name) + cached)
    lens_q = [rng.randint(lo, hi) for _ in range(S)]
    lens_q[rng.randrange(S)] = 0 if S > 3 else lens_q[0]           # an empty sequence
    lens_k = [q + (256 * rng.randint(0, 6) if cached and q else 0) for q in lens_q]
    n, items = schedule(lib, lens_q, lens_k, Hq, Hkv)
    toks_per_pair = 256 // G
    want_pairs = {(s, p) for s, q in enumerate(lens_q) for p in range(-(-q // toks_per_pair))}
    got_pairs = [(int(r[0]), int(r[1])) for r in items]
    assert n == len(want_pairs)
    assert len(set(got_pairs)) == n and set(got_pairs) == want_pairs, "an item is missing or doubled"
    for s, p, r00, r01, n0, n1, cost, srt in items.tolist():
        e0 = expected_sub_tile(p, 0, lens_q[s], lens_k[s], G)
        e1 = expected_sub_tile(p, 1, lens_q[s], lens_k[s], G)
        assert (r00, n0) == e0 and (r01, n1) == e1, (name, s, p)
        assert cost == max(n0, n1)
        assert srt == 1
    costs = items[:, 6].tolist()
    assert costs == sorted(costs, reverse=True), "heaviest items first"
    # ties: ascending (sequence, pair) = ascending item index (what makes the order unique)
    for a, b in zip(items[:-1].tolist(), items[1:].tolist()):
        if a[6] == b[6]:
            assert (a[0], a[1]) < (b[0], b[1])


def test_single_sequence_walks_pairs_from_the_end(lib):
    n, items = schedule(lib, [5000], [5000], 32, 8)
    assert n == -(-5000 // 64) and (items[:, 7] == 0).all()
    assert items[:, 1].tolist() == list(range(n - 1, -1, -1))
    assert items[:, 6].tolist() == sorted(items[:, 6].tolist(), reverse=True)


def test_more_items_than_sort_keys_fall_back_to_index_order(lib):
    lens = [66000, 66000]                      # 32 tokens per pair at heads 8/1: 4126 items > 4096
    n, items = schedule(lib, lens, lens, 8, 1)
    assert n == 2 * -(-66000 // 32) and (items[:, 7] == 0).all()
    assert [(int(r[0]), int(r[1])) for r in items[:3]] == [(1, n // 2 - 1), (1, n // 2 - 2),
                                                            (1, n // 2 - 3)]
    assert len({(int(r[0]), int(r[1])) for r in items}) == n


def test_bounds_and_errors(lib):
    n, items = schedule(lib, [300, 700], [300, 700], 16, 8, max_items=3)
    assert n == 3 + 6 and len(items) == 3                     # count returned, only 3 written
    cu = np.zeros(2, dtype=np.int32)
    assert lib.b200attn_prefill_schedule_host(cu.ctypes.data, cu.ctypes.data, 1, 24, 2, None, 0) < 0
    assert lib.b200attn_prefill_schedule_host(None, None, 1, 16, 8, None, 0) < 0
    assert lib.b200attn_prefill_schedule_host(None, None, 0, 16, 8, None, 0) == 0
    # the grid the launcher sizes from host-visible numbers bounds the real item count
    rng = random.Random(5)
    for _ in range(50):
        S = rng.randint(1, 40)
        lens = [rng.randint(0, 900) for _ in range(S)]
        n, _ = schedule(lib, lens, lens, 16, 8)
        assert n <= sum(lens) // 128 + S
