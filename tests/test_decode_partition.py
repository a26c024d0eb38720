"""The decode kernel's work partition, checked on the CPU through the C-ABI.

b200attn_decode_partition_host runs the kernel's own partition code (csrc/decode_plan.cuh: the
unit stream, the ranges, the piece iterator of a work item) on host arrays, so the invariants the
kernel's correctness rests on are pinned without a GPU:
  * every 64-token tile of every (sequence, KV head) row is attended exactly once, in order;
  * padding rows (context_lens <= 0) and the cascade's shared pages produce no work;
  * a row cut across ranges is held by the items of consecutive ranges range_first..range_last on
    its KV head, one piece each - the arrival count the in-launch merge waits for - and every
    piece reports the same pair;
  * an item writes at most one piece per workspace slot (2i = continues an earlier range's row,
    2i+1 = continued by the next range), and the slots fit the workspace the ABI sizes;
  * every KV head sees the same cut; ranges are balanced (static regime) or fine enough for the
    atomic item counter to balance them (dynamic regime), and short rows are never cut there.
"""
import ctypes
import random

import numpy as np
import pytest


@pytest.fixture(scope="module")
def lib(built_lib):
    from byox_vllm_b200 import _lib

    return _lib.load()


def partition(lib, lens, Hkv, n_cta, shared_pages=0, block_size=256):
    lens = np.asarray(lens, dtype=np.int32)
    cap = 16 * (len(lens) * Hkv + n_cta) + 16
    out = np.zeros((cap, 12), dtype=np.int32)
    n = lib.b200attn_decode_partition_host(
        lens.ctypes.data_as(ctypes.c_void_p), len(lens), Hkv, n_cta, shared_pages, block_size,
        out.ctypes.data_as(ctypes.c_void_p), cap)
    assert 0 <= n <= cap, n
    return out[:n]


WAVES = 8  # B200ATTN_DECODE_WAVES default


def check(lib, lens, Hkv, n_cta, shared_pages=0, block_size=256):
    pieces = partition(lib, lens, Hkv, n_cta, shared_pages, block_size)
    t_skip = shared_pages * (block_size // 64)
    covered = {}
    slots = set()
    per_row = {}
    prev_item = -1
    for item, b, kvh, t0, t1, total, first, last, g_first, g_last, rng_, per in pieces.tolist():
        assert item >= prev_item, "pieces must come in item order"
        prev_item = item
        assert 0 <= b < len(lens) and 0 <= kvh < Hkv
        assert item == rng_ * Hkv + kvh
        assert lens[b] > 0, "padding rows carry no work"
        assert total == -(-lens[b] // 64)
        assert t_skip <= t0 < t1 <= total
        assert first == (t0 == t_skip) and last == (t1 == total)
        row = covered.setdefault((b, kvh), [])
        assert (row[-1][1] if row else t_skip) == t0, "tiles attended in order, exactly once"
        row.append((t0, t1))
        if not (first and last):
            assert g_first <= rng_ <= g_last
            per_row.setdefault((b, kvh), []).append((rng_, g_first, g_last))
            slot = 2 * item + (1 if first else 0)
            assert slot not in slots, "two partial pieces in one workspace slot"
            assert slot < 2 * (n_cta + Hkv), "slot beyond the workspace the ABI sizes"
            slots.add(slot)
    for b, ctx in enumerate(lens):
        for kvh in range(Hkv):
            if ctx > 0:
                assert covered[(b, kvh)][-1][1] == -(-ctx // 64), f"row {(b, kvh)} incomplete"
            else:
                assert (b, kvh) not in covered
    for key, lst in per_row.items():
        g_first, g_last = lst[0][1], lst[0][2]
        assert g_last > g_first
        assert [g for g, _, _ in lst] == list(range(g_first, g_last + 1)), key
        assert all(gf == g_first and gl == g_last for _, gf, gl in lst), key
    # every KV head sees the same cut (items with adjacent ids stream the same tokens)
    by_head = {}
    for item, b, kvh, t0, t1, *_rest in pieces.tolist():
        by_head.setdefault(kvh, []).append((item // Hkv, b, t0, t1))
    ref = by_head.get(0, [])
    for kvh in range(Hkv):
        assert by_head.get(kvh, []) == ref
    return pieces


def regime(pieces):
    per = int(pieces[0, 11])
    return per & ~(1 << 30), bool(per >> 30)


def test_uniform_and_ragged_batches(lib):
    rng = random.Random(0)
    for _ in range(40):
        B = rng.choice([1, 2, 3, 7, 64, 256, 512])
        Hkv = rng.choice([1, 2, 4, 8])
        n_cta = rng.choice([1, 2, 3, 5, 37, 148, 296])
        lens = [rng.choice([0, 1, 63, 64, 65, 255, 256, 257, rng.randint(1, 5000)])
                for _ in range(B)]
        check(lib, lens, Hkv, n_cta)


def test_baseline_shapes(lib):
    check(lib, [4096] * 128, 8, 296)               # config 3
    check(lib, [8192] * 256, 1, 296)               # config 4 per rank, plain
    check(lib, [8192] * 256, 1, 296, shared_pages=30)  # config 4 with the cascade armed
    check(lib, [131072], 8, 296)                   # config 5 corner: one row across many CTAs
    check(lib, [131072], 1, 296)
    check(lib, [1024] * 512, 8, 296)


def test_padding_rows_and_empty_batch(lib):
    check(lib, [0, 0, 0], 8, 296)
    check(lib, [300, 0, 700, 0, 0, 90], 8, 5)
    check(lib, [0] * 7 + [1], 1, 296)
    assert len(partition(lib, [], 8, 4)) == 0


def test_shared_prefix_is_skipped(lib):
    pieces = check(lib, [900, 1000, 1100, 0, 800], 2, 7, shared_pages=3)
    assert pieces[:, 3].min() == 12  # no piece starts before the 3 shared pages (12 tiles)


def test_static_regime_is_balanced(lib):
    """bench.py's c2 shape at TP=8 / TP=4 (Hkv 1 / 2): ragged 169..1502 tokens, about 12 / 24 units
    per CTA group: one range per group, equal to within the boundary snap, and the longest row
    (24 tiles) is cut instead of being the tail."""
    rng = random.Random(1)
    lens = [rng.randint(169, 1502) for _ in range(256)]
    total = sum(-(-c // 64) + 1 for c in lens)
    for Hkv, n_cta in ((1, 296), (2, 296)):
        pieces = check(lib, lens, Hkv, n_cta)
        per, dynamic = regime(pieces)
        n_groups = n_cta // Hkv
        assert not dynamic and per == max(-(-total // n_groups), 8)
        units = np.zeros(n_groups, dtype=np.int64)
        for item, b, kvh, t0, t1, tot, first, *_ in pieces.tolist():
            if kvh == 0:
                units[item // Hkv] += t1 - t0 + (1 if first else 0)  # row cost on the first piece
        assert units.sum() == total
        busy = units[units > 0]
        assert busy.max() <= per + 2 * 2, "a range grows by less than the snap at each end"
        if Hkv == 1:
            assert busy.max() < 24 + 1


def test_dynamic_regime_never_cuts_a_row(lib):
    """c2 at TP=1 (Hkv 8, ~97 units per CTA group): ranges of half the longest row or 1/8 of a
    group's share, every boundary on a row boundary: no row is cut, nothing is merged, and there
    are several items per CTA for the atomic counter to balance. Few long rows fall back to the
    static regime (cutting them finer would only add pieces to merge)."""
    rng = random.Random(1)
    lens = [rng.randint(169, 1502) for _ in range(256)]
    units = [-(-c // 64) + 1 for c in lens]
    pieces = check(lib, lens, 8, 296)
    per, dynamic = regime(pieces)
    assert dynamic and per == max(-(-sum(units) // (37 * WAVES)), -(-max(units) // 2), 8)
    assert all(f and l for f, l in pieces[:, 6:8].tolist()), "no row may be cut"
    n_items = len({int(i) for i in pieces[:, 0]})
    assert 2 * 296 <= n_items <= WAVES * 296
    # config 3 (128 x 4096): rows of 65 units -> ranges of 33: one row per item, never cut
    pieces = check(lib, [4096] * 128, 8, 296)
    assert regime(pieces) == (33, True)
    assert all(f and l for f, l in pieces[:, 6:8].tolist()) and len(pieces) == 128 * 8
    # one 128k-token sequence / 8 x 16k / 16 x 4096: too few rows to balance uncut -> static
    for lens_, groups in (([131072], 37), ([16384] * 8, 37), ([4096] * 16, 37)):
        pieces = check(lib, lens_, 8, 296)
        per, dynamic = regime(pieces)
        assert not dynamic and per == -(-sum(-(-c // 64) + 1 for c in lens_) // groups)
    # a single CTA (no counters registered) walks every item statically
    pieces = check(lib, lens, 8, 1)
    assert not regime(pieces)[1] and all(f and l for f, l in pieces[:, 6:8].tolist())


def test_boundary_snap_avoids_tiny_pieces(lib):
    """A boundary with fewer than `snap` (2) tiles of a row on one side moves onto the row
    boundary: no one-tile slivers next to a row end."""
    pieces = check(lib, [1024] * 16, 8, 296)
    assert min((pieces[:, 4] - pieces[:, 3]).tolist()) >= 2
    pieces = check(lib, [1024], 8, 17)
    assert min((pieces[:, 4] - pieces[:, 3]).tolist()) >= 2
