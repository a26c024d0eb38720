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
  * every KV head sees the same cut; ranges are balanced (static regime) or whole sequences taken
    from the atomic item counter (dynamic regime: no row is ever cut there).
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


DYN_PCT = 60  # B200ATTN_DECODE_DYN_PCT default


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


MIN_UNITS, DYN_PCT, STATIC_PCT = 16, 200, 50  # B200ATTN_DECODE_MIN_TILES / _DYN_PCT / _STATIC_PCT


def units_of(lens):
    return [-(-c // 64) + 1 if c > 0 else 0 for c in lens]


def expect_regime(lens, Hkv, n_cta):
    """(dynamic, per) as csrc/decode_plan.cuh stream_init decides them."""
    u = units_of(lens)
    groups = max(n_cta // Hkv, 1)
    share = -(-sum(u) // groups)
    if n_cta > 1 and sum(u) > 0 and max(u) * 100 <= DYN_PCT * share:
        return True, max(share, MIN_UNITS)
    g_static = max(groups * STATIC_PCT // 100, 1)
    return False, max(-(-sum(u) // g_static), MIN_UNITS)


def test_static_regime_is_balanced(lib):
    """Few long rows (the longest more than twice a CTA group's share): equal ranges over half the
    CTA groups (one CTA per SM streams evenly; fewer pieces to merge), each range equal to within
    the boundary snap, so a long row is cut instead of being the tail."""
    for lens, Hkv, n_cta in (([16384] * 8, 8, 296), ([32768] * 2, 8, 296), ([131072], 8, 296),
                             ([131072], 1, 296), ([60000] + [300] * 255, 8, 296),
                             ([4096] * 4, 8, 256)):
        pieces = check(lib, lens, Hkv, n_cta)
        per, dynamic = regime(pieces)
        assert (dynamic, per) == expect_regime(lens, Hkv, n_cta)
        assert not dynamic
        n_ranges = int(pieces[:, 10].max()) + 1
        assert n_ranges <= max((n_cta // Hkv) * STATIC_PCT // 100, 1)
        units = np.zeros(n_ranges, dtype=np.int64)
        for item, b, kvh, t0, t1, tot, first, *_ in pieces.tolist():
            if kvh == 0:
                units[item // Hkv] += t1 - t0 + (1 if first else 0)  # row cost on the first piece
        assert units.sum() == sum(units_of(lens))
        busy = units[units > 0]
        assert busy.max() <= per + 2 * 2, "a range grows by less than the snap at each end"


def test_dynamic_regime_never_cuts_a_row(lib):
    """Many rows (the longest at most twice a CTA group's share): a range is one sequence, an item
    one whole (sequence, KV head) row in batch order: no row is cut, nothing is merged, no search
    is needed, and CTAs beyond their first item take rows from the atomic counter."""
    rng = random.Random(1)
    lens = [rng.randint(169, 1502) for _ in range(256)]
    for Hkv in (8, 2, 1):  # bench.py's c2 shape at TP = 1, 4, 8
        pieces = check(lib, lens, Hkv, 296)
        per, dynamic = regime(pieces)
        assert dynamic and (dynamic, per) == expect_regime(lens, Hkv, 296)
        assert all(f and l for f, l in pieces[:, 6:8].tolist()), "no row may be cut"
        assert pieces[:, 0].tolist() == list(range(256 * Hkv)), "item = sequence * Hkv + kv head"
        assert pieces[:, 10].tolist() == pieces[:, 1].tolist(), "range id = sequence"
    # padding rows are items without a piece
    pieces = check(lib, [0 if i % 5 == 0 else l for i, l in enumerate(lens)], 8, 296)
    assert regime(pieces)[1] and len(pieces) == 8 * sum(1 for i in range(256) if i % 5)
    # config 3 (128 x 4096), 64 x 4096, 32 x 4096 (65 units against a share of 57): dynamic
    for B in (128, 64, 32):
        pieces = check(lib, [4096] * B, 8, 296)
        assert regime(pieces)[1] and len(pieces) == B * 8
    # 16 x 4096 (share 29): rows are cut
    assert not regime(check(lib, [4096] * 16, 8, 296))[1]
    # a single CTA (no counters registered) walks every item statically
    pieces = check(lib, lens, 8, 1)
    assert not regime(pieces)[1] and all(f and l for f, l in pieces[:, 6:8].tolist())


def test_closed_form_first_and_last_range(lib):
    """b200attn_decode_partition_host cross-checks the O(1) first / last range of every cut row
    against the definition (search over the snapped boundaries) and fails with < -1000 on a
    mismatch; check() asserts n >= 0. Ragged batches with many cuts, all snap cases."""
    rng = random.Random(7)
    for _ in range(300):
        B = rng.choice([1, 2, 3, 5, 9, 17, 40])
        Hkv = rng.choice([1, 2, 8])
        n_cta = rng.choice([8, 16, 37, 74, 148, 296])
        lens = [rng.choice([0, 64, 65, 128, 129, 1000, 1023, 1024, 1025, rng.randint(1, 70000)])
                for _ in range(B)]
        check(lib, lens, Hkv, n_cta)


def test_boundary_snap_avoids_tiny_pieces(lib):
    """A boundary with fewer than `snap` (2) tiles of a row on one side moves onto the row
    boundary: no one-tile slivers next to a row end."""
    pieces = check(lib, [1024] * 16, 8, 296)
    assert min((pieces[:, 4] - pieces[:, 3]).tolist()) >= 2
    pieces = check(lib, [1024], 8, 17)
    assert min((pieces[:, 4] - pieces[:, 3]).tolist()) >= 2
