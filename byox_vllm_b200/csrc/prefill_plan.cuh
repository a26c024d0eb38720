// The work items of the persistent tcgen05 prefill kernel (prefill_tc.cu) and their order.
// Host + device code: the kernel runs these functions on the cu_seqlens it copied to shared
// memory, and b200attn_prefill_schedule_host runs the very same functions on the host so that the
// item geometry and the schedule can be checked without a GPU (tests/test_prefill_schedule.py).
//
//   * sequence s contributes ceil(Lq_s / (256 / G)) items per KV head ("pairs": two 128-row
//     sub-tiles; a row is a (query token, group head) pair, 128 / G tokens per sub-tile);
//   * sub-tile i of pair p covers query tokens [row0_i, row0_i + 128 / G) and visits the 128-token
//     KV tiles [0, n_kv_i): causal (bottom-right) skipping, n_kv_i = 0 when it has no row or no
//     visible key;
//   * the cost of an item is the tile count of its heavier sub-tile; items are ordered by
//     descending cost, ties by ascending item index (keys are unique: every CTA sorts to the same
//     list). One sequence (cost is monotone in the pair index) or more items than sort keys:
//     descending item index instead.
#pragma once
#include <cstdint>

#include "decode_plan.cuh"  // B200_HD, imin, imax

namespace b200 {

constexpr int kPrefillRows = 128;     // query rows per sub-tile (= TMEM lanes)
constexpr int kPrefillKvTile = 128;   // KV tokens per tile
constexpr int kSchedKeyBits = 13;     // item index inside a sort key
constexpr uint32_t kSchedCostMax = (1u << (32 - kSchedKeyBits)) - 1;

// 256-row pairs of a sequence with Lq query tokens.
B200_HD int prefill_pairs(int Lq, int group) {
    const int toks_per_pair = 2 * kPrefillRows / group;
    return Lq > 0 ? (Lq + toks_per_pair - 1) / toks_per_pair : 0;
}

struct PrefillSubTiles {
    int row0[2];  // first query token of each sub-tile
    int n_kv[2];  // KV tiles each sub-tile visits
};

B200_HD PrefillSubTiles prefill_sub_tiles(int pair, int Lq, int Lk, int group) {
    PrefillSubTiles t;
    const int toks_per_sub = kPrefillRows / group;
    const int offset = Lk - Lq;  // token i sees keys j <= i + offset
    for (int i = 0; i < 2; ++i) {
        t.row0[i] = (pair * 2 + i) * toks_per_sub;
        const int toks_end = imin(Lq, t.row0[i] + toks_per_sub);
        const int vis = t.row0[i] < Lq ? imin(Lk, toks_end + offset) : 0;  // visible keys [0, vis)
        t.n_kv[i] = vis > 0 ? (vis + kPrefillKvTile - 1) / kPrefillKvTile : 0;
    }
    return t;
}

// KV tiles the heavier sub-tile of a pair visits (the sort key's cost).
B200_HD int prefill_pair_cost(int pair, int Lq, int Lk, int group) {
    const int toks_per_sub = kPrefillRows / group;
    const int toks_end = imin(Lq, (pair * 2 + 2) * toks_per_sub);
    const int vis = imin(Lk, toks_end + (Lk - Lq));
    return vis > 0 ? (vis + kPrefillKvTile - 1) / kPrefillKvTile : 0;
}

// Ascending key order = descending cost, then ascending item index t.
B200_HD uint32_t prefill_sched_key(int cost, int t) {
    const uint32_t c = static_cast<uint32_t>(cost) < kSchedCostMax - 1 ? static_cast<uint32_t>(cost)
                                                                       : kSchedCostMax - 1;
    return ((kSchedCostMax - 1 - c) << kSchedKeyBits) | static_cast<uint32_t>(t);
}
B200_HD int prefill_key_item(uint32_t key) {
    return static_cast<int>(key & ((1u << kSchedKeyBits) - 1u));
}

}  // namespace b200
