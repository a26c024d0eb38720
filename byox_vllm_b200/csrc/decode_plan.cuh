// The unit stream of the decode kernel and its partition into work items and pieces
// (decode_attn.cu). Host + device code: the kernel runs it on the scan it builds in shared memory,
// and b200attn_decode_partition_host runs the very same functions on the host so the partition can
// be checked without a GPU (tests/test_decode_partition.py).
//
//   * unit stream (one per KV head, identical for all heads): sequence b contributes
//     units_b = row_cost + (its 64-token tiles after the cascade's shared pages), 0 for a padding
//     row; cum[] is the exclusive scan, U = cum[B].
//   * the stream is cut into R contiguous ranges; work item i = (range i / Hkv, KV head i % Hkv),
//     so items with adjacent ids cover the SAME tokens of adjacent KV heads (the heads of a token
//     are contiguous in the cache row: CTAs running side by side use whole DRAM pages).
//   * two regimes, chosen on the device from the scan (U, the longest row) and groups = CTAs / Hkv:
//       dynamic: ranges of per = max(ceil(U / (waves x groups)), ceil(longest row / 2), min_units)
//               units, every boundary moved to the nearer end of the row it falls in - so NO row
//               is ever cut, nothing is merged - and CTAs take items beyond their first from an
//               atomic counter: the faster SMs simply take more items. Used when that leaves at
//               least two ranges per CTA group to balance with.
//       static:  otherwise (few long rows): one range of per = max(ceil(U / groups), min_units)
//               units per group, a boundary moved onto a row boundary only when fewer than `snap`
//               tiles would lie on one side of the cut; CTA c walks items c, c + CTAs, ...
//   * a range boundary cuts at most one row; a cut row is held by the items of consecutive ranges
//     g_first .. g_last on its KV head, one piece each, merged by the piece that arrives last.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define B200_HD __host__ __device__ __forceinline__
#else
#define B200_HD inline
#endif

namespace b200 {

constexpr int kDecodeTile = 64;  // tokens per KV tile (one pipeline stage)

B200_HD int imin(int a, int b) { return a < b ? a : b; }
B200_HD int imax(int a, int b) { return a > b ? a : b; }

// Units of sequence b: its streamed tiles + the fixed row cost; 0 for a padding row (ctx <= 0).
B200_HD int decode_row_units(int ctx, int t_skip, int row_cost) {
    return ctx > 0 ? (ctx + kDecodeTile - 1) / kDecodeTile - t_skip + row_cost : 0;
}

struct StreamTunables {
    int min_units;  // fewest units per range
    int row_cost;   // virtual units charged per row
    int snap;       // static regime: boundary snapping distance in tiles
    int waves;      // dynamic regime: ranges per CTA group aimed for (1 = always static)
    int dyn_min;    // dynamic regime: fewest ranges per CTA group worth balancing dynamically
};

struct Stream {
    const int32_t* cum;  // exclusive scan of the per-sequence unit counts, [batch + 1]
    int batch;
    int num_kv_heads;
    int row_cost;
    int t_skip;          // 64-token tiles of the shared prefix skipped in every active row
    int per;             // nominal units per range
    int snap;            // boundaries closer than this many tiles to a row boundary move onto it
    int n_ranges;        // R
    int total_items;     // R * Hkv
    bool dynamic;
};

B200_HD void stream_init(Stream& s, const int32_t* cum, int batch, int num_kv_heads, int n_cta,
                         const StreamTunables& t, int t_skip, int max_row_units) {
    s.cum = cum;
    s.batch = batch;
    s.num_kv_heads = num_kv_heads;
    s.row_cost = t.row_cost;
    s.t_skip = t_skip;
    const int groups = imax(n_cta / num_kv_heads, 1);
    const int U = cum[batch];
    const int min_units = imax(t.min_units, 1);
    s.dynamic = false;
    if (t.waves > 1) {
        const int per = imax(imax((U + groups * t.waves - 1) / (groups * t.waves),
                                  (max_row_units + 1) / 2), min_units);
        if ((U + per - 1) / per >= t.dyn_min * groups) {
            s.dynamic = true;
            s.per = per;
            s.snap = 0;  // unused: every boundary moves to the nearer row end
        }
    }
    if (!s.dynamic) {
        s.per = imax((U + groups - 1) / groups, min_units);
        s.snap = t.snap;
    }
    s.n_ranges = (U + s.per - 1) / s.per;
    s.total_items = s.n_ranges * num_kv_heads;
}

// Largest b with cum[b] <= u, for 0 <= u < cum[batch] (a row with units, by construction).
B200_HD int stream_find_seq(const Stream& s, int u) {
    int lo = 0, hi = s.batch;  // invariant: cum[lo] <= u < cum[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s.cum[mid] <= u) lo = mid; else hi = mid;
    }
    return lo;
}

// Start of range g; *row = a sequence index from which the row holding that unit is reached by
// skipping rows that end at or before it.
B200_HD int stream_bnd(const Stream& s, int g, int* row) {
    const int U = s.cum[s.batch];
    const long long raw = static_cast<long long>(g) * s.per;
    *row = 0;
    if (raw <= 0) return 0;
    if (raw >= U) {
        *row = s.batch;
        return U;
    }
    const int x = static_cast<int>(raw);
    const int b = stream_find_seq(s, x);
    *row = b;
    const int lo = s.cum[b], hi = s.cum[b + 1];
    // dynamic regime: no row is longer than two ranges and none is cut (nearer row boundary)
    if (s.dynamic) return x - lo < hi - x ? lo : hi;
    if (x - lo - s.row_cost < s.snap) return lo;  // fewer than `snap` tiles left of the cut
    if (hi - x < s.snap) return hi;               // ... or right of it
    return x;
}

// Range whose [bnd(g), bnd(g+1)) holds unit x (0 <= x < U).
B200_HD int stream_group_of(const Stream& s, int x) {
    int row;
    int g = imin(x / s.per, s.n_ranges - 1);  // bnd() is monotone and moves a boundary < per
    while (g > 0 && stream_bnd(s, g, &row) > x) --g;
    while (g + 1 < s.n_ranges && stream_bnd(s, g + 1, &row) <= x) ++g;
    return g;
}

// One piece of a (sequence, KV head) row.
struct Piece {
    int b, kvh;
    int range;           // the range (work item = range * Hkv + kvh) this piece belongs to
    int t_begin, t_end;  // 64-token tiles of the sequence
    int tiles_total;     // ceil(ctx / 64)
    bool first, last;    // starts at the row's first streamed tile / ends at its last tile
};

// Ranges that hold the first and the last tile of sequence b's row (only needed for cut rows).
B200_HD void piece_ranges(const Stream& s, int b, int* g_first, int* g_last) {
    *g_first = stream_group_of(s, s.cum[b] + s.row_cost);
    *g_last = stream_group_of(s, s.cum[b + 1] - 1);
}

// Iterates the pieces of one work item. Producer warp and consumer warps run the same iterator
// over the same item sequence, so they agree on the piece list without communicating.
struct ItemIter {
    int hi, u, b, kvh, range;
    B200_HD void init(const Stream& s, int item) {
        range = item / s.num_kv_heads;
        kvh = item - range * s.num_kv_heads;
        int row_hi;
        u = stream_bnd(s, range, &b);
        hi = stream_bnd(s, range + 1, &row_hi);
    }
    // Next non-empty piece; false when the item is exhausted.
    B200_HD bool next(const Stream& s, Piece& pc) {
        while (u < hi) {
            while (s.cum[b + 1] <= u) ++b;  // padding rows, or a boundary snapped to a row end
            const int row_lo = s.cum[b], row_hi = s.cum[b + 1];
            const int seg_hi = imin(hi, row_hi);
            const int ta = imax(u - row_lo - s.row_cost, 0);
            const int tb = imax(seg_hi - row_lo - s.row_cost, 0);
            pc.b = b;
            pc.kvh = kvh;
            pc.range = range;
            pc.t_begin = s.t_skip + ta;
            pc.t_end = s.t_skip + tb;
            pc.tiles_total = s.t_skip + row_hi - row_lo - s.row_cost;
            pc.first = ta == 0;
            pc.last = seg_hi == row_hi;
            u = seg_hi;
            if (tb > ta) return true;
        }
        return false;
    }
};

}  // namespace b200
