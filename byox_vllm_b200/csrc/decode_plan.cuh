// The unit stream of the decode kernel and its partition into work items and pieces
// (decode_attn.cu). Host + device code: the kernel runs it on the scan it builds in shared memory,
// and b200attn_decode_partition_host runs the very same functions on the host so the partition can
// be checked without a GPU (tests/test_decode_partition.py).
//
//   * unit stream (one per KV head, identical for all heads): sequence b contributes
//     units_b = row_cost + (its 64-token tiles after the cascade's shared pages), 0 for a padding
//     row; cum[] is the exclusive scan, U = cum[B].
//   * work item i = (range i / Hkv, KV head i % Hkv), so items with adjacent ids cover the SAME
//     tokens of adjacent KV heads (the heads of a token are contiguous in the cache row: CTAs
//     running side by side use whole DRAM pages).
//   * two regimes, chosen on the device from the scan (U, the longest row) and groups = CTAs / Hkv:
//       dynamic: a range is one whole sequence (range id = b): NO row is cut, nothing is merged,
//               no search is needed to find an item's row, and CTAs take items beyond their first
//               from an atomic counter - the faster SMs simply take more rows. Used when the
//               longest row is at most dyn_pct % of a CTA group's share ceil(U / groups): list
//               scheduling then ends within one longest row of the balanced finish.
//       static:  otherwise (few or long rows): the stream is cut into one range of
//               per = max(ceil(U / groups), min_units) units per group, a boundary moved onto a
//               row boundary only when fewer than `snap` tiles would lie on one side of the cut;
//               CTA c walks items c, c + CTAs, ...
//   * static regime: a range boundary cuts at most one row; a cut row is held by the items of
//     consecutive ranges g_first .. g_last on its KV head, one piece each, merged by the piece
//     that arrives last.
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

// x / d for 0 <= x < 2^24, d >= 1 with a precomputed float reciprocal: a handful of instructions
// instead of the ~25 of an integer division (these sit on the latency path of every CTA's start).
B200_HD int idiv_rcp(int x, int d, float rcp) {
    int q = static_cast<int>(static_cast<float>(x) * rcp);
    const int r = x - q * d;
    if (r < 0) --q;
    else if (r >= d) ++q;
    return q;
}

// Units of sequence b: its streamed tiles + the fixed row cost; 0 for a padding row (ctx <= 0).
B200_HD int decode_row_units(int ctx, int t_skip, int row_cost) {
    return ctx > 0 ? (ctx + kDecodeTile - 1) / kDecodeTile - t_skip + row_cost : 0;
}

struct StreamTunables {
    int min_units;  // fewest units per range
    int row_cost;   // virtual units charged per row
    int snap;       // static regime: boundary snapping distance in tiles
    int dyn_ok;     // 0: always static (no scheduler counters registered, or switched off)
    int dyn_pct;    // dynamic regime when longest row * 100 <= dyn_pct * ceil(U / groups)
    int static_pct; // static regime: percent of the CTA groups that get a range (fewer, longer
                    // pieces: less to merge, and one CTA per SM streams more evenly than two)
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
    float rcp_per, rcp_hkv;
};

B200_HD void stream_init(Stream& s, const int32_t* cum, int batch, int num_kv_heads, int n_cta,
                         const StreamTunables& t, int t_skip, int max_row_units) {
    s.cum = cum;
    s.batch = batch;
    s.num_kv_heads = num_kv_heads;
    s.row_cost = t.row_cost;
    s.t_skip = t_skip;
    s.rcp_hkv = 1.0f / static_cast<float>(num_kv_heads);
    const int groups = imax(idiv_rcp(n_cta, num_kv_heads, s.rcp_hkv), 1);
    const int U = cum[batch];
    const int min_units = imax(t.min_units, 1);
    const float rcp_groups = 1.0f / static_cast<float>(groups);
    const int share = idiv_rcp(U + groups - 1, groups, rcp_groups);
    s.snap = t.snap;
    // (float compare: both sides are far below 2^24 * 100)
    s.dynamic = t.dyn_ok != 0 && U > 0 &&
                static_cast<float>(max_row_units) * 100.f <=
                    static_cast<float>(t.dyn_pct) * static_cast<float>(share);
    if (s.dynamic) {
        s.per = imax(share, min_units);  // reported only
        s.n_ranges = batch;              // one range per sequence
    } else {
        const int g_static = imax(groups * t.static_pct / 100, 1);
        const int share_s = idiv_rcp(U + g_static - 1, g_static, 1.0f / static_cast<float>(g_static));
        s.per = imax(share_s, min_units);
        s.rcp_per = 1.0f / static_cast<float>(s.per);
        s.n_ranges = idiv_rcp(U + s.per - 1, s.per, s.rcp_per);
    }
    if (s.dynamic) s.rcp_per = 1.0f / static_cast<float>(s.per);
    s.total_items = s.n_ranges * num_kv_heads;
}

// Largest b with cum[b] <= u, for 0 <= u < cum[batch] (a row with units, by construction).
B200_HD int stream_find_seq(const Stream& s, int u) {
    int lo = 0, hi = s.batch;  // invariant: cum[lo] <= u < cum[hi]
#if defined(__CUDA_ARCH__)
    // Device callers are whole converged warps with identical arguments (the producer warp): a
    // 33-ary search, every lane probing one position per round (2 rounds for 1024 rows).
    const int lane_ = threadIdx.x & 31;
    while (hi - lo > 1) {
        const int step = (hi - lo + 32) / 33;                  // probes at lo + (k+1) * step
        const int pos = lo + (lane_ + 1) * step;
        const bool le = pos < hi && s.cum[pos] <= u;           // monotone in lane
        const int n = __popc(__ballot_sync(0xffffffffu, le));  // probes at or below u
        const int new_lo = lo + n * step;
        hi = min(hi, new_lo + step);
        lo = new_lo;
    }
    return lo;
#endif
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
    if (x - lo - s.row_cost < s.snap) return lo;  // fewer than `snap` tiles left of the cut
    if (hi - x < s.snap) return hi;               // ... or right of it
    return x;
}

// Range whose [bnd(g), bnd(g+1)) holds unit x (0 <= x < U).
B200_HD int stream_group_of(const Stream& s, int x) {
    int row;
    int g = imin(idiv_rcp(x, s.per, s.rcp_per), s.n_ranges - 1);  // bnd() moves a boundary < per
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

// Ranges that hold the first and the last tile of sequence b's row (static regime, cut rows only).
B200_HD void piece_ranges(const Stream& s, int b, int* g_first, int* g_last) {
    *g_first = stream_group_of(s, s.cum[b] + s.row_cost);
    *g_last = stream_group_of(s, s.cum[b + 1] - 1);
}

// Iterates the pieces of one work item (producer warp; the consumers get finished descriptors).
struct ItemIter {
    int hi, u, b, kvh, range;
    B200_HD void init(const Stream& s, int item) {
        range = idiv_rcp(item, s.num_kv_heads, s.rcp_hkv);
        kvh = item - range * s.num_kv_heads;
        if (s.dynamic) {  // the whole row of sequence `range` (empty for a padding row)
            b = range;
            u = s.cum[b];
            hi = s.cum[b + 1];
            return;
        }
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

// What the producer warp hands to the consumers for one piece (shared-memory queue, 32 bytes).
// b < 0 ends the CTA. whole row <=> g_first == g_last.
struct PieceDesc {
    int b, kvh, t_begin, t_end, tiles_total, range, g_first, g_last;
};

// Range that holds the first tile of row b (static regime), in O(1): with g0 = x0 / per for the
// tile's unit x0 = cum[b] + row_cost, the start of g0 is at or before x0 (a boundary inside row b
// at or before its first tile snaps onto the row start), and the start of g0 + 1 lies in row b
// after x0 - unless it snaps back onto the row start, in which case g0 + 1 holds the tile.
B200_HD int row_first_range(const Stream& s, int b) {
    const int lo = s.cum[b], hi = s.cum[b + 1];
    const int g0 = idiv_rcp(lo + s.row_cost, s.per, s.rcp_per);
    const int raw1 = (g0 + 1) * s.per;
    if (raw1 < hi && raw1 - lo - s.row_cost < s.snap) return imin(g0 + 1, s.n_ranges - 1);
    return imin(g0, s.n_ranges - 1);
}
// Range that holds the last tile of row b: g1 = (cum[b+1] - 1) / per, unless the start of g1 lies
// in row b within `snap` tiles of its end and is therefore moved onto the row end.
B200_HD int row_last_range(const Stream& s, int b) {
    const int lo = s.cum[b], hi = s.cum[b + 1];
    const int g1 = imin(idiv_rcp(hi - 1, s.per, s.rcp_per), s.n_ranges - 1);
    const int raw = g1 * s.per;
    if (raw > lo && !(raw - lo - s.row_cost < s.snap) && hi - raw < s.snap) return g1 - 1;
    return g1;
}

B200_HD PieceDesc piece_desc(const Stream& s, const Piece& pc) {
    PieceDesc d;
    d.b = pc.b;
    d.kvh = pc.kvh;
    d.t_begin = pc.t_begin;
    d.t_end = pc.t_end;
    d.tiles_total = pc.tiles_total;
    d.range = pc.range;
    d.g_first = d.g_last = pc.range;
    if (!(pc.first && pc.last)) {
        // a piece that holds the row's first (last) tile is itself the first (last) range
        if (!pc.first) d.g_first = row_first_range(s, pc.b);
        if (!pc.last) d.g_last = row_last_range(s, pc.b);
    }
    return d;
}

}  // namespace b200
