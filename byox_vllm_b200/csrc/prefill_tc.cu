// Varlen causal prefill on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), sm_100a.
// Replaces flash_attn_varlen_func on the prefill branch of Attention.forward
// (reference nanovllm/layers/attention.py:72-86), dense and paged-prefix K/V.
//
// Work item = 256 query rows (two 128-row sub-tiles) of one (sequence, KV head); a row is a (query
// token, group head) pair. PERSISTENT: a CTA walks a cost-ordered item list that every CTA derives
// identically from cu_seqlens in its prologue (pair counts -> scan -> bitonic sort of (KV tiles,
// item) keys in shared memory, heaviest first). Which positions of the list a CTA takes:
//   clc     (default) the grid has one CTA per possible item; a running CTA takes its own
//           blockIdx first and then STEALS the index of a not-yet-launched CTA with
//           clusterlaunchcontrol.try_cancel (the hardware work queue of sm_100: dynamic, needs no
//           global counter, nothing to allocate or reset) until the grid is drained;
//   static  grid = SM count, CTA c takes positions c, 2P-1-c, 2P+c, ... (zig-zag);
//   grid    one item per CTA, no stealing (the round-1 behaviour; B200ATTN_PREFILL_SCHED picks).
// All barriers, the TMEM allocation and the K/V rings live across items:
// the producer prefetches the next item's first K/V tiles and (once the last S MMA of the current
// item has retired) its Q tiles while the current item's last softmax / PV / epilogue run.
// 12 warps (the last warpgroup gives most of its registers to the softmax groups, setmaxnreg):
//   warps 0-3   softmax group 0: one thread per query row of sub-tile 0 (TMEM lane = row)
//   warps 4-7   softmax group 1: sub-tile 1
//   warp  8     TMA producer: Q per item, 128-token K and V tiles through two 2-stage rings
//               (cp.async.bulk.tensor, SWIZZLE_128B boxes of 64 dims x 128 rows, mbarrier tx)
//   warp  9     MMA issuer (one elected lane) + TMEM owner
// TMEM (512 columns x 128 lanes of fp32): S0 | S1 | O0 | O1, 128 columns each.
//   S_i = Q_i . K_j^T      tcgen05.mma SS, M128 N128 K16 x 8, both operands K-major in smem
//   P_i = exp2(S_i*c - m)  softmax thread reads its S row with tcgen05.ld, writes bf16 P over the
//                          first 64 columns of S_i with tcgen05.st (fp32 max/sum in registers)
//   O_i += P_i . V_j       tcgen05.mma TS: A = P from TMEM, B = V tile MN-major in smem
// The issuer interleaves the two sub-tiles (PV_0[j], S_0[j+1], PV_1[j], S_1[j+1], ...) so the
// tensor pipe works on one sub-tile while the other one is in softmax. tcgen05.mma executes in
// issue order, so S_i[j+1] may overwrite P_i[j] as soon as it is issued after PV_i[j]; the same
// argument carries over item boundaries (S_i[0] of the next item is issued after PV_i[last], and
// its first PV — which overwrites O_i — only after the softmax group has finished the epilogue
// reads of O_i and signalled the next P).
// O is only rescaled when a row maximum grows by more than 2^8 (lazy rescale); the row sum uses
// the same stale maximum, so O / l is exact at the end.
//
// Roofline: bf16 tensor pipe. Algorithmic flops = 4 * 128 * (visible q-k pairs) per q-head.
#include "common.h"
#include "prefill_plan.cuh"
#include "ptx.cuh"
#include "tcgen05.cuh"

#include <algorithm>
#include <cstdlib>
#include <type_traits>

namespace b200 {

namespace {

constexpr int kTM = kPrefillRows;              // query rows per sub-tile (= TMEM lanes)
constexpr int kTN = kPrefillKvTile;            // KV tokens per tile
constexpr int kStages = 2;                     // K ring and V ring depth
constexpr int kHalfBytes = 128 * 128;          // [128 rows][64 dims] bf16, 128-byte rows, SW128
constexpr int kTileBytes = 2 * kHalfBytes;     // 32 KiB: dims 0-63 | dims 64-127
constexpr int kProducerWarp = 8;
constexpr int kMmaWarp = 9;
constexpr int kTcThreads = 384;                // 3 warpgroups: softmax 0, softmax 1, {TMA, MMA, 2 idle}
// setmaxnreg budgets. The pool is what the CTA got at launch (168 regs x 384 threads), not the
// whole register file: an inc that does not fit waits forever.
constexpr int kRegsLaunch = 168;
constexpr int kRegsSoftmax = 216;
constexpr int kRegsControl = 72;
static_assert(256 * kRegsSoftmax + 128 * kRegsControl <= kTcThreads * kRegsLaunch,
              "setmaxnreg budgets exceed the registers the CTA owns");
constexpr uint32_t kTmemCols = 512;
constexpr float kRescaleThreshold = 8.0f;      // log2 domain

// Item schedule (persistent varlen prefill): exclusive scan of the 256-row pairs per sequence and
// the sorted keys, both in shared memory. Batches beyond kTcMaxSeqs take the shape fallback
// (api.cu); beyond kTcMaxItems pairs they run unsorted.
constexpr uint32_t kSmemWindowBase = 0x400;  // first 1 KiB of the window is reserved by the system
constexpr int kSmemQ = 0;
constexpr int kSmemK = kSmemQ + 2 * kTileBytes;
constexpr int kSmemV = kSmemK + kStages * kTileBytes;
constexpr int kSmemBar = kSmemV + kStages * kTileBytes;
constexpr int kClcSlots = 4;                                               // ticket ring
constexpr int kSmemClc = kSmemBar + 256;                                    // 16-byte responses
// cu_seqlens_q | cu_seqlens_k | pair_start, int32[kTcMaxSeqs + 1] each: the item walk never touches
// global memory (a load issued behind an epilogue's stores waits for all of them in the LSU)
constexpr int kSchedStride = kTcMaxSeqs + 1;
constexpr int kItemSlots = 4;                                               // decoded-item ring
constexpr int kItemWords = 20;                                              // int32 per slot
constexpr int kSmemItems = kSmemClc + kClcSlots * 16;
constexpr int kSmemPairStart = kSmemItems + kItemSlots * kItemWords * 4;
constexpr int kSmemKeys = kSmemPairStart + 3 * kSchedStride * 4;            // uint32[kTcMaxItems]
constexpr int kTcSmemBytes = kSmemKeys + kTcMaxItems * 4 + 1024;
static_assert(kTcMaxItems <= (1 << kSchedKeyBits), "item index does not fit the key");
static_assert(kTcSmemBytes <= 232448, "more shared memory than an sm_100 CTA can opt into");

struct Barriers {
    uint64_t q_full, q_empty;
    uint64_t k_full[kStages], k_empty[kStages];
    uint64_t v_full[kStages], v_empty[kStages];
    uint64_t s_full[2], p_full[2], o_full[2];
    uint64_t sched_full[kClcSlots];   // try_cancel response k+1 has landed in slot k % kClcSlots
    uint64_t item_full[kItemSlots];   // the producer has published item k in slot k % kItemSlots
    uint64_t item_empty[kItemSlots];  // ... and the issuer + the 8 softmax warps have read it
    uint32_t tmem_base;
    int32_t num_pairs;   // 256-row pairs over all sequences (items per KV head)
    int32_t use_keys;    // 1: keys[] holds the cost-sorted order; 0: descending item index
};
constexpr int kItemReaders = 1 + kProducerWarp;  // MMA warp + softmax warps (lane 0 arrives)
static_assert(sizeof(Barriers) <= 256, "barrier block overflows its shared-memory slot");

// Bounded wait: a protocol bug traps (launch error) after ~20 s instead of hanging the GPU
// (mbarrier.try_wait itself may block for a while, so the bound is on the clock, not on spins).
__device__ __forceinline__ void wait_bar(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++spins & 1023u) == 0 && clock64() - t0 > 40000000000LL) __trap();
    }
}

// One work item: two 128-row sub-tiles of one (sequence, KV head) — or, in cascade mode, of the
// batch rows of one KV head against one chunk of the shared prefix.
struct Item {
    int seq, kvh, q_begin, Lq, Lk, k_begin, offset;
    int j0;       // first 128-token KV tile (cascade: start of the chunk)
    int chunk;    // cascade: prefix chunk = workspace slot ws_slot0 + chunk
    int q_c1;     // TMA head coordinate of the Q boxes
    int row0[2], n_kv[2], q_c2[2];
    int n_max;
};

// clusterlaunchcontrol: ask the hardware scheduler to cancel a CTA of this grid that has not
// started yet; the 16-byte response (cancelled? + its blockIdx) lands in shared memory and
// completes 16 bytes on the mbarrier.
__device__ __forceinline__ void clc_try_cancel(void* resp, uint64_t* bar) {
    asm volatile(
        "clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.b128 "
        "[%0], [%1];" ::"r"(smem_u32(resp)), "r"(smem_u32(bar))
        : "memory");
}
// Block index handed over by a successful cancellation, -1 when nothing was left to cancel.
__device__ __forceinline__ int clc_decode(const void* resp) {
    const ulonglong2 r = *reinterpret_cast<const ulonglong2*>(resp);
    uint32_t ok, x;
    asm volatile(
        "{\n\t.reg .b128 r;\n\t.reg .pred p;\n\t"
        "mov.b128 r, {%2, %3};\n\t"
        "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p, r;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t"
        "mov.b32 %1, 0;\n\t"
        "@p clusterlaunchcontrol.query_cancel.get_first_ctaid::x.b32.b128 %1, r;\n\t}"
        : "=r"(ok), "=r"(x)
        : "l"(r.x), "l"(r.y));
    return ok ? static_cast<int>(x) : -1;
}

enum : int { kSchedClc = 0, kSchedStatic = 1, kSchedGrid = 2 };

// The k-th item of this CTA (false: no more); an item may be empty (n_max == 0: a ticket beyond
// the real item count, or no visible key). Only the producer thread runs this: it publishes the
// decoded item to the other roles through a shared-memory ring (publish_item / consume_item), so
// the ticket decode, the searches and their (cold) code are paid once per item, ahead of time.
template <bool kCascade>
__device__ __forceinline__ bool fetch_item(int k, const PrefillParams& p, Barriers* bar,
                                           const uint8_t* clc, const int32_t* pair_start,
                                           const uint32_t* keys, Item& it) {
    if constexpr (kCascade) {
        if (k > 0) return false;
        const PrefixChunks pc = prefix_chunks_of(p.plan[0], p.block_size, p.prefix_chunks);
        it.chunk = blockIdx.y;  // < pc.valid (checked at kernel entry)
        it.kvh = blockIdx.z;
        it.seq = p.plan[1];     // row whose table names the shared pages
        it.j0 = it.chunk * pc.per;
        it.q_begin = it.Lq = it.k_begin = it.offset = 0;
        it.Lk = 0;
        const int n = min(pc.per, pc.tiles - it.j0);
        const int rows_per_sub = kTM / p.group;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            it.row0[i] = (static_cast<int>(blockIdx.x) * 2 + i) * rows_per_sub;  // first batch row
            it.n_kv[i] = it.row0[i] < p.batch ? n : 0;
            it.q_c2[i] = it.row0[i];
        }
    } else {
        const int T = bar->num_pairs;
        const int64_t total = static_cast<int64_t>(T) * p.num_kv_heads;
        int64_t g;
        if (p.sched_mode == kSchedStatic) {
            // zig-zag: round k runs left to right over the CTAs, round k+1 right to left
            const int P = gridDim.x, c = blockIdx.x;
            g = static_cast<int64_t>(k) * P + ((k & 1) ? P - 1 - c : c);
            if (g >= total) return false;
        } else if (k == 0) {
            g = blockIdx.x;
        } else {
            if (p.sched_mode == kSchedGrid) return false;
            const int slot = (k - 1) % kClcSlots;
            wait_bar(&bar->sched_full[slot], ((k - 1) / kClcSlots) & 1u);
            const int t = clc_decode(clc + slot * 16);
            if (t < 0) return false;  // grid drained
            g = t;
        }
        if (g >= total) {  // surplus ticket (the grid is sized from an upper bound): no rows
            it.n_kv[0] = it.n_kv[1] = it.n_max = 0;
            it.kvh = it.q_begin = it.Lq = it.Lk = it.offset = 0;
            it.row0[0] = it.row0[1] = 0;
            return true;
        }
        const int li = static_cast<int>(g / p.num_kv_heads);
        it.kvh = static_cast<int>(g - static_cast<int64_t>(li) * p.num_kv_heads);
        const int t = bar->use_keys ? prefill_key_item(keys[li]) : T - 1 - li;
        const int32_t* cu_q = pair_start - 2 * kSchedStride;  // shared-memory copies
        const int32_t* cu_k = pair_start - kSchedStride;
        int lo = 0, hi = p.num_seqs;  // largest seq with pair_start[seq] <= t
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (pair_start[mid] <= t) lo = mid; else hi = mid;
        }
        it.seq = lo;
        const int pair = t - pair_start[lo];
        it.q_begin = cu_q[lo];
        it.Lq = cu_q[lo + 1] - it.q_begin;
        it.k_begin = cu_k[lo];
        it.Lk = cu_k[lo + 1] - it.k_begin;
        it.offset = it.Lk - it.Lq;  // token i sees keys j <= i + offset
        it.j0 = 0;
        it.chunk = 0;
        // GQA packing: a 128-row sub-tile = 128/G query tokens x the G heads of one KV head
        // (row r = token r / G, head r % G; one TMA box 64 dims x G heads x 128/G tokens), so a
        // K/V tile is loaded once per KV head and short query blocks still fill the tile.
        const PrefillSubTiles st = prefill_sub_tiles(pair, it.Lq, it.Lk, p.group);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            it.row0[i] = st.row0[i];
            it.n_kv[i] = st.n_kv[i];
            it.q_c2[i] = it.q_begin + st.row0[i];
        }
    }
    it.q_c1 = it.kvh * p.group;
    it.n_max = max(it.n_kv[0], it.n_kv[1]);
    return true;
}

// Producer thread: item k (or the end marker) into slot k % kItemSlots, once the slot's previous
// item has been read by every consumer warp.
__device__ __forceinline__ void publish_item(int k, bool more, const Item& it, Barriers* bar,
                                             int32_t* items) {
    const int slot = k % kItemSlots;
    if (k >= kItemSlots) wait_bar(&bar->item_empty[slot], ((k / kItemSlots) - 1) & 1u);
    int4* dst = reinterpret_cast<int4*>(items + slot * kItemWords);
    dst[0] = make_int4(it.seq, it.kvh, it.q_begin, it.Lq);
    dst[1] = make_int4(it.Lk, it.k_begin, it.offset, it.j0);
    dst[2] = make_int4(it.chunk, it.q_c1, it.row0[0], it.row0[1]);
    dst[3] = make_int4(it.n_kv[0], it.n_kv[1], it.q_c2[0], it.q_c2[1]);
    dst[4] = make_int4(it.n_max, more ? 1 : 0, 0, 0);
    mbar_arrive(&bar->item_full[slot]);  // release: the stores above are visible to the waiters
}
// Consumer warps (all lanes call): item k of this CTA, false at the end marker.
__device__ __forceinline__ bool consume_item(int k, Barriers* bar, const int32_t* items, Item& it) {
    const int slot = k % kItemSlots;
    wait_bar(&bar->item_full[slot], (k / kItemSlots) & 1u);
    const int4* src = reinterpret_cast<const int4*>(items + slot * kItemWords);
    const int4 a = src[0], b = src[1], c = src[2], d = src[3], e = src[4];
    it.seq = a.x; it.kvh = a.y; it.q_begin = a.z; it.Lq = a.w;
    it.Lk = b.x; it.k_begin = b.y; it.offset = b.z; it.j0 = b.w;
    it.chunk = c.x; it.q_c1 = c.y; it.row0[0] = c.z; it.row0[1] = c.w;
    it.n_kv[0] = d.x; it.n_kv[1] = d.y; it.q_c2[0] = d.z; it.q_c2[1] = d.w;
    it.n_max = e.x;
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(&bar->item_empty[slot]);
    return e.y != 0;
}

// Prologue of the persistent prefill: pair_start[0..S] = exclusive scan of the 256-row pairs per
// sequence; keys[] = item indices ordered by descending cost (KV tiles), ties by index — a
// bitonic sort of unique keys, so every CTA ends with the identical list. All threads call it.
__device__ void build_schedule(const PrefillParams& p, Barriers* bar, int32_t* pair_start,
                               uint32_t* keys) {
    const int tid = threadIdx.x, lane = tid & 31;
    const int S = p.num_seqs;
    const int toks_per_pair = 2 * kTM / p.group;
    int32_t* cu_q = pair_start - 2 * kSchedStride;
    int32_t* cu_k = pair_start - kSchedStride;
    for (int s = tid; s <= S; s += kTcThreads) {
        cu_q[s] = __ldg(p.cu_seqlens_q + s);
        cu_k[s] = __ldg(p.cu_seqlens_k + s);
    }
    __syncthreads();
    if (tid < 32) {
        int running = 0;
        for (int base = 0; base < S; base += 32) {
            const int s = base + lane;
            int n = 0;
            if (s < S) {
                const int lq = cu_q[s + 1] - cu_q[s];
                n = prefill_pairs(lq, p.group);
            }
            int incl = n;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int up = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += up;
            }
            if (s < S) pair_start[s] = running + incl - n;
            running += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) {
            pair_start[S] = running;
            bar->num_pairs = running;
            bar->use_keys = (S > 1 && running <= kTcMaxItems) ? 1 : 0;
        }
    }
    __syncthreads();
    const int T = bar->num_pairs;
    if (!bar->use_keys) return;  // one sequence (cost is monotone in the pair index) or too many
    int n2 = 1;
    while (n2 < T) n2 <<= 1;
    for (int t = tid; t < max(n2, 4); t += kTcThreads) {  // (>= 4: the rank sort reads uint4)
        uint32_t key = 0xffffffffu;
        if (t < T) {
            int lo = 0, hi = S;
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (pair_start[mid] <= t) lo = mid; else hi = mid;
            }
            const int lq = cu_q[lo + 1] - cu_q[lo];
            const int lk = cu_k[lo + 1] - cu_k[lo];
            key = prefill_sched_key(prefill_pair_cost(t - pair_start[lo], lq, lk, p.group), t);
        }
        keys[t] = key;
    }
    __syncthreads();
    if (T <= 2 * kTcThreads) {
        // Few items (every BASELINE prefill step): rank sort, no barrier rounds. A key's position
        // is the number of smaller keys (keys are unique); all threads read the same address in
        // each iteration (shared-memory broadcast).
        uint32_t mine[2] = {0xffffffffu, 0xffffffffu};
        int rank[2] = {0, 0};
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (tid + h * kTcThreads < T) mine[h] = keys[tid + h * kTcThreads];
        if (tid < T) {  // (keys[T .. n2) hold the 0xffffffff padding: never smaller)
            const uint4* k4 = reinterpret_cast<const uint4*>(keys);
            for (int u = 0; u < n2 / 4 + (n2 < 4); ++u) {
                const uint4 o = k4[u];
#pragma unroll
                for (int h = 0; h < 2; ++h)
                    rank[h] += (o.x < mine[h]) + (o.y < mine[h]) + (o.z < mine[h]) + (o.w < mine[h]);
            }
        }
        __syncthreads();
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (tid + h * kTcThreads < T) keys[rank[h]] = mine[h];
        __syncthreads();
        return;
    }
    for (int k = 2; k <= n2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int idx = tid; idx < n2; idx += kTcThreads) {
                const int ixj = idx ^ j;
                if (ixj > idx) {
                    const uint32_t a = keys[idx], b = keys[ixj];
                    if ((a > b) == ((idx & k) == 0)) {
                        keys[idx] = b;
                        keys[ixj] = a;
                    }
                }
            }
            __syncthreads();
        }
    }
}

// kCascade = false: varlen causal prefill, persistent over (pair of 128-row tiles, KV head,
//   sequence) items; a tile row is (query token r / G, group head r % G).
// kCascade = true:  shared-prefix pass of the cascade decode (cascade.cu). CTA = (pair of 128-row
//   tiles of the B*G query rows of one KV head, prefix chunk, KV head), one item per CTA. A tile
//   row r is (batch row r / G, group head r % G): the TMA box 64 dims x G heads x 128/G batch
//   rows of q [B, Hq, 128] lands in shared memory in exactly that order. Keys are the tiles of
//   this chunk of the P shared pages (block ids from row b0's table), all visible: no masking. The
//   epilogue writes normalised fp32 partials + log2-sum-exp for the decode merge.
// kEmu of every 4 exponential pairs are evaluated on the FMA pipe (tc::exp2_fma_pipe), the rest on
// the MUFU.
// kTrace: debug build (b200attn_debug_prefill_trace) — the first kTraceCtas CTAs stamp clock64 at
// the phase boundaries of their first kTraceSteps tile steps: stream 0/1 = softmax group 0/1
// {S ready, S row loaded, P computed, P signalled}, stream 2 = issuer {tiles ready, P_0 ready,
// PV_0 issued, S_0 of the next tile issued}; per ITEM: stream 3 = issuer {item fetched, Q landed,
// first S issued, last step issued}, stream 4 = producer {item fetched, first K/V issued, Q buffer
// free, all tiles issued}, stream 5 = softmax group 0 {item fetched, last P signalled, O complete,
// rows stored}; layout [cta][stream][step][4] uint64 (tools/trace_prefill.py).
// kF16: fp16 storage / MMA operands (dtype code 1 of the C-ABI) instead of bf16.
template <bool kCascade, int kEmu, bool kTrace = false, bool kF16 = false>
__global__ void __launch_bounds__(kTcThreads, 1)
prefill_tc_kernel(const __grid_constant__ CUtensorMap q_map,
                  const __grid_constant__ CUtensorMap k_map,
                  const __grid_constant__ CUtensorMap v_map, const PrefillParams p) {
    if constexpr (kCascade) {
        // Let the dependent grid (the streaming decode over the per-row remainders, launched
        // with programmatic stream serialization) start now; it waits on this grid only where it
        // reads the partials written below.
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        const PrefixChunks pc = prefix_chunks_of(p.plan[0], p.block_size, p.prefix_chunks);
        if (static_cast<int>(blockIdx.y) >= pc.valid) return;  // uniform (also: no shared prefix)
    }
    const int g_shift = __ffs(p.group) - 1;  // G is a power of two (128 % G == 0)
    const long long t_entry = kTrace ? clock64() : 0;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    Barriers* bar = reinterpret_cast<Barriers*>(smem + kSmemBar);
    const uint8_t* clc = smem + kSmemClc;
    int32_t* items = reinterpret_cast<int32_t*>(smem + kSmemItems);
    const int32_t* pair_start =
        reinterpret_cast<const int32_t*>(smem + kSmemPairStart) + 2 * kSchedStride;
    const uint32_t* keys = reinterpret_cast<const uint32_t*>(smem + kSmemKeys);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        mbar_init(&bar->q_full, 1);
        mbar_init(&bar->q_empty, 1);
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&bar->k_full[s], 1);
            mbar_init(&bar->k_empty[s], 1);
            mbar_init(&bar->v_full[s], 1);
            mbar_init(&bar->v_empty[s], 1);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            mbar_init(&bar->s_full[i], 1);
            mbar_init(&bar->p_full[i], kTM);
            mbar_init(&bar->o_full[i], 1);
        }
#pragma unroll
        for (int t = 0; t < kClcSlots; ++t) {
            mbar_init(&bar->sched_full[t], 1);
        }
#pragma unroll
        for (int t = 0; t < kItemSlots; ++t) {
            mbar_init(&bar->item_full[t], 1);
            mbar_init(&bar->item_empty[t], kItemReaders);
        }
        fence_mbar_init();
    }
    if constexpr (!kCascade)
        build_schedule(p, bar, reinterpret_cast<int32_t*>(smem + kSmemPairStart) + 2 * kSchedStride,
                       reinterpret_cast<uint32_t*>(smem + kSmemKeys));
    if (warp == kMmaWarp) tc::tmem_alloc(&bar->tmem_base, kTmemCols);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = bar->tmem_base;
    unsigned long long* trace = nullptr;  // this CTA's stamps (kTrace only)
    if constexpr (kTrace) {
        if (p.trace != nullptr && blockIdx.x < kTraceCtas)
            trace = p.trace + static_cast<size_t>(blockIdx.x) * kTraceStreams * kTraceSteps * 4;
    }
    uint32_t tr_step = 0;  // tile steps this role has stamped
    if constexpr (kTrace) {  // last slot of the producer's item stream: {CTA entry, roles start}
        if (trace != nullptr && threadIdx.x == 0) {
            trace[(4 * kTraceSteps + kTraceSteps - 1) * 4 + 0] = t_entry;
            trace[(4 * kTraceSteps + kTraceSteps - 1) * 4 + 1] = clock64();
        }
    }

    // setmaxnreg sits inside each role branch so ptxas budgets every role separately.
    if (warp == kProducerWarp) {
        // ============================ TMA producer ============================
        tc::reg_dec<kRegsControl>();
        if (lane == 0) {
            tma_prefetch_desc(&q_map);
            tma_prefetch_desc(&k_map);
            tma_prefetch_desc(&v_map);
            const bool paged = kCascade || p.block_tables != nullptr;
            uint32_t kv_it = 0;   // K/V tiles issued so far (ring position, across items)
            uint32_t q_it = 0;    // items with work so far (Q buffer phase)
            Item it;
            for (int k = 0;; ++k) {
                const bool more = fetch_item<kCascade>(k, p, bar, clc, pair_start, keys, it);
                publish_item(k, more, it, bar, items);
                if (!more) break;
                if (!kCascade && p.sched_mode == kSchedClc) {
                    // ticket k was good: ask for ticket k + 1 now, it lands while item k runs.
                    // (Never after a failed request: the loop ends on the first one. This thread
                    // is the only reader of the responses, so the slot is free.)
                    const int slot = k % kClcSlots;
                    mbar_expect_tx(&bar->sched_full[slot], 16);
                    clc_try_cancel(smem + kSmemClc + slot * 16, &bar->sched_full[slot]);
                }
                if (it.n_max == 0) continue;
                unsigned long long* tri = nullptr;
                if constexpr (kTrace) {
                    if (trace != nullptr && q_it < kTraceSteps)
                        tri = trace + (4 * kTraceSteps + q_it) * 4;
                    if (tri) tri[0] = clock64();
                }
                const int32_t* table =
                    paged ? p.block_tables + static_cast<int64_t>(it.seq) * p.table_row_stride
                          : nullptr;
                for (int j = 0; j < it.n_max; ++j, ++kv_it) {
                    const int s = kv_it % kStages;
                    const uint32_t ph = (kv_it / kStages) & 1u;
                    const int tok0 = (it.j0 + j) * kTN;
                    int slot;
                    if (paged) {
                        // a 128-token tile never straddles a page (block_size % 128 == 0)
                        slot = __ldg(table + tok0 / p.block_size) * p.block_size +
                               tok0 % p.block_size;
                    } else {
                        slot = it.k_begin + tok0;
                    }
                    wait_bar(&bar->k_empty[s], ph ^ 1u);
                    mbar_expect_tx(&bar->k_full[s], kTileBytes);
                    uint8_t* kd = smem + kSmemK + s * kTileBytes;
                    tma_load_3d(kd, &k_map, &bar->k_full[s], 0, it.kvh, slot);
                    tma_load_3d(kd + kHalfBytes, &k_map, &bar->k_full[s], 64, it.kvh, slot);
                    wait_bar(&bar->v_empty[s], ph ^ 1u);
                    mbar_expect_tx(&bar->v_full[s], kTileBytes);
                    uint8_t* vd = smem + kSmemV + s * kTileBytes;
                    tma_load_3d(vd, &v_map, &bar->v_full[s], 0, it.kvh, slot);
                    tma_load_3d(vd + kHalfBytes, &v_map, &bar->v_full[s], 64, it.kvh, slot);
                    if (kTrace && tri && j == 0) tri[1] = clock64();
                    if (j == 0) {
                        // Q after the item's first K/V tile: those ring slots free up earlier
                        // than the Q buffer (released by the previous item's LAST S MMA).
                        wait_bar(&bar->q_empty, (q_it & 1u) ^ 1u);
                        if (kTrace && tri) tri[2] = clock64();
                        const int n_q = it.n_kv[1] > 0 ? 2 : 1;  // sub-tile 0 always has rows
                        mbar_expect_tx(&bar->q_full, n_q * kTileBytes);
                        for (int i = 0; i < n_q; ++i) {
                            uint8_t* dst = smem + kSmemQ + i * kTileBytes;
                            tma_load_3d(dst, &q_map, &bar->q_full, 0, it.q_c1, it.q_c2[i]);
                            tma_load_3d(dst + kHalfBytes, &q_map, &bar->q_full, 64, it.q_c1,
                                        it.q_c2[i]);
                        }
                        ++q_it;
                    }
                }
                if (kTrace && tri) tri[3] = clock64();
            }
        }
    } else if (warp == kMmaWarp) {
        // ============================ MMA issuer ============================
        tc::reg_dec<kRegsControl>();
        constexpr uint32_t idesc_s = tc::idesc_f32acc(kTM, kTN, false, kF16);
        constexpr uint32_t idesc_o = tc::idesc_f32acc(kTM, kHeadDim, true, kF16);
        // Every tcgen05.mma operand below is a COMPILE-TIME constant, so an MMA is a few uniform
        // moves + UTCHMMA. With run-time descriptors ptxas cannot prove warp-uniformity and emits
        // ~13 dependent vector instructions per MMA (ELECT / R2UR broadcast loop), 75-84 clk for
        // a 64-clk MMA — which made this warp the bottleneck of the kernel
        // (profiles/r02_trace_prefill_*.json). Hence: the ring stage and the accumulate flag are
        // template parameters selected by (warp-uniform) branches; the TMEM base is 0 (this CTA
        // owns all 512 columns of the SM's TMEM); the tiles sit at kSmemWindowBase, where the
        // dynamic shared memory of a kernel without static shared memory starts on sm_100
        // (descriptors only carry address bits 4..17). Both are checked: a different layout
        // traps instead of computing garbage.
        if (tmem != 0 || (smem_u32(smem) & 0x3FFFFu) != kSmemWindowBase) __trap();
        constexpr uint32_t smem_base = kSmemWindowBase;
        auto issue_s = [&](auto i_c, auto stage_c) {  // S_i = Q_i . K^T (head_dim 128 = 2 x 4 x 16)
            constexpr int i = decltype(i_c)::value, stage = decltype(stage_c)::value;
            {
                const uint64_t a0 = tc::desc_kmajor(smem_base + kSmemQ + i * kTileBytes);
                const uint64_t b0 = tc::desc_kmajor(smem_base + kSmemK + stage * kTileBytes);
                // (bytes >> 4) added to the 14-bit start-address field: smem < 256 KiB, no carry
                auto off = [](int kk) -> uint64_t {
                    return static_cast<uint64_t>(((kk >> 2) * kHalfBytes + (kk & 3) * 32) >> 4);
                };
                tc::mma_ss_static<i * kTN, false>(a0 + off(0), b0 + off(0), idesc_s);
#pragma unroll
                for (int kk = 1; kk < 8; ++kk)
                    tc::mma_ss_static<i * kTN, true>(a0 + off(kk), b0 + off(kk), idesc_s);
                if (lane == 0) tc::commit(&bar->s_full[i]);
            }
            __syncwarp();
        };
        auto issue_pv = [&](auto i_c, auto stage_c, auto acc_c) {  // O_i (+)= P_i . V
            constexpr int i = decltype(i_c)::value, stage = decltype(stage_c)::value;
            constexpr bool acc = decltype(acc_c)::value;
            {
                const uint64_t b0 =
                    tc::desc_mnmajor(smem_base + kSmemV + stage * kTileBytes, kHalfBytes);
                constexpr uint32_t d = 2 * kTN + i * kHeadDim, a = i * kTN;
                constexpr uint64_t step = (16 * 128) >> 4;  // 16 tokens per instruction
                tc::mma_ts_static<d, a + 0, acc>(b0 + 0 * step, idesc_o);
                tc::mma_ts_static<d, a + 8, true>(b0 + 1 * step, idesc_o);
                tc::mma_ts_static<d, a + 16, true>(b0 + 2 * step, idesc_o);
                tc::mma_ts_static<d, a + 24, true>(b0 + 3 * step, idesc_o);
                tc::mma_ts_static<d, a + 32, true>(b0 + 4 * step, idesc_o);
                tc::mma_ts_static<d, a + 40, true>(b0 + 5 * step, idesc_o);
                tc::mma_ts_static<d, a + 48, true>(b0 + 6 * step, idesc_o);
                tc::mma_ts_static<d, a + 56, true>(b0 + 7 * step, idesc_o);
            }
            __syncwarp();
        };
        using std::integral_constant;
        auto issue_s_rt = [&](auto i_c, int stage) {
            if (stage == 0) issue_s(i_c, integral_constant<int, 0>{});
            else issue_s(i_c, integral_constant<int, 1>{});
        };
        auto issue_pv_rt = [&](auto i_c, int stage, bool acc) {
            if (stage == 0) {
                if (acc) issue_pv(i_c, integral_constant<int, 0>{}, integral_constant<bool, true>{});
                else issue_pv(i_c, integral_constant<int, 0>{}, integral_constant<bool, false>{});
            } else {
                if (acc) issue_pv(i_c, integral_constant<int, 1>{}, integral_constant<bool, true>{});
                else issue_pv(i_c, integral_constant<int, 1>{}, integral_constant<bool, false>{});
            }
        };
        static_assert(kStages == 2, "the issue helpers enumerate two ring stages");

        uint32_t kv_it = 0, q_it = 0;
        uint32_t sp_it[2] = {0, 0};  // S/P rounds per sub-tile so far (s_full / p_full phases)
        // Items are picked up one ahead: the next item is read from the ring at the top of the
        // current item's LAST tile step (its loads do not depend on that step, so it has been
        // published or is about to be), and its first S MMAs go out right behind the last PV
        // instead of a ring wait + ~1000 clk of cold code later.
        Item it, nxt;
        bool have = consume_item(0, bar, items, it);
        for (int k = 0; have; ++k) {
            bool got_next = false, more = false;
            auto take_next = [&]() {
                if (!got_next) {
                    more = consume_item(k + 1, bar, items, nxt);
                    got_next = true;
                }
            };
            const int n_max = it.n_max;
            if (n_max == 0) {
                take_next();
                it = nxt;
                have = more;
                continue;
            }
            unsigned long long* tri = nullptr;
            if constexpr (kTrace) {
                if (trace != nullptr && lane == 0 && q_it < kTraceSteps)
                    tri = trace + (3 * kTraceSteps + q_it) * 4;
                if (tri) tri[0] = clock64();
            }
            wait_bar(&bar->q_full, q_it & 1u);
            if (kTrace && tri) tri[1] = clock64();
            ++q_it;
            wait_bar(&bar->k_full[kv_it % kStages], (kv_it / kStages) & 1u);
            tc::fence_after_sync();
            if (it.n_kv[0] > 0) issue_s_rt(integral_constant<int, 0>{}, kv_it % kStages);
            if (it.n_kv[1] > 0) issue_s_rt(integral_constant<int, 1>{}, kv_it % kStages);
            if (lane == 0) {
                tc::commit(&bar->k_empty[kv_it % kStages]);
                if (n_max == 1) tc::commit(&bar->q_empty);  // the item's last S has been issued
            }
            if (kTrace && tri) tri[2] = clock64();
            __syncwarp();
            for (int j = 0; j < n_max; ++j, ++kv_it) {
                const int s = kv_it % kStages;
                unsigned long long* tr = nullptr;
                if constexpr (kTrace) {
                    if (trace != nullptr && lane == 0 && tr_step < kTraceSteps)
                        tr = trace + (2 * kTraceSteps + tr_step) * 4;
                    ++tr_step;
                }
                if (j == n_max - 1) take_next();
                wait_bar(&bar->v_full[s], (kv_it / kStages) & 1u);
                // Rows at or beyond Lk inside the tile are masked (P = 0), but 0 * NaN would still
                // poison O: clear those V rows (unwritten cache slots may hold anything).
                const int valid = kCascade ? kTN : it.Lk - j * kTN;  // shared pages are full
                if (valid < kTN) {
                    uint8_t* vs = smem + kSmemV + s * kTileBytes;
                    const int r_lo = max(valid, 0);
                    for (int idx = r_lo * 16 + lane; idx < kTN * 16; idx += 32) {
                        const int r = idx >> 4, c = idx & 15;
                        *reinterpret_cast<int4*>(vs + (c >> 3) * kHalfBytes + r * 128 +
                                                 (c & 7) * 16) = make_int4(0, 0, 0, 0);
                    }
                    fence_proxy_async();
                    __syncwarp();
                }
                const int sn = (kv_it + 1) % kStages;
                if (j + 1 < n_max) wait_bar(&bar->k_full[sn], ((kv_it + 1) / kStages) & 1u);
                if (kTrace && tr) tr[0] = clock64();
                auto sub_tile = [&](auto i_c) {
                    constexpr int i = decltype(i_c)::value;
                    if (j < it.n_kv[i]) {
                        wait_bar(&bar->p_full[i], sp_it[i] & 1u);
                        if (kTrace && tr && i == 0) tr[1] = clock64();
                        ++sp_it[i];
                        tc::fence_after_sync();
                        issue_pv_rt(i_c, s, j > 0);
                        if (kTrace && tr && i == 0) tr[2] = clock64();
                        if (j == it.n_kv[i] - 1) {
                            if (lane == 0) tc::commit(&bar->o_full[i]);
                            __syncwarp();
                        }
                    }
                    if (j + 1 < it.n_kv[i]) issue_s_rt(i_c, sn);
                    if (kTrace && tr && i == 0) tr[3] = clock64();
                };
                sub_tile(integral_constant<int, 0>{});
                sub_tile(integral_constant<int, 1>{});
                if (lane == 0) {
                    tc::commit(&bar->v_empty[s]);
                    if (j + 1 < n_max) tc::commit(&bar->k_empty[sn]);
                    if (j + 2 == n_max) tc::commit(&bar->q_empty);  // last S issued in this step
                }
                __syncwarp();
            }
            if (kTrace && tri) tri[3] = clock64();
            take_next();
            it = nxt;
            have = more;
        }
    } else if (warp < kProducerWarp) {
        // ============================ softmax / correction / epilogue ============================
        tc::reg_inc<kRegsSoftmax>();
        const int i = warp >> 2;                       // sub-tile
        const int r = (warp & 3) * 32 + lane;          // row inside the sub-tile = TMEM lane
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const uint32_t s_tmem = tmem + lane_base + i * kTN;
        const uint32_t o_tmem = tmem + lane_base + 2 * kTN + i * kHeadDim;
        const float c = p.scale_log2;
        uint32_t sp_it = 0;   // S/P rounds of this sub-tile so far
        uint32_t o_it = 0;    // items of this sub-tile with work so far
        // The next item is read from the ring between the last P signal and the epilogue, i.e.
        // while this item's last PV MMAs run.
        Item it, nxt;
        bool have = consume_item(0, bar, items, it);
        for (int k = 0; have; ++k) {
        // (selects, not it.row0[i]: a runtime index would push the item into local memory)
        const int row = (i ? it.row0[1] : it.row0[0]) + (r >> g_shift);  // token (cascade: batch row)
        const int head = it.kvh * p.group + (r & (p.group - 1));
        const int n = i ? it.n_kv[1] : it.n_kv[0];
        const int Lk = it.Lk, offset = it.offset;
        float m_ref = -INFINITY;  // scaled (log2-domain) reference maximum of this row
        float l = 0.f;
        unsigned long long* tri = nullptr;
        if constexpr (kTrace) {
            if (trace != nullptr && i == 0 && r == 0 && n > 0 && o_it < kTraceSteps)
                tri = trace + (5 * kTraceSteps + o_it) * 4;
            if (tri) tri[0] = clock64();
        }

        for (int j = 0; j < n; ++j, ++sp_it) {
            unsigned long long* tr = nullptr;
            if constexpr (kTrace) {
                if (trace != nullptr && r == 0 && tr_step < kTraceSteps)
                    tr = trace + (i * kTraceSteps + tr_step) * 4;
                ++tr_step;
            }
            wait_bar(&bar->s_full[i], sp_it & 1u);
            tc::fence_after_sync();
            if (kTrace && tr) tr[0] = clock64();
            uint32_t sr[128];
            tc::tmem_ld128(s_tmem, sr);
            if (kTrace && tr) tr[1] = clock64();

            const int key0 = j * kTN;
            // columns <= lim are visible (cascade: every key of the shared prefix is)
            const int lim = kCascade ? kTN - 1 : min(row + offset, Lk - 1) - key0;
            if (lim < kTN - 1) {
#pragma unroll
                for (int t = 0; t < 128; ++t)
                    sr[t] = t <= lim ? sr[t] : 0xff800000u;  // -inf
            }
            float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
            for (int t = 0; t < 128; t += 8) {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    mx4[u] = tc::fmax3(mx4[u], __uint_as_float(sr[t + 2 * u]),
                                       __uint_as_float(sr[t + 2 * u + 1]));
            }
            const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
            const float m_new = fmaxf(m_ref, mx * c);
            if (j == 0) {
                m_ref = m_new;
            } else if (const bool grow = m_new - m_ref > kRescaleThreshold;
                       __any_sync(0xffffffffu, grow)) {
                // PV_i[j-1] has completed (it was issued before S_i[j], whose commit we waited on).
                // The branch is warp-uniform (TMEM access is .sync.aligned) but the decision is
                // per row: rows that did not grow are multiplied by exactly 1, so a row's result
                // never depends on its neighbours.
                const float alpha = grow ? fast_exp2(m_ref - m_new) : 1.f;
                l *= alpha;
                m_ref = grow ? m_new : m_ref;
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) {
                    uint32_t orow[32];
                    tc::tmem_ld32(o_tmem + q4 * 32, orow);
#pragma unroll
                    for (int t = 0; t < 32; ++t)
                        orow[t] = __float_as_uint(__uint_as_float(orow[t]) * alpha);
                    tc::tmem_st32(o_tmem + q4 * 32, orow);
                }
            }
            const float mu = m_ref == -INFINITY ? 0.f : m_ref;
            // p = exp2(s*c - mu): packed f32x2 FMA / ADD halve the issue slots next to MUFU.ex2
            const float2 c2 = make_float2(c, c), nmu2 = make_float2(-mu, -mu);
            float2 ls[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {  // 32 packed columns at a time: P regs die early
                uint32_t pr[32];
#pragma unroll
                for (int t = 0; t < 32; ++t) {
                    const int e = hh * 64 + 2 * t;
                    float2 x = make_float2(__uint_as_float(sr[e]), __uint_as_float(sr[e + 1]));
                    x = __ffma2_rn(x, c2, nmu2);
                    if ((t & 3) < kEmu) {
                        x = tc::exp2_fma_pipe(x);
                    } else {
                        x.x = fast_exp2(x.x);
                        x.y = fast_exp2(x.y);
                    }
                    ls[t & 1] = __fadd2_rn(ls[t & 1], x);
                    pr[t] = pack_16x2<kF16>(x.x, x.y);
                }
                tc::tmem_st32(s_tmem + hh * 32, pr);
            }
            l += (ls[0].x + ls[0].y) + (ls[1].x + ls[1].y);
            tc::wait_st();
            if (kTrace && tr) tr[2] = clock64();
            tc::fence_before_sync();
            mbar_arrive(&bar->p_full[i]);
            if (kTrace && tr) tr[3] = clock64();
        }

        const bool more = consume_item(k + 1, bar, items, nxt);
        if constexpr (kCascade) {
            // ---- epilogue: normalised fp32 partial + log2-sum-exp of this prefix chunk ----
            const int b = row;
            if (n > 0) {  // warp-uniform: n depends on the sub-tile only
                wait_bar(&bar->o_full[i], o_it & 1u);
                ++o_it;
                tc::fence_after_sync();
                const int64_t qrow =
                    static_cast<int64_t>(b) * p.num_q_heads + head;
                const int64_t slot = qrow * p.ws_slots + p.ws_slot0 + it.chunk;
                const float inv = l > 0.f ? 1.f / l : 0.f;
                float* og = p.ws_o + slot * kHeadDim;
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) {
                    uint32_t orow[32];
                    tc::tmem_ld32(o_tmem + q4 * 32, orow);
                    if (b < p.batch) {
#pragma unroll
                        for (int v4 = 0; v4 < 8; ++v4) {
                            float4 f;
                            f.x = __uint_as_float(orow[v4 * 4 + 0]) * inv;
                            f.y = __uint_as_float(orow[v4 * 4 + 1]) * inv;
                            f.z = __uint_as_float(orow[v4 * 4 + 2]) * inv;
                            f.w = __uint_as_float(orow[v4 * 4 + 3]) * inv;
                            *reinterpret_cast<float4*>(og + q4 * 32 + v4 * 4) = f;
                        }
                    }
                }
                if (b < p.batch) p.ws_lse[slot] = l > 0.f ? m_ref + log2f(l) : -INFINITY;
            }
        } else {
        // ---- epilogue: O / l -> bf16 -> global (row-predicated 16-byte stores) ----
        if (kTrace && tri) tri[1] = clock64();
        const bool row_ok = row < it.Lq;  // (surplus ticket: Lq = 0)
        uint8_t* orow_g = static_cast<uint8_t*>(p.out) +
                          ((static_cast<int64_t>(it.q_begin) + row) * p.num_q_heads + head) * 256;
        if (n > 0) {
            wait_bar(&bar->o_full[i], o_it & 1u);
            ++o_it;
            tc::fence_after_sync();
            if (kTrace && tri) tri[2] = clock64();
            const float inv = l > 0.f ? 1.f / l : 0.f;
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
                uint32_t orow[32];
                tc::tmem_ld32(o_tmem + q4 * 32, orow);
                if (row_ok) {
#pragma unroll
                    for (int v8 = 0; v8 < 2; ++v8) {  // 32-byte stores: whole sectors, half the
                        uint32_t pk[8];               // LSU work of 16-byte ones
#pragma unroll
                        for (int w = 0; w < 8; ++w)
                            pk[w] = pack_16x2<kF16>(__uint_as_float(orow[v8 * 16 + 2 * w]) * inv,
                                                __uint_as_float(orow[v8 * 16 + 2 * w + 1]) * inv);
                        st_v8(orow_g + q4 * 64 + v8 * 32, pk);
                    }
                }
            }
            // the next item's first PV overwrites O_i: order these TMEM reads before the P
            // signal that releases it
            tc::fence_before_sync();
            if (kTrace && tri) tri[3] = clock64();
        } else if (row_ok) {
            // no visible key for this sub-tile (only possible when Lk < Lq): zeros, as flash-attn
#pragma unroll
            for (int v = 0; v < 16; ++v) st_v4(orow_g + v * 16, make_int4(0, 0, 0, 0));
        }
        }  // !kCascade
        it = nxt;
        have = more;
        }  // items
    }

    else {
        tc::reg_dec<kRegsControl>();  // idle warps of the control warpgroup
    }

    tc::fence_before_sync();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem, kTmemCols);
    }
}

}  // namespace

namespace {

// B200ATTN_EXP_EMU=0..3 overrides how many of every 4 exponential pairs leave the MUFU.
int exp_emu_from_env(int dflt) {
    static int v = [] {
        const char* e = std::getenv("B200ATTN_EXP_EMU");
        return (e && e[0] >= '0' && e[0] <= '3') ? e[0] - '0' : -1;
    }();
    return v < 0 ? dflt : v;
}

constexpr int kMaxDevices = 64;
constexpr int kDefaultExpEmu = 0;

template <bool kCascade, int kEmu, bool kTrace = false, bool kF16 = false>
cudaError_t launch_tc(dim3 grid, const PrefillParams& p, const CUtensorMap& q_map,
                      const CUtensorMap& k_map, const CUtensorMap& v_map, cudaStream_t stream) {
    // the opt-in shared-memory size is a per-device function attribute
    static bool configured[kMaxDevices] = {};
    int dev = 0;
    if (cudaError_t e = cudaGetDevice(&dev); e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices || !configured[dev]) {
        cudaError_t e = cudaFuncSetAttribute(prefill_tc_kernel<kCascade, kEmu, kTrace, kF16>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kTcSmemBytes);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < kMaxDevices) configured[dev] = true;
    }
    prefill_tc_kernel<kCascade, kEmu, kTrace, kF16><<<grid, kTcThreads, kTcSmemBytes, stream>>>(
        q_map, k_map, v_map, p);
    count_launch();
    return cudaGetLastError();
}

template <bool kCascade>
cudaError_t launch_tc_emu(int emu, dim3 grid, const PrefillParams& p, const CUtensorMap& q_map,
                          const CUtensorMap& k_map, const CUtensorMap& v_map,
                          cudaStream_t stream) {
    if (p.f16)  // fp16: the default exponentials only
        return launch_tc<kCascade, kDefaultExpEmu, false, true>(grid, p, q_map, k_map, v_map,
                                                                stream);
    switch (emu) {
        case 0: return launch_tc<kCascade, 0>(grid, p, q_map, k_map, v_map, stream);
        case 1: return launch_tc<kCascade, 1>(grid, p, q_map, k_map, v_map, stream);
        case 3: return launch_tc<kCascade, 3>(grid, p, q_map, k_map, v_map, stream);
        default: return launch_tc<kCascade, 2>(grid, p, q_map, k_map, v_map, stream);
    }
}

int device_sm_count() {
    const int n = current_device_sm_count();
    return n > 0 ? n : 148;
}

}  // namespace

// B200ATTN_PREFILL_SCHED=clc|static|grid (default clc), see the header comment.
int sched_mode_from_env() {
    static int v = [] {
        const char* e = std::getenv("B200ATTN_PREFILL_SCHED");
        if (e && e[0] == 's') return static_cast<int>(kSchedStatic);
        if (e && e[0] == 'g') return static_cast<int>(kSchedGrid);
        return static_cast<int>(kSchedClc);
    }();
    return v;
}

cudaError_t launch_prefill_tc(const PrefillParams& p_in, const CUtensorMap& q_map,
                              const CUtensorMap& k_map, const CUtensorMap& v_map,
                              cudaStream_t stream) {
    // Items = 256-row pairs over all sequences (<= total_q / tokens-per-pair + num_seqs) x KV
    // heads; the exact count is device data. clc / grid: one CTA per possible item (surplus
    // tickets are empty); static: one CTA per SM.
    PrefillParams p = p_in;
    p.sched_mode = sched_mode_from_env();
    const int toks_per_pair = 2 * kTM / p.group;
    const int64_t bound =
        (static_cast<int64_t>(p.total_q) / toks_per_pair + p.num_seqs) * p.num_kv_heads;
    int64_t ctas = bound;
    if (p.sched_mode == kSchedStatic && ctas > device_sm_count()) ctas = device_sm_count();
    dim3 grid(static_cast<unsigned>(ctas > 0 ? ctas : 1), 1, 1);
    if (p.trace != nullptr)  // debug build, default exponentials only
        return launch_tc<false, kDefaultExpEmu, true>(grid, p, q_map, k_map, v_map, stream);
    return launch_tc_emu<false>(exp_emu_from_env(kDefaultExpEmu), grid, p, q_map, k_map, v_map,
                                stream);
}

// The kernel's schedule on the host (b200attn_prefill_schedule_host): the same functions of
// prefill_plan.cuh on host copies of cu_seqlens; std::sort of the unique keys = the device's rank /
// bitonic sort. items: 8 int32 per item {seq, pair, row0_0, row0_1, n_kv_0, n_kv_1, cost, sorted}
// in the order the CTAs take them (per KV head; the kernel's global index g walks item g / Hkv
// for KV head g % Hkv). Returns the item count per KV head (also when it exceeds max_items).
int prefill_schedule_host(const int32_t* cu_q, const int32_t* cu_k, int32_t num_seqs,
                          int32_t group, int32_t* items, int32_t max_items) {
    int T = 0;
    for (int s = 0; s < num_seqs; ++s) T += prefill_pairs(cu_q[s + 1] - cu_q[s], group);
    const bool sorted = num_seqs > 1 && T <= kTcMaxItems;
    uint32_t* keys = new uint32_t[T > 0 ? T : 1];
    int32_t* seq_of = new int32_t[T > 0 ? T : 1];
    int32_t* pair_of = new int32_t[T > 0 ? T : 1];
    int t = 0;
    for (int s = 0; s < num_seqs; ++s) {
        const int lq = cu_q[s + 1] - cu_q[s], lk = cu_k[s + 1] - cu_k[s];
        for (int pr = 0; pr < prefill_pairs(lq, group); ++pr, ++t) {
            seq_of[t] = s;
            pair_of[t] = pr;
            keys[t] = prefill_sched_key(prefill_pair_cost(pr, lq, lk, group), t);
        }
    }
    if (sorted) std::sort(keys, keys + T);
    for (int li = 0; li < T && li < max_items; ++li) {
        const int item = sorted ? prefill_key_item(keys[li]) : T - 1 - li;
        const int s = seq_of[item];
        const int lq = cu_q[s + 1] - cu_q[s], lk = cu_k[s + 1] - cu_k[s];
        const PrefillSubTiles st = prefill_sub_tiles(pair_of[item], lq, lk, group);
        int32_t* o = items + static_cast<int64_t>(li) * 8;
        o[0] = s;
        o[1] = pair_of[item];
        o[2] = st.row0[0];
        o[3] = st.row0[1];
        o[4] = st.n_kv[0];
        o[5] = st.n_kv[1];
        o[6] = prefill_pair_cost(pair_of[item], lq, lk, group);
        o[7] = sorted ? 1 : 0;
    }
    delete[] keys;
    delete[] seq_of;
    delete[] pair_of;
    return T;
}

cudaError_t launch_cascade_prefix(const PrefillParams& p, const CUtensorMap& q_map,
                                  const CUtensorMap& k_map, const CUtensorMap& v_map,
                                  cudaStream_t stream) {
    const int rows = p.batch * p.group;  // query rows per KV head
    dim3 grid((rows + 2 * kTM - 1) / (2 * kTM), p.prefix_chunks, p.num_kv_heads);
    return launch_tc_emu<true>(exp_emu_from_env(kDefaultExpEmu), grid, p, q_map, k_map, v_map,
                               stream);
}

}  // namespace b200
