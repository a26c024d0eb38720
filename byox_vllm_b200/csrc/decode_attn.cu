// Paged decode attention for sm_100a: a persistent, work-balanced stream over the KV tiles of the
// whole batch, TMA-staged K/V tiles, GQA packing. Replaces flash_attn_with_kvcache on the decode
// branch of Attention.forward (reference nanovllm/layers/attention.py:87-96).
//
// Scheduling (the reference's kernel, flash-attn's split-KV, picks one split count on the host
// from sizes only and is blind to the ragged context_lens it cannot read):
//   * A "row" is a (sequence b, KV head) pair; its work is the 64-token KV tiles of the sequence
//     (after the shared-prefix pages the cascade already attended) plus `row_cost` virtual units
//     for its fixed overhead (q load, epilogue). Rows are laid end to end in (b, kv head) order
//     into one unit stream of U units.
//   * Every CTA re-derives the exclusive scan of the per-sequence unit counts from context_lens
//     in shared memory (B <= 1024: one block scan, ~1 us), cuts the stream into gridDim.x
//     contiguous ranges of `per = max(ceil(U / gridDim.x), min_units)` units and walks its own
//     range. So the partition is a function of DEVICE data (ragged lengths balance to within one
//     tile) while the grid and the workspace are functions of host-visible sizes only: the launch
//     is CUDA-graph capturable and needs no plan launch of its own.
//   * A range boundary cuts at most one row, so a CTA produces at most two partial pieces: the
//     tail of a row the previous CTA started ("head" slot 2c) and the head of a row the next CTA
//     continues ("tail" slot 2c+1). Partial pieces go to the workspace as normalised fp32 O plus
//     log2-sum-exp; the piece that arrives last at the row's counter (fence + atomic) merges the
//     row in the same launch. Rows that lie inside one range are written straight to `out`.
//   * One CTA = 4 consumer warps + 1 TMA producer warp, 2 CTAs / SM. The producer walks the same
//     piece list and streams 64-token tiles of K and V (4 SWIZZLE_128B boxes = 32 KiB / stage,
//     3 stages, mbarrier expect-tx, L2 evict_first) WITHOUT draining the ring at row boundaries:
//     while the consumers run a row's epilogue the next row's tiles are already landing.
//   * GQA packing: the G = Hq/Hkv query heads that share the KV head are the N dimension of
//     mma.sync.m16n8k16 (N = 8 >= G), tokens are M: S^T[tok, head] = K[tok, :] . Q^T[:, head], so
//     every K/V byte is read from HBM once per KV head. Each consumer warp owns 16 tokens of every
//     tile with its own online-softmax state; P^T goes from accumulator to B-operand layout with
//     movmatrix.trans and O^T += V^T . P^T uses ldmatrix.trans on the V tile. The 4 warps are
//     combined by a two-level tree through 8 KiB of shared memory that is NOT part of the ring.
//
// Roofline: HBM. Algorithmic bytes per (b, kv-head) = ctx * 128 * 2 B * 2 (K and V), read once.
#include "common.h"
#include "decode_plan.cuh"
#include "ptx.cuh"

#include <cstdlib>

namespace b200 {

namespace {

constexpr int kTile = kDecodeTile;           // tokens per stage (64)
constexpr int kConsumerWarps = 4;           // 16 tokens each
constexpr int kDecodeThreads = (kConsumerWarps + 1) * 32;
constexpr int kSubTileBytes = kTile * 128;  // [64 tok][64 dims] bf16 = 8 KiB, 128-byte rows
constexpr int kStageBytes = 4 * kSubTileBytes;  // K lo/hi + V lo/hi = 32 KiB
constexpr int kScratchFloats = 2 * 8 * kHeadDim;  // two partial O^T sets of [8 heads][128 dims]
constexpr int kPieceQueue = 4;  // piece descriptors in flight from the producer warp to the consumers

// Programmatic dependent launch: block until the primary grid (the cascade prefix pass) has
// completed and its global writes are visible. A no-op when launched without the attribute.
__device__ __forceinline__ void grid_dependency_wait() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ void consumer_bar_sync() {
    asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory");
}

// Scratch index of (set, head h, dim d): rows of 128 floats, XOR-swizzled on bits 3-4 by the head
// pair so that the fragment-layout accesses (4 head pairs x 8 consecutive dims per instruction)
// hit 32 distinct banks; the swizzle keeps 16-byte groups intact for the float4 reads.
__device__ __forceinline__ int scratch_idx(int set, int h, int d) {
    return (set * 8 + h) * kHeadDim + (d ^ ((h >> 1) << 3));
}

// ---- merging partial pieces ----------------------------------------------------------------------
// Sources of one (b, q-head) row: `n_piece` pieces written by the CTAs of consecutive groups
// g_first .. on this KV head (slot pair 2 * (g * Hkv + kvh): the first piece in the "continued by
// the next group" slot +1, the others in the "continues an earlier group's row" slot +0) followed
// by `n_prefix` shared-prefix chunks of the cascade. All are normalised fp32 O plus a log2 lse.
struct MergeSrc {
    const float* piece_o;    // [2 * groups * Hkv][8][128]
    const float* piece_lse;  // [2 * groups * Hkv][8]
    const float* prefix_o;   // [B * Hq][chunks][128]
    const float* prefix_lse; // [B * Hq][chunks]
    int g_first, kvh, num_kv_heads, n_piece, h_local;
    int rows_base;           // rows engine: slot of the row's split 0 (>= 0); stream engine: -1
    int64_t row;             // b * Hq + q head
    int prefix_chunks, n_prefix;
    __device__ __forceinline__ int piece_slot(int i) const {
        if (rows_base >= 0) return rows_base + i;
        return 2 * ((g_first + i) * num_kv_heads + kvh) + (i == 0 ? 1 : 0);
    }
    __device__ __forceinline__ const float* lse_ptr(int i) const {
        return i < n_piece ? piece_lse + piece_slot(i) * 8 + h_local
                           : prefix_lse + row * prefix_chunks + (i - n_piece);
    }
    __device__ __forceinline__ const float4* o_ptr(int i) const {
        return reinterpret_cast<const float4*>(
            i < n_piece ? piece_o + (static_cast<int64_t>(piece_slot(i)) * 8 + h_local) * kHeadDim
                        : prefix_o + (row * prefix_chunks + (i - n_piece)) * kHeadDim);
    }
};

// One warp combines the sources of a row into out[row]: lane l holds the weight of source l of
// each group of 32; the O loads of 8 sources are issued together so their latencies overlap.
// Loads go through L2 (ld.global.cg): the partials were written by other SMs a moment ago.
template <bool kF16>
__device__ __forceinline__ void merge_row(const MergeSrc& src, __nv_bfloat16* __restrict__ out,
                                          int lane) {
    const int n = src.n_piece + src.n_prefix;
    float mx = -INFINITY;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const float v = i0 + lane < n ? __ldcg(src.lse_ptr(i0 + lane)) : -INFINITY;
        mx = fmaxf(mx, v);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    float wsum = 0.f;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i0 = 0; i0 < n; i0 += 32) {
        const float v = i0 + lane < n ? __ldcg(src.lse_ptr(i0 + lane)) : -INFINITY;
        const float my_w = v == -INFINITY ? 0.f : fast_exp2(v - mx);  // empty piece: weight 0
        wsum += my_w;
        const int cnt = min(32, n - i0);
        for (int s0 = 0; s0 < cnt; s0 += 8) {
            float4 x[8];
            float w[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float wl = __shfl_sync(0xffffffffu, my_w, (s0 + u) & 31);
                w[u] = s0 + u < cnt ? wl : 0.f;
                x[u] = w[u] != 0.f ? __ldcg(src.o_ptr(i0 + s0 + u) + lane)
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                acc.x += w[u] * x[u].x;
                acc.y += w[u] * x[u].y;
                acc.z += w[u] * x[u].z;
                acc.w += w[u] * x[u].w;
            }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, off);
    const float inv = wsum > 0.f ? 1.f / wsum : 0.f;
    uint2 pk;
    pk.x = pack_16x2<kF16>(acc.x * inv, acc.y * inv);
    pk.y = pack_16x2<kF16>(acc.z * inv, acc.w * inv);
    *reinterpret_cast<uint2*>(out + src.row * kHeadDim + lane * 4) = pk;
}

// Every piece of a cut row calls this after its partial is written (the 128 consumer threads, or
// all 160 threads of a rows-engine CTA with an empty split).
// Returns true in the CTA that arrived last - that one merges. The partial stores of all threads
// are ordered before thread 0's arrival by the CTA barrier; the arrival is an acq_rel atomic at gpu
// scope (release is cumulative over what the barrier ordered before it, acquire makes the other
// pieces' partials visible to the threads released by the second barrier). The counter is left at
// zero for the next launch.
__device__ __forceinline__ int atom_add_acq_rel_gpu(int32_t* addr, int v) {
    int old;
    asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], %2;"
                 : "=r"(old)
                 : "l"(addr), "r"(v)
                 : "memory");
    return old;
}
template <int kThreads, int kBarId>
__device__ __forceinline__ bool piece_arrive_is_last(int32_t* counter, int n_pieces,
                                                     int* flag_smem) {
    asm volatile("bar.sync %0, %1;" ::"n"(kBarId), "n"(kThreads) : "memory");
    if (threadIdx.x == 0) {
        const int old = atom_add_acq_rel_gpu(counter, 1);
        const int last = old == n_pieces - 1;
        if (last) *counter = 0;
        *flag_smem = last;
    }
    asm volatile("bar.sync %0, %1;" ::"n"(kBarId), "n"(kThreads) : "memory");
    return *flag_smem != 0;
}

// ---- producer helpers ----------------------------------------------------------------------------
// Fused scatter (reference attention.py:70-71): the piece that ends at the newest token writes this
// step's K/V row of its KV head into the cache slot, then orders the generic-proxy stores before
// its own TMA reads of that slot (fence.proxy.async). No other piece reads the slot: it is token
// ctx-1 of sequence b, inside the last tile of the row. The loads are issued one piece ahead
// (scatter_prefetch) so that their latency hides behind the previous piece's tile stream.
struct ScatterPre {
    int4 row;      // lanes 0-15: 16-byte chunk of the K row, lanes 16-31: of the V row
    int64_t slot;
};
__device__ __forceinline__ ScatterPre scatter_prefetch(const DecodeParams& p, int b, int kvh,
                                                       int lane) {
    ScatterPre s;
    const bool is_v = lane >= 16;
    const int c = lane & 15;  // 16-byte chunk of the 256-byte head row
    const uint8_t* src =
        static_cast<const uint8_t*>(is_v ? p.new_value : p.new_key) +
        static_cast<int64_t>(b) * (is_v ? p.value_row_stride : p.key_row_stride) * 2 + kvh * 256 +
        c * 16;
    s.slot = p.slot_mapping[b];
    s.row = ld_nc_v4(src);
    return s;
}
__device__ __forceinline__ void scatter_finish(const DecodeParams& p, const ScatterPre& s, int b,
                                               int kvh, int lane) {
    const bool ok = s.slot >= 0 && s.slot < p.num_slots;  // warp-uniform
    if (ok) {
        const bool is_v = lane >= 16;
        const int c = lane & 15;
        uint8_t* dst = static_cast<uint8_t*>(is_v ? p.v_cache : p.k_cache) +
                       (s.slot * p.num_kv_heads + kvh) * 256 + c * 16;
        int4 row = s.row;
        if (p.nr.positions != nullptr && !is_v) {
            // fused k_norm + RoPE of the raw K row: lanes 0-15 only (half-warp shuffles)
            const bool has_norm = p.nr.k_weight != nullptr;
            uint4 wv = make_uint4(0, 0, 0, 0);
            if (has_norm)
                wv = *reinterpret_cast<const uint4*>(
                    static_cast<const uint8_t*>(p.nr.k_weight) + c * 16);
            const uint4 rot = norm_rope_row16(
                make_uint4(row.x, row.y, row.z, row.w), wv, has_norm, p.nr.eps,
                load_cos_sin8(cos_sin_row(p.nr, b), c), c, 0x0000ffffu);
            row = make_int4(rot.x, rot.y, rot.z, rot.w);
        }
        st_v4(dst, row);
        asm volatile("fence.proxy.async.global;" ::: "memory");
    }
    __syncwarp();
}

// ---- consumer helpers ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_nc_u32(const void* p) {
    uint32_t r;
    asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ int ld_nc_s32(const void* p) {
    int r;
    asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}

// Q^T as the B operand (k = dim, n = head): b0 holds dims 2*(lane%4)+{0,1}, b1 the same +8, for
// head n = lane/4. The raw loads of a piece are issued one piece ahead (volatile asm: they stay
// where they are written) and finished - optional q_norm + RoPE, heads >= G zeroed - when the
// piece starts, so a row boundary costs the consumers no global-memory round trip.
struct QPre {
    uint32_t qf[8][2];
    int ctx;
};
__device__ __forceinline__ void q_prefetch(const DecodeParams& p, int b, int kvh, int G, int lane,
                                           QPre& q) {
    const int head = lane >> 2;
    const int hl = head < G ? head : 0;  // heads >= G: load row 0 (convergent), zeroed later
    const __nv_bfloat16* qrow = static_cast<const __nv_bfloat16*>(p.q) +
                                static_cast<int64_t>(b) * p.q_row_stride +
                                (static_cast<int64_t>(kvh) * G + hl) * kHeadDim + 2 * (lane & 3);
    q.ctx = ld_nc_s32(p.context_lens + b);
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
        q.qf[ks][0] = ld_nc_u32(qrow + ks * 16);
        q.qf[ks][1] = ld_nc_u32(qrow + ks * 16 + 8);
    }
}
// With p.nr.positions set, q rows are raw QKV-projection outputs and q_norm + RoPE
// (norm_rope.cuh) are applied in the fragment layout.
__device__ __forceinline__ void q_finish(const DecodeParams& p, int b, int G, int lane,
                                         uint32_t (&qf)[8][2]) {
    const int head = lane >> 2;
    if (p.nr.positions != nullptr) {
        // the 4 lanes of a head hold its whole row, lane (lane&3) the dims
        // ks*16 + 2*(lane&3) + {0, 1, 8, 9}; the rotation partner (dim +- 64) is register ks +- 4
        // of the same thread.
        const float* cs = cos_sin_row(p.nr, b) + 2 * (lane & 3);
        float x[8][4];
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            x[ks][0] = bf16_lo(qf[ks][0]);
            x[ks][1] = bf16_hi(qf[ks][0]);
            x[ks][2] = bf16_lo(qf[ks][1]);
            x[ks][3] = bf16_hi(qf[ks][1]);
        }
        if (p.nr.q_weight != nullptr) {
            float ss = 0.f;
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
#pragma unroll
                for (int i = 0; i < 4; ++i) ss += x[ks][i] * x[ks][i];
            ss += __shfl_xor_sync(0xffffffffu, ss, 1);
            ss += __shfl_xor_sync(0xffffffffu, ss, 2);
            const float r = rsqrtf(ss * (1.0f / 128.0f) + p.nr.eps);
            const __nv_bfloat16* wrow =
                static_cast<const __nv_bfloat16*>(p.nr.q_weight) + 2 * (lane & 3);
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                const uint32_t w0 = *reinterpret_cast<const uint32_t*>(wrow + ks * 16);
                const uint32_t w1 = *reinterpret_cast<const uint32_t*>(wrow + ks * 16 + 8);
                const float w[4] = {bf16_lo(w0), bf16_hi(w0), bf16_lo(w1), bf16_hi(w1)};
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    x[ks][i] = bf16_round(__fmul_rn(bf16_round(__fmul_rn(x[ks][i], r)), w[i]));
            }
        }
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
            // rotation index j = ks*16 + 2*(lane&3) + {0, 1, 8, 9}
            const float2 c01 = *reinterpret_cast<const float2*>(cs + ks * 16);
            const float2 c89 = *reinterpret_cast<const float2*>(cs + ks * 16 + 8);
            const float2 s01 = *reinterpret_cast<const float2*>(cs + 64 + ks * 16);
            const float2 s89 = *reinterpret_cast<const float2*>(cs + 64 + ks * 16 + 8);
            const float cv[4] = {c01.x, c01.y, c89.x, c89.y};
            const float sv[4] = {s01.x, s01.y, s89.x, s89.y};
            float y1[4], y2[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float x1 = x[ks][i], x2 = x[ks + 4][i];
                y1[i] = __fsub_rn(__fmul_rn(x1, cv[i]), __fmul_rn(x2, sv[i]));
                y2[i] = __fadd_rn(__fmul_rn(x2, cv[i]), __fmul_rn(x1, sv[i]));
            }
            qf[ks][0] = bf16_pack(y1[0], y1[1]);
            qf[ks][1] = bf16_pack(y1[2], y1[3]);
            qf[ks + 4][0] = bf16_pack(y2[0], y2[1]);
            qf[ks + 4][1] = bf16_pack(y2[2], y2[3]);
        }
    }
    if (head >= G) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) qf[ks][0] = qf[ks][1] = 0u;
    }
}

template <int kStages>
struct DecodeSmem {
    static constexpr int kRing = kStages * kStageBytes;
    static constexpr int kBars = kRing;                              // full[kStages] | empty[kStages]
    static constexpr int kMl = kBars + 2 * kStages * 8;              // [2][4 warps][8 heads][2] fp32
    static constexpr int kScratch = kMl + 2 * kConsumerWarps * 8 * 2 * 4;
    static constexpr int kCum = kScratch + kScratchFloats * 4;       // int32[kDecodeMaxBatch + 1]
    static constexpr int kWarpSums = kCum + (kDecodeMaxBatch + 2) * 4;  // int32[8] sums | int32[8] maxima
    static constexpr int kFlag = kWarpSums + 16 * 4;
    static constexpr int kQueueBars = kFlag + 16;                    // q_full[kPieceQueue] | q_empty[..]
    static constexpr int kQueue = (kQueueBars + 2 * kPieceQueue * 8 + 15) & ~15;  // PieceDesc[kPieceQueue]
    static constexpr int kBytes = kQueue + kPieceQueue * 32;
    static_assert(kBars % 8 == 0 && kQueueBars % 8 == 0, "mbarriers need 8-byte alignment");
    static_assert(kScratch % 16 == 0 && kQueue % 16 == 0, "float4 / int4 shared-memory accesses");
    static_assert(sizeof(PieceDesc) == 32, "queue slots are two int4");
};

// Debug trace (b200attn_debug_decode_trace): 16 uint64 per CTA.
//   0 globaltimer at entry   1 clock at entry        2 clock after the scan
//   3 producer: first piece known                    4 producer: first TMA issued
//   5 consumer: first tile landed                    6 consumer: first piece's tiles done
//   7 consumer: first piece's epilogue done          8 producer: all tiles issued (clock)
//   9 consumer: exit (clock)  10 pieces  11 tiles    12 SM id   13 clocks waiting for tiles
//  14 globaltimer at consumer exit                   15 clocks in epilogues
enum { kTrEntryNs = 0, kTrEntry, kTrScan, kTrPiece0, kTrTma0, kTrTile0, kTrTilesDone0, kTrEpi0,
       kTrProdExit, kTrConsExit, kTrPieces, kTrTiles, kTrSm, kTrWait, kTrExitNs, kTrEpi };
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// kRows = false: the persistent stream engine described above (grid = CTAs, work from the scan).
// kRows = true: the rows engine - grid (KV head, split, sequence), one piece per CTA, the hardware
//   block scheduler hands out the rows. No scan, no queue: every CTA derives its piece from its
//   own context_lens entry, so its first TMA is two dependent loads from kernel entry.
// kF16: fp16 storage / MMA operands (dtype code 1) instead of bf16; the 16-bit zero pattern, the
// tile geometry and the fp32 softmax are the same. The fused q/k norm + RoPE prologue
// (b200attn_qkv_decode) exists for bf16 only.
template <int kStages, bool kTrace, bool kRows, bool kF16 = false>
__global__ void __launch_bounds__(kDecodeThreads, 2)
decode_stream_kernel(const __grid_constant__ CUtensorMap k_map,
                     const __grid_constant__ CUtensorMap v_map, const DecodeParams p) {
    using L = DecodeSmem<kStages>;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBars);
    uint64_t* empty_bar = full_bar + kStages;
    float* ml_smem = reinterpret_cast<float*>(smem + L::kMl);
    float* scratch = reinterpret_cast<float*>(smem + L::kScratch);
    int32_t* cum = reinterpret_cast<int32_t*>(smem + L::kCum);
    int32_t* warp_sums = reinterpret_cast<int32_t*>(smem + L::kWarpSums);
    int* last_flag = reinterpret_cast<int*>(smem + L::kFlag);
    uint64_t* q_full = reinterpret_cast<uint64_t*>(smem + L::kQueueBars);
    uint64_t* q_empty = q_full + kPieceQueue;
    int4* piece_q = reinterpret_cast<int4*>(smem + L::kQueue);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int G = p.group;
    const int Hq = p.num_q_heads;
    const int B = p.batch;
    const int cta_linear = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    unsigned long long* trace = kTrace ? p.trace + static_cast<size_t>(cta_linear) * 16 : nullptr;
    if (kTrace && threadIdx.x == 0) {
        trace[kTrEntryNs] = global_timer_ns();
        trace[kTrEntry] = clock64();
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        trace[kTrSm] = smid;
    }

    // Let a programmatic dependent (the next decode launch with overlap on) start its own prologue
    // while this grid streams; it waits for this grid's completion before it touches data.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    int shared_pages = 0, t_skip = 0;
    if constexpr (!kRows) {
    // The block tables are read right after the scan (first piece's page window): start them on
    // their way into L2 now, one 128-byte line per thread, spread over the CTAs.
    {
        const int64_t lines =
            (static_cast<int64_t>(B) * p.table_row_stride * 4 + 127) / 128;
        const int64_t mine = static_cast<int64_t>(blockIdx.x) * kDecodeThreads + threadIdx.x;
        if (mine < lines && mine < 8192)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(
                reinterpret_cast<const uint8_t*>(p.block_tables) + mine * 128));
    }

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kConsumerWarps);
        }
#pragma unroll
        for (int s = 0; s < kPieceQueue; ++s) {
            mbar_init(&q_full[s], 1);
            mbar_init(&q_empty[s], kConsumerWarps);
        }
        fence_mbar_init();
    }

    // ---- exclusive scan of the per-sequence unit counts (every CTA, identical result) ----
    // Shared-prefix cascade: the first P pages of every active row were attended by the prefix
    // kernel; this kernel streams the remainder only. P pages lie entirely before token ctx-1 of
    // every active row, so the remainder is never empty.
    shared_pages = p.plan != nullptr ? p.plan[0] : 0;
    t_skip = shared_pages * (p.block_size / kTile);
    {
        const int per_thread = (B + kDecodeThreads - 1) / kDecodeThreads;
        const int i0 = threadIdx.x * per_thread;
        int units[(kDecodeMaxBatch + kDecodeThreads - 1) / kDecodeThreads];
        int local = 0, local_max = 0;
#pragma unroll
        for (int j = 0; j < (kDecodeMaxBatch + kDecodeThreads - 1) / kDecodeThreads; ++j) {
            const int i = i0 + j;
            units[j] = (j < per_thread && i < B)
                           ? decode_row_units(p.context_lens[i], t_skip, p.tun.row_cost) : 0;
            local += units[j];
            local_max = max(local_max, units[j]);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1)
            local_max = max(local_max, __shfl_xor_sync(0xffffffffu, local_max, off));
        int incl = local;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += v;
        }
        if (lane == 31) {
            warp_sums[warp] = incl;
            warp_sums[8 + warp] = local_max;
        }
        __syncthreads();  // also publishes the mbarrier inits
        int base = incl - local;
        for (int w = 0; w < warp; ++w) base += warp_sums[w];
#pragma unroll
        for (int j = 0; j < (kDecodeMaxBatch + kDecodeThreads - 1) / kDecodeThreads; ++j) {
            const int i = i0 + j;
            if (j < per_thread && i < B) cum[i] = base;
            base += units[j];
        }
        if (threadIdx.x == kDecodeThreads - 1) cum[B] = base;  // all threads before + own
        __syncthreads();
    }
    }
    // ---- rows engine: this CTA's single piece, from its own context_lens entry ----
    PieceDesc own{};
    if constexpr (kRows) {
        const int kvh = blockIdx.x, split = blockIdx.y, b = blockIdx.z;
        const int S = p.num_splits;
        const int ctx = p.context_lens[b];
        const int tiles_total = ctx > 0 ? (ctx + kTile - 1) / kTile : 0;
        // Shared-prefix cascade: the first P pages were attended by the prefix kernel.
        const int sp = (p.plan != nullptr && ctx > 0) ? p.plan[0] : 0;
        shared_pages = sp;
        t_skip = sp * (p.block_size / kTile);
        const int per_split = (tiles_total - t_skip + S - 1) / S;
        own.b = b;
        own.kvh = kvh;
        own.t_begin = t_skip + split * per_split;
        own.t_end = min(own.t_begin + per_split, tiles_total);
        own.tiles_total = tiles_total;
        own.range = split;
        own.g_first = 0;
        own.g_last = S - 1;
        if (own.t_begin >= own.t_end) {
            // Nothing to attend in this split (ctx == 0 padding row of a CUDA-graph batch, or a
            // split beyond the sequence). No table entry is read.
            if (p.overlap) grid_dependency_wait();
            const int64_t q_row0 = static_cast<int64_t>(b) * Hq + static_cast<int64_t>(kvh) * G;
            if (S == 1) {
                __nv_bfloat16* out = static_cast<__nv_bfloat16*>(p.out) + q_row0 * kHeadDim;
                for (int i = threadIdx.x; i < G * kHeadDim; i += kDecodeThreads)
                    out[i] = __float2bfloat16(0.f);
            } else {
                const int slot = (b * p.num_kv_heads + kvh) * S + split;
                if (threadIdx.x < 8) p.piece_lse[slot * 8 + threadIdx.x] = -INFINITY;
                if (piece_arrive_is_last<kDecodeThreads, 2>(
                        p.split_counters + b * p.num_kv_heads + kvh, S, last_flag)) {
                    // this empty split arrived last: it merges the row group
                    const int n_pre = sp > 0 ? prefix_chunks_of(sp, p.block_size,
                                                                p.prefix_chunks).valid : 0;
                    if (n_pre > 0) grid_dependency_wait();
                    MergeSrc src;
                    src.piece_o = p.piece_o;
                    src.piece_lse = p.piece_lse;
                    src.prefix_o = p.prefix_o;
                    src.prefix_lse = p.prefix_lse;
                    src.rows_base = (b * p.num_kv_heads + kvh) * S;
                    src.g_first = 0;
                    src.kvh = kvh;
                    src.num_kv_heads = p.num_kv_heads;
                    src.n_piece = S;
                    src.prefix_chunks = p.prefix_chunks;
                    src.n_prefix = n_pre;
                    for (int h = warp; h < G; h += kConsumerWarps + 1) {
                        src.h_local = h;
                        src.row = q_row0 + h;
                        merge_row<kF16>(src, static_cast<__nv_bfloat16*>(p.out), lane);
                    }
                }
            }
            if (p.plan != nullptr) grid_dependency_wait();
            return;
        }
        if (threadIdx.x == 0) {
#pragma unroll
            for (int s = 0; s < kStages; ++s) {
                mbar_init(&full_bar[s], 1);
                mbar_init(&empty_bar[s], kConsumerWarps);
            }
            fence_mbar_init();
        }
        __syncthreads();
    }
    if (kTrace && threadIdx.x == 0) trace[kTrScan] = clock64();

    const int n_prefix =
        shared_pages > 0 ? prefix_chunks_of(shared_pages, p.block_size, p.prefix_chunks).valid : 0;

    if (warp == kConsumerWarps) {
        // ===================== TMA producer (and the CTA's scheduler) =====================
        if (lane == 0) {
            tma_prefetch_desc(&k_map);
            tma_prefetch_desc(&v_map);
        }
        Stream st{};
        if constexpr (!kRows) {
            int max_row_units = 0;
#pragma unroll
            for (int w = 0; w < kConsumerWarps + 1; ++w)
                max_row_units = max(max_row_units, warp_sums[8 + w]);
            stream_init(st, cum, B, p.num_kv_heads, gridDim.x, p.tun, t_skip, max_row_units);
        }
        // Work items: every CTA starts with item = its id. Static regime: it goes on with items
        // id + CTAs, ... (only when there are fewer CTAs than ranges x KV heads). Dynamic regime:
        // further items come from the atomic counter sched[0]. The number is requested two tiles
        // before the current piece has been issued and used one tile later: early enough to hide
        // the round trip, late enough that no CTA sits on a reserved row while others go idle.
        // The consumers never see items, only finished piece descriptors (shared-memory queue).
        const bool dynamic = st.dynamic;
        const int tiles_per_page = p.block_size / kTile;
        const float rcp_tpp = 1.0f / static_cast<float>(tiles_per_page);
        const uint64_t policy = l2_policy_evict_first();
        const bool with_store = p.new_key != nullptr;
        ItemIter iter{};
        int item = blockIdx.x;
        bool item_live = !kRows && item < st.total_items;
        if (item_live) iter.init(st, item);
        int pre_item = 0;        // lane 0: the number requested from sched[0]
        bool pre_valid = false;  // a request is outstanding (warp-uniform)
        // Overlap mode: the preceding grid may be another decode launch that is still drawing
        // items from sched[0] (and whose last CTA resets it). The counter belongs to the stream
        // order: no request goes out before the wait. A CTA whose first item is an empty padding
        // row (ctx == 0) asks for its next item before it has issued a tile, i.e. in the prologue.
        bool waited = !p.overlap;
        auto request_item = [&]() {
            if (dynamic && !pre_valid) {
                if (!waited) {
                    grid_dependency_wait();
                    waited = true;
                }
                if (lane == 0) pre_item = static_cast<int>(gridDim.x) + atomicAdd(p.sched, 1);
                pre_valid = true;
            }
        };
        int n_pushed = 0;
        auto push = [&](const PieceDesc& d) {
            if (lane == 0) {
                const int qs = n_pushed % kPieceQueue;
                mbar_wait(&q_empty[qs], ((n_pushed / kPieceQueue) & 1u) ^ 1u);
                piece_q[2 * qs] = make_int4(d.b, d.kvh, d.t_begin, d.t_end);
                piece_q[2 * qs + 1] = make_int4(d.tiles_total, d.range, d.g_first, d.g_last);
                mbar_arrive(&q_full[qs]);
            }
            ++n_pushed;
        };
        // Next piece of this CTA (all lanes compute it redundantly), pushed to the consumers; the
        // end of the work is pushed as a descriptor with b = -1.
        auto next_piece = [&](Piece& pc) -> bool {
            for (;;) {
                if (!item_live) {
                    PieceDesc end{};
                    end.b = -1;
                    push(end);
                    return false;
                }
                if (iter.next(st, pc)) {
                    push(piece_desc(st, pc));
                    return true;
                }
                if (dynamic) {
                    request_item();  // normally requested a tile ago: the shuffle waits for it
                    item = __shfl_sync(0xffffffffu, pre_item, 0);
                    pre_valid = false;
                } else {
                    item += gridDim.x;
                }
                item_live = item < st.total_items;
                if (item_live) iter.init(st, item);
            }
        };
        Piece cur, nxt;
        bool have;
        if constexpr (kRows) {  // the one piece of this CTA; nothing is handed over
            cur.b = own.b;
            cur.kvh = own.kvh;
            cur.range = own.range;
            cur.t_begin = own.t_begin;
            cur.t_end = own.t_end;
            cur.tiles_total = own.tiles_total;
            cur.first = own.t_begin == t_skip;
            cur.last = own.t_end == own.tiles_total;
            have = true;
        } else {
            have = next_piece(cur);
        }
        if (kTrace && lane == 0) trace[kTrPiece0] = clock64();
        // block-table window of a piece: 32 pages from its first page, one entry per lane
        auto load_window = [&](const Piece& pc) {
            const int32_t* table =
                p.block_tables + static_cast<int64_t>(pc.b) * p.table_row_stride;
            const int page0 = idiv_rcp(pc.t_begin, tiles_per_page, rcp_tpp);
            const int last_page = idiv_rcp(pc.t_end - 1, tiles_per_page, rcp_tpp);
            return page0 + lane <= last_page ? ld_nc_s32(table + page0 + lane) : 0;
        };
        int cur_block = 0;
        ScatterPre cur_sc{};
        if (have) cur_block = load_window(cur);
        // Overlap mode: up to here only step metadata was read. The new K/V rows and the scheduler
        // counters belong to the stream order: wait for the preceding grid — the stream engine
        // here, the rows engine only in front of the ONE tile of a row that holds the newest token
        // (written by this step's scatter, fused or not). Every older tile was written by earlier
        // steps, so its loads go out while the preceding grid (the previous layer's decode) is
        // still draining: the ring is full when the consumers are released, and the first-tile
        // latency (~2-4 us of a ~20 us launch at the TP=8 shapes) is off the critical path.
        if (!kRows && !waited) {
            grid_dependency_wait();
            waited = true;
        }
        if (have && with_store && cur.last && waited)
            cur_sc = scatter_prefetch(p, cur.b, cur.kvh, lane);
        int it = 0;
        while (have) {
            if (with_store && cur.last && waited)
                scatter_finish(p, cur_sc, cur.b, cur.kvh, lane);
            const int32_t* table =
                p.block_tables + static_cast<int64_t>(cur.b) * p.table_row_stride;
            const int last_page = idiv_rcp(cur.t_end - 1, tiles_per_page, rcp_tpp);
            int page_base = idiv_rcp(cur.t_begin, tiles_per_page, rcp_tpp);
            int page = page_base;
            int in_page = cur.t_begin - page_base * tiles_per_page;
            int my_block = cur_block;
            bool have_next = false;
            int nxt_block = 0;
            ScatterPre nxt_sc{};
            const int request_t = max(cur.t_begin, cur.t_end - 2);
            for (int t = cur.t_begin; t < cur.t_end; ++t, ++it) {
                const int stage = it % kStages;
                const uint32_t phase = (it / kStages) & 1u;
                if (!kRows && t == request_t && iter.u >= iter.hi) request_item();
                if (!kRows && t == cur.t_end - 1) {
                    // before the last tile goes out: the next piece, handed to the consumers (they
                    // take it when this tile has landed), its table window and new K/V row
                    have_next = next_piece(nxt);
                    if (have_next) {
                        nxt_block = load_window(nxt);
                        if (with_store && nxt.last)
                            nxt_sc = scatter_prefetch(p, nxt.b, nxt.kvh, lane);
                    }
                }
                if (kRows && !waited && t == cur.tiles_total - 1) {
                    grid_dependency_wait();
                    waited = true;
                    if (with_store && cur.last) {
                        cur_sc = scatter_prefetch(p, cur.b, cur.kvh, lane);
                        scatter_finish(p, cur_sc, cur.b, cur.kvh, lane);
                    }
                }
                if (page - page_base >= 32) {
                    page_base = page;
                    my_block = (page_base + lane <= last_page) ? table[page_base + lane] : 0;
                }
                const int block_id = __shfl_sync(0xffffffffu, my_block, page - page_base);
                const int slot = block_id * p.block_size + in_page * kTile;
                if (lane == 0) {
                    uint8_t* dst = smem + stage * kStageBytes;
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    mbar_expect_tx(&full_bar[stage], kStageBytes);
                    tma_load_3d_hint(dst, &k_map, &full_bar[stage], 0, cur.kvh, slot, policy);
                    tma_load_3d_hint(dst + kSubTileBytes, &k_map, &full_bar[stage], 64, cur.kvh,
                                     slot, policy);
                    tma_load_3d_hint(dst + 2 * kSubTileBytes, &v_map, &full_bar[stage], 0, cur.kvh,
                                     slot, policy);
                    tma_load_3d_hint(dst + 3 * kSubTileBytes, &v_map, &full_bar[stage], 64,
                                     cur.kvh, slot, policy);
                    if (kTrace && it == 0) trace[kTrTma0] = clock64();
                }
                __syncwarp();
                if (++in_page == tiles_per_page) {
                    in_page = 0;
                    ++page;
                }
            }
            cur = nxt;
            cur_block = nxt_block;
            cur_sc = nxt_sc;
            have = have_next;
        }
        if (kTrace && lane == 0) trace[kTrProdExit] = clock64();
        // Dynamic regime: the CTA that finishes last puts the counters back to zero for the next
        // launch (every CTA passes here exactly once, after its last request has been answered).
        if (dynamic && lane == 0) {
            asm volatile("" ::"r"(pre_item) : "memory");
            const int old = atomicAdd(p.sched + 1, 1);
            if (old == static_cast<int>(gridDim.x) - 1) {
                p.sched[0] = 0;
                p.sched[1] = 0;
            }
        }
        // With the cascade armed this grid is a programmatic dependent of the prefix pass: make
        // sure it cannot finish before its primary did.
        if (p.plan != nullptr) grid_dependency_wait();
        return;
    }

    // ===================== consumers =====================
    // Overlap mode: everything above read only step metadata (context_lens, block tables), which
    // the caller promises is not written by the kernel right before this one. From here on the
    // consumers read q and write out / workspace / counters: wait for the preceding grid.
    if (p.overlap) grid_dependency_wait();
    // Padding rows of a CUDA-graph batch (ctx == 0) carry no units: their output rows are zeroed
    // here, spread over the CTAs. No table entry of such a row is ever read. (Rows engine: the
    // CTAs of such a row did it above.)
    for (int b = blockIdx.x; !kRows && b < B; b += gridDim.x) {
        if (cum[b + 1] == cum[b]) {
            uint32_t* o32 = reinterpret_cast<uint32_t*>(static_cast<__nv_bfloat16*>(p.out) +
                                                        static_cast<int64_t>(b) * Hq * kHeadDim);
            for (int i = threadIdx.x; i < Hq * kHeadDim / 2; i += kConsumerWarps * 32) o32[i] = 0u;
        }
    }
    const uint32_t smem_base = smem_u32(smem);
    // K A-fragment addressing (rows = tokens): lane -> row lane%16, 16-byte chunk +(lane/16)
    const uint32_t k_row = warp * 16 + (lane & 15);
    const uint32_t k_csel = lane >> 4;
    // V^T A-fragment addressing via ldmatrix.trans: matrix j = lane/8 -> tokens (j/2)*8 + lane%8,
    // dim chunk + (j%2)
    const uint32_t v_row = warp * 16 + ((lane >> 4) << 3) + (lane & 7);
    const uint32_t v_csel = (lane >> 3) & 1;

    int n_popped = 0;
    auto pop = [&]() -> PieceDesc {
        const int qs = n_popped % kPieceQueue;
        mbar_wait(&q_full[qs], (n_popped / kPieceQueue) & 1u);
        const int4 lo = piece_q[2 * qs], hi = piece_q[2 * qs + 1];
        __syncwarp();
        if (lane == 0) mbar_arrive(&q_empty[qs]);
        ++n_popped;
        PieceDesc d;
        d.b = lo.x;
        d.kvh = lo.y;
        d.t_begin = lo.z;
        d.t_end = lo.w;
        d.tiles_total = hi.x;
        d.range = hi.y;
        d.g_first = hi.z;
        d.g_last = hi.w;
        return d;
    };
    PieceDesc pc;
    if constexpr (kRows) pc = own; else pc = pop();
    QPre qpre;
    if (pc.b >= 0) q_prefetch(p, pc.b, pc.kvh, G, lane, qpre);
    int it = 0;
    int parity = 0;
    unsigned long long tr_wait = 0, tr_epi = 0, tr_t0 = 0;
    int tr_pieces = 0;
    while (pc.b >= 0) {
        const int b = pc.b, kvh = pc.kvh;
        const int ctx = qpre.ctx;
        uint32_t qf[8][2];
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            qf[ks][0] = qpre.qf[ks][0];
            qf[ks][1] = qpre.qf[ks][1];
        }
        q_finish(p, b, G, lane, qf);
        PieceDesc nxt;
        nxt.b = -1;
        // A cut row is merged by the piece that arrives last at its counter: the ranges that hold
        // the row's first and last tile give the number of pieces and their workspace slots.
        const int g_first = pc.g_first, g_last = pc.g_last;
        const bool whole = g_first == g_last;
        const bool first_piece = pc.range == g_first;


        float m_run[2] = {-INFINITY, -INFINITY};  // running max (log2 domain), heads 2*(lane%4)+{0,1}
        float l_run[2] = {0.f, 0.f};              // per-thread partial row sums
        float o[8][4];
#pragma unroll
        for (int mt = 0; mt < 8; ++mt)
#pragma unroll
            for (int c = 0; c < 4; ++c) o[mt][c] = 0.f;

        for (int t = pc.t_begin; t < pc.t_end; ++t, ++it) {
            const int stage = it % kStages;
            const uint32_t phase = (it / kStages) & 1u;
            if (kTrace && threadIdx.x == 0) {
                const unsigned long long w0 = clock64();
                mbar_wait(&full_bar[stage], phase);
                const unsigned long long w1 = clock64();
                tr_wait += w1 - w0;
                if (it == 0) trace[kTrTile0] = w1;
            } else {
                mbar_wait(&full_bar[stage], phase);
            }
            if (!kRows && t == pc.t_end - 1) {
                // The producer pushed the next descriptor before it issued this (last) tile, so
                // it is there now. Its q loads run under this tile and the epilogue.
                nxt = pop();
                if (nxt.b >= 0) q_prefetch(p, nxt.b, nxt.kvh, G, lane, qpre);
            }
            const uint32_t kb = smem_base + stage * kStageBytes;
            const uint32_t vb = kb + 2 * kSubTileBytes;

            // ---- S^T[16 tok x 8 heads] = K . Q^T ----
            float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                uint32_t a[4];
                ldmatrix_x4(a, kb + (ks >> 2) * kSubTileBytes +
                                   swz128(k_row, (ks & 3) * 2 + k_csel));
                mma_16816<kF16>(s, a, qf[ks][0], qf[ks][1]);
            }

            // ---- mask + online softmax (log2 domain) ----
            const int tok = t * kTile + warp * 16 + (lane >> 2);
            const bool v0 = tok < ctx, v1 = tok + 8 < ctx;
            s[0] = v0 ? s[0] * p.scale_log2 : -INFINITY;
            s[1] = v0 ? s[1] * p.scale_log2 : -INFINITY;
            s[2] = v1 ? s[2] * p.scale_log2 : -INFINITY;
            s[3] = v1 ? s[3] * p.scale_log2 : -INFINITY;
            float mx0 = fmaxf(s[0], s[2]), mx1 = fmaxf(s[1], s[3]);
#pragma unroll
            for (int off = 4; off < 32; off <<= 1) {
                mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, off));
                mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, off));
            }
            const float mn0 = fmaxf(m_run[0], mx0), mn1 = fmaxf(m_run[1], mx1);
            const float mu0 = mn0 == -INFINITY ? 0.f : mn0, mu1 = mn1 == -INFINITY ? 0.f : mn1;
            const float alpha0 = fast_exp2(m_run[0] - mu0), alpha1 = fast_exp2(m_run[1] - mu1);
            m_run[0] = mn0;
            m_run[1] = mn1;
            const float p0 = fast_exp2(s[0] - mu0), p1 = fast_exp2(s[1] - mu1);
            const float p2 = fast_exp2(s[2] - mu0), p3 = fast_exp2(s[3] - mu1);
            l_run[0] = l_run[0] * alpha0 + p0 + p2;
            l_run[1] = l_run[1] * alpha1 + p1 + p3;

            // Tokens beyond ctx inside the last page are whatever the allocator left there; P is
            // 0 for them but 0 * NaN would poison O, so their V rows are cleared in shared memory.
            if (t == pc.tiles_total - 1) {
                const int valid_rows = ctx - (t * kTile + warp * 16);  // rows of this warp's 16
                if (valid_rows < 16) {
                    const int r_lo = max(valid_rows, 0);
                    if (lane < 16) {
                        uint8_t* vsm = smem + stage * kStageBytes + 2 * kSubTileBytes +
                                       (lane >> 3) * kSubTileBytes;
                        for (int r = r_lo; r < 16; ++r)
                            *reinterpret_cast<int4*>(vsm + (warp * 16 + r) * 128 +
                                                     (lane & 7) * 16) = make_int4(0, 0, 0, 0);
                    }
                    __syncwarp();
                }
            }

            // ---- O^T[128 dims x 8 heads] = alpha * O^T + V^T . P^T ----
            const uint32_t pb0 = movmatrix_trans(pack_16x2<kF16>(p0, p1));
            const uint32_t pb1 = movmatrix_trans(pack_16x2<kF16>(p2, p3));
#pragma unroll
            for (int mt = 0; mt < 8; ++mt) {
                uint32_t a[4];
                ldmatrix_x4_trans(
                    a, vb + (mt >> 2) * kSubTileBytes + swz128(v_row, (mt & 3) * 2 + v_csel));
                o[mt][0] *= alpha0;
                o[mt][1] *= alpha1;
                o[mt][2] *= alpha0;
                o[mt][3] *= alpha1;
                mma_16816<kF16>(o[mt], a, pb0, pb1);
            }

            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[stage]);
        }

        if (kTrace && threadIdx.x == 0) {
            tr_t0 = clock64();
            if (tr_pieces == 0) trace[kTrTilesDone0] = tr_t0;
        }
        // ---- finish the row sums (tokens are spread over lane/4) ----
#pragma unroll
        for (int off = 4; off < 32; off <<= 1) {
            l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], off);
            l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], off);
        }

        // ---- combine the 4 warps: (m, l) through ml[parity], O^T through a two-level tree in
        // `scratch` (never part of the ring: the producer keeps streaming the next piece). The
        // parity double buffer lets a fast warp start the next piece's epilogue while a slow one
        // still reads this piece's (m, l); every scratch write sits behind a barrier that all
        // warps reach only after their reads of the previous piece. ----
        float* ml = ml_smem + parity * (kConsumerWarps * 8 * 2);
        parity ^= 1;
        if (lane < 4) {
            float* dst = ml + (warp * 8 + 2 * lane) * 2;
            dst[0] = m_run[0];
            dst[1] = l_run[0];
            dst[2] = m_run[1];
            dst[3] = l_run[1];
        }
        consumer_bar_sync();
        float fac[2];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int h = 2 * (lane & 3) + i;
            float m = -INFINITY;
#pragma unroll
            for (int w = 0; w < kConsumerWarps; ++w) m = fmaxf(m, ml[(w * 8 + h) * 2]);
            fac[i] = m == -INFINITY ? 0.f : fast_exp2(m_run[i] - m);
        }
#pragma unroll
        for (int mt = 0; mt < 8; ++mt)
#pragma unroll
            for (int c = 0; c < 4; ++c) o[mt][c] *= fac[c & 1];
        if (warp >= 2) {
#pragma unroll
            for (int mt = 0; mt < 8; ++mt)
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    scratch[scratch_idx(warp - 2, 2 * (lane & 3) + (c & 1),
                                        mt * 16 + (lane >> 2) + (c >> 1) * 8)] = o[mt][c];
        }
        consumer_bar_sync();
        if (warp < 2) {
#pragma unroll
            for (int mt = 0; mt < 8; ++mt)
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int idx = scratch_idx(warp, 2 * (lane & 3) + (c & 1),
                                                mt * 16 + (lane >> 2) + (c >> 1) * 8);
                    scratch[idx] += o[mt][c];
                }
        }
        consumer_bar_sync();

        // Final reduction: warp w owns heads w, w+4; lane owns dims 4*lane..4*lane+3 (float4
        // shared-memory reads, 8-byte bf16 / 16-byte fp32 global stores, a head row is one line).
        const int64_t q_row0 = static_cast<int64_t>(b) * Hq + static_cast<int64_t>(kvh) * G;
        const int rows_base = kRows ? (b * p.num_kv_heads + kvh) * p.num_splits : -1;
        const int slot = kRows ? rows_base + pc.range
                               : 2 * (pc.range * p.num_kv_heads + kvh) + (first_piece ? 1 : 0);
        for (int h = warp; h < G; h += kConsumerWarps) {
            float m = -INFINITY;
#pragma unroll
            for (int w = 0; w < kConsumerWarps; ++w) m = fmaxf(m, ml[(w * 8 + h) * 2]);
            float l = 0.f;
#pragma unroll
            for (int w = 0; w < kConsumerWarps; ++w) {
                const float mw = ml[(w * 8 + h) * 2];
                l += mw == -INFINITY ? 0.f : ml[(w * 8 + h) * 2 + 1] * fast_exp2(mw - m);
            }
            const int d4 = (4 * lane) ^ ((h >> 1) << 3);
            const float4 x0 = *reinterpret_cast<const float4*>(&scratch[(0 * 8 + h) * kHeadDim + d4]);
            const float4 x1 = *reinterpret_cast<const float4*>(&scratch[(1 * 8 + h) * kHeadDim + d4]);
            float4 acc = make_float4(x0.x + x1.x, x0.y + x1.y, x0.z + x1.z, x0.w + x1.w);
            const int64_t row = q_row0 + h;
            if (whole) {
                if (n_prefix > 0) {
                    grid_dependency_wait();  // the prefix pass (PDL primary) has written its slots
                    // fold in the shared-prefix partials (normalised O_c, lse_c) of this row:
                    // out = (acc 2^(m-M) + sum_c O_c 2^(lse_c-M)) / (l 2^(m-M) + sum_c 2^(lse_c-M))
                    // lane c holds lse_c; the O_c loads are independent and unrolled.
                    const float my_lse =
                        lane < n_prefix ? p.prefix_lse[row * p.prefix_chunks + lane] : -INFINITY;
                    float M = fmaxf(m, my_lse);
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1)
                        M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, off));
                    const float f = m == -INFINITY ? 0.f : fast_exp2(m - M);
                    const float my_w = my_lse == -INFINITY ? 0.f : fast_exp2(my_lse - M);
                    acc.x *= f;
                    acc.y *= f;
                    acc.z *= f;
                    acc.w *= f;
                    l *= f;
                    const float4* o_c = reinterpret_cast<const float4*>(
                                            p.prefix_o + row * p.prefix_chunks * kHeadDim) + lane;
                    for (int c0 = 0; c0 < n_prefix; c0 += 8) {
                        float4 x[8];
#pragma unroll
                        for (int u = 0; u < 8; ++u)
                            x[u] = c0 + u < n_prefix ? o_c[(c0 + u) * (kHeadDim / 4)]
                                                     : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            const float w = __shfl_sync(0xffffffffu, my_w, (c0 + u) & 31);
                            const float wu = c0 + u < n_prefix ? w : 0.f;
                            acc.x += wu * x[u].x;
                            acc.y += wu * x[u].y;
                            acc.z += wu * x[u].z;
                            acc.w += wu * x[u].w;
                            l += wu;
                        }
                    }
                }
                const bool live = l > 0.f;
                uint2 pk;
                pk.x = pack_16x2<kF16>(live ? acc.x / l : 0.f, live ? acc.y / l : 0.f);
                pk.y = pack_16x2<kF16>(live ? acc.z / l : 0.f, live ? acc.w / l : 0.f);
                *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(p.out) + row * kHeadDim +
                                          4 * lane) = pk;
            } else {
                const bool live = l > 0.f;
                acc.x = live ? acc.x / l : 0.f;
                acc.y = live ? acc.y / l : 0.f;
                acc.z = live ? acc.z / l : 0.f;
                acc.w = live ? acc.w / l : 0.f;
                *reinterpret_cast<float4*>(
                    p.piece_o + (static_cast<int64_t>(slot) * 8 + h) * kHeadDim + 4 * lane) = acc;
                if (lane == 0) p.piece_lse[slot * 8 + h] = live ? m + log2f(l) : -INFINITY;
            }
        }
        if (!whole &&
            piece_arrive_is_last<kConsumerWarps * 32, 1>(
                p.split_counters + b * p.num_kv_heads + kvh, g_last - g_first + 1, last_flag)) {
            // last piece of this (sequence, KV head) to finish: merge its G rows here
            if (n_prefix > 0) grid_dependency_wait();
            MergeSrc src;
            src.piece_o = p.piece_o;
            src.piece_lse = p.piece_lse;
            src.prefix_o = p.prefix_o;
            src.prefix_lse = p.prefix_lse;
            src.rows_base = rows_base;
            src.g_first = g_first;
            src.kvh = kvh;
            src.num_kv_heads = p.num_kv_heads;
            src.n_piece = g_last - g_first + 1;
            src.prefix_chunks = p.prefix_chunks;
            src.n_prefix = n_prefix;
            for (int h = warp; h < G; h += kConsumerWarps) {
                src.h_local = h;
                src.row = q_row0 + h;
                merge_row<kF16>(src, static_cast<__nv_bfloat16*>(p.out), lane);
            }
        }
        if (kTrace && threadIdx.x == 0) {
            const unsigned long long now = clock64();
            if (tr_pieces == 0) trace[kTrEpi0] = now;
            tr_epi += now - tr_t0;
            ++tr_pieces;
        }
        pc = nxt;
    }
    if (kTrace && threadIdx.x == 0) {
        trace[kTrConsExit] = clock64();
        trace[kTrExitNs] = global_timer_ns();
        trace[kTrPieces] = tr_pieces;
        trace[kTrTiles] = it;
        trace[kTrWait] = tr_wait;
        trace[kTrEpi] = tr_epi;
    }
    if (p.plan != nullptr) grid_dependency_wait();
}

template <int kStages, bool kTrace, bool kRows, bool kF16 = false>
cudaError_t launch_stream(const DecodeParams& p, const CUtensorMap& k_map, const CUtensorMap& v_map,
                          cudaStream_t stream) {
    constexpr int smem = DecodeSmem<kStages>::kBytes;
    // per device, not per process: the attribute belongs to the function on the current device
    static bool configured[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        cudaError_t e = cudaFuncSetAttribute(decode_stream_kernel<kStages, kTrace, kRows, kF16>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    dim3 grid(p.num_ctas);
    if (kRows) grid = dim3(p.num_kv_heads, p.num_splits, p.batch);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kDecodeThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    if (p.plan != nullptr || p.overlap) {
        // Overlap mode: dependent of whatever kernel precedes it in the stream (see DecodeParams).
        // Programmatic dependent launch behind the prefix pass: this grid may start while the
        // prefix CTAs are still running (they signal launch_dependents at once) and only waits
        // (griddepcontrol.wait) where it consumes their partials. It touches nothing the prefix
        // pass reads or writes before that point: the fused scatter writes token ctx-1, which
        // lies after the shared pages, and the piece partials go to another workspace region.
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
    }
    cudaError_t e =
        cudaLaunchKernelEx(&cfg, decode_stream_kernel<kStages, kTrace, kRows, kF16>, k_map, v_map, p);
    count_launch();
    return e != cudaSuccess ? e : cudaGetLastError();
}

int env_int(const char* name, int dflt, int lo, int hi) {
    const char* e = std::getenv(name);
    if (!e || !e[0]) return dflt;
    const int v = std::atoi(e);
    return v < lo || v > hi ? dflt : v;
}

int decode_stages_from_env() {
    static int stages = [] {
        const int s = env_int("B200ATTN_DECODE_STAGES", 3, 2, 3);
        return s;
    }();
    return stages;
}

}  // namespace

// Tunables of the unit stream (read once): B200ATTN_DECODE_MIN_TILES = fewest units a CTA takes
// before another CTA is started (below that the partials + merge cost more than the extra
// parallelism hides); B200ATTN_DECODE_ROW_COST = virtual units charged per row for its fixed cost.
int decode_min_units() {
    static int v = env_int("B200ATTN_DECODE_MIN_TILES", 16, 1, 4096);
    return v;
}
int decode_row_cost() {
    static int v = env_int("B200ATTN_DECODE_ROW_COST", 1, 0, 64);
    return v;
}

int decode_snap() {
    static int v = env_int("B200ATTN_DECODE_SNAP", 2, 0, 64);
    return v;
}
// Dynamic regime (whole rows from an atomic counter): B200ATTN_DECODE_DYNAMIC=0 switches it off;
// it is used when the longest row is at most B200ATTN_DECODE_DYN_PCT % of a CTA group's share.
int decode_dyn_ok() {
    static int v = env_int("B200ATTN_DECODE_DYNAMIC", 1, 0, 1);
    return v;
}
int decode_dyn_pct() {
    static int v = env_int("B200ATTN_DECODE_DYN_PCT", 200, 0, 100000);
    return v;
}
// Static regime: percent of the CTA groups that get a range.
int decode_static_pct() {
    static int v = env_int("B200ATTN_DECODE_STATIC_PCT", 50, 1, 100);
    return v;
}

bool decode_engine_rows() {
    static const bool rows = [] {
        const char* e = std::getenv("B200ATTN_DECODE_ENGINE");
        return !(e && e[0] == 's');  // "stream" selects the persistent engine
    }();
    return rows;
}

// Slot pairs of the piece region: one per work item (range, KV head) of the STATIC regime (the
// dynamic regime never cuts a row, so it writes no partials): ranges * Hkv <= max(ctas, Hkv).
static size_t piece_pairs(int32_t num_q_heads, int32_t num_ctas) {
    return static_cast<size_t>(num_ctas) + static_cast<size_t>(num_q_heads);
}

// Partial-piece slots ([8 heads][128] fp32 O + [8] lse each). Stream engine: two per work item
// that can hold a cut row. Rows engine: one per (sequence, KV head, split); the KV-head count is
// not an argument of the ABI's size query, so the bound num_q_heads >= num_kv_heads is used.
static size_t piece_slots(int32_t batch, int32_t num_q_heads, int32_t num_splits) {
    if (decode_engine_rows())
        return static_cast<size_t>(batch) * num_q_heads * num_splits;
    return piece_pairs(num_q_heads, num_splits) * 2;
}

size_t decode_workspace_bytes(int32_t batch, int32_t num_q_heads, int32_t num_ctas,
                              int32_t prefix_chunks) {
    size_t floats = 0;
    if (num_ctas > 1) floats += piece_slots(batch, num_q_heads, num_ctas) * 8 * (kHeadDim + 1);
    floats = (floats + 3) & ~static_cast<size_t>(3);  // keep the prefix region 16-byte aligned
    if (prefix_chunks > 0)
        floats += static_cast<size_t>(batch) * num_q_heads * prefix_chunks * (kHeadDim + 1);
    return floats * sizeof(float);
}

void decode_workspace_layout(void* workspace, int32_t batch, int32_t num_q_heads, int32_t num_ctas,
                             int32_t prefix_chunks, DecodeParams* p) {
    float* w = static_cast<float*>(workspace);
    size_t at = 0;
    if (num_ctas > 1) {
        const size_t slots = piece_slots(batch, num_q_heads, num_ctas);
        p->piece_o = w;
        p->piece_lse = w + slots * 8 * kHeadDim;
        at = slots * 8 * (kHeadDim + 1);
    }
    at = (at + 3) & ~static_cast<size_t>(3);
    if (prefix_chunks > 0) {
        p->prefix_o = w + at;
        p->prefix_lse =
            p->prefix_o + static_cast<size_t>(batch) * num_q_heads * prefix_chunks * kHeadDim;
    }
}

// Rows engine: equal pieces per row, from host-visible sizes only (graph-capture rule).
// Calibrated on B200 (profiles/r01_sweep_splits.log): one streaming CTA per SM already saturates
// HBM, so splitting only pays while (sequence, KV head) rows are fewer than ~0.7 x SMs; then aim
// for about one CTA per SM and never cut the longest possible context into pieces shorter than 8
// tiles (512 tokens) - below that the partials + merge cost more than they hide.
static int decode_pick_row_splits(int32_t batch, int32_t num_kv_heads, int32_t table_width,
                                  int32_t block_size, int32_t sm_count) {
    const int64_t base = static_cast<int64_t>(batch) * num_kv_heads;
    if (base <= 0 || base * 10 >= static_cast<int64_t>(sm_count) * 7) return 1;
    const int64_t max_tiles = static_cast<int64_t>(table_width) * block_size / kTile;
    int64_t s = (sm_count + base / 2) / base;  // round(sm_count / base)
    const int64_t cap = max_tiles / 8 > 1 ? max_tiles / 8 : 1;
    if (s > cap) s = cap;
    if (s > 64) s = 64;
    return static_cast<int>(s < 1 ? 1 : s);
}

int decode_pick_ctas(int32_t batch, int32_t num_kv_heads, int32_t table_width, int32_t block_size,
                     int32_t sm_count) {
    if (decode_engine_rows())
        return decode_pick_row_splits(batch, num_kv_heads, table_width, block_size, sm_count);
    // Host-visible sizes only (graph-capture rule): two resident CTAs per SM, fewer when even the
    // longest contexts the table can describe would leave CTAs with less than min_units of work;
    // a multiple of the KV-head count (CTAs = groups x head lanes). How the work is cut is decided
    // on the device from context_lens.
    const int64_t rows = static_cast<int64_t>(batch) * num_kv_heads;
    if (rows <= 0) return 1;
    const int64_t max_units =
        rows * (static_cast<int64_t>(table_width) * block_size / kTile + decode_row_cost());
    int64_t n = (max_units + decode_min_units() - 1) / decode_min_units();
    static const int per_sm = env_int("B200ATTN_DECODE_CTAS_PER_SM", 2, 1, 2);
    const int64_t cap = per_sm * static_cast<int64_t>(sm_count);
    if (n < rows) n = rows;  // never fewer CTAs than rows (while they fit): a row is never queued
    if (n > cap) n = cap;
    if (n >= num_kv_heads) n -= n % num_kv_heads;
    return static_cast<int>(n < 1 ? 1 : n);
}

StreamTunables decode_tunables() {
    StreamTunables t;
    t.min_units = decode_min_units();
    t.row_cost = decode_row_cost();
    t.snap = decode_snap();
    t.dyn_ok = decode_dyn_ok();
    t.dyn_pct = decode_dyn_pct();
    t.static_pct = decode_static_pct();
    return t;
}

int decode_partition_host(const int32_t* context_lens, int32_t batch, int32_t num_kv_heads,
                          int32_t num_ctas, int32_t shared_pages, int32_t block_size,
                          int32_t* pieces, int32_t max_pieces) {
    // The kernel's own partition code (decode_plan.cuh) run on the host: scan, ranges, pieces.
    const int t_skip = shared_pages * (block_size / kTile);
    StreamTunables tun = decode_tunables();
    if (num_ctas == 1) tun.dyn_ok = 0;  // as the decode entry points do without counters
    int32_t* cum = new int32_t[batch + 1];
    cum[0] = 0;
    int max_row_units = 0;
    for (int b = 0; b < batch; ++b) {
        const int u = decode_row_units(context_lens[b], t_skip, tun.row_cost);
        cum[b + 1] = cum[b] + u;
        max_row_units = imax(max_row_units, u);
    }
    Stream st;
    stream_init(st, cum, batch, num_kv_heads, num_ctas, tun, t_skip, max_row_units);
    int n = 0;
    for (int item = 0; item < st.total_items; ++item) {
        ItemIter iter;
        iter.init(st, item);
        Piece pc;
        while (iter.next(st, pc)) {
            if (n < max_pieces) {
                int32_t* o = pieces + static_cast<int64_t>(n) * 12;
                const PieceDesc d = piece_desc(st, pc);  // what the consumers are handed
                const int g_first = d.g_first, g_last = d.g_last;
                if (!(pc.first && pc.last)) {
                    // the closed forms must agree with the definition (search over the boundaries)
                    int gf, gl;
                    piece_ranges(st, pc.b, &gf, &gl);
                    if (gf != g_first || gl != g_last) {
                        delete[] cum;
                        return -1000 - item;
                    }
                }
                o[0] = item;
                o[1] = pc.b;
                o[2] = pc.kvh;
                o[3] = pc.t_begin;
                o[4] = pc.t_end;
                o[5] = pc.tiles_total;
                o[6] = pc.first;
                o[7] = pc.last;
                o[8] = g_first;
                o[9] = g_last;
                o[10] = pc.range;
                o[11] = st.per | (st.dynamic ? (1 << 30) : 0);
            }
            ++n;
        }
    }
    delete[] cum;
    return n;
}

cudaError_t launch_decode(const DecodeParams& p, const CUtensorMap& k_map, const CUtensorMap& v_map,
                          cudaStream_t stream) {
    if (p.batch == 0) return cudaSuccess;
    if (p.f16) return launch_stream<3, false, true, true>(p, k_map, v_map, stream);  // rows engine
    if (decode_engine_rows()) {
        if (p.trace != nullptr) return launch_stream<3, true, true>(p, k_map, v_map, stream);
        switch (decode_stages_from_env()) {
            case 2: return launch_stream<2, false, true>(p, k_map, v_map, stream);
            default: return launch_stream<3, false, true>(p, k_map, v_map, stream);
        }
    }
    if (p.trace != nullptr) return launch_stream<3, true, false>(p, k_map, v_map, stream);
    switch (decode_stages_from_env()) {
        case 2: return launch_stream<2, false, false>(p, k_map, v_map, stream);
        default: return launch_stream<3, false, false>(p, k_map, v_map, stream);
    }
}

}  // namespace b200
