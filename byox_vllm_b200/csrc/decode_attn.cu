// Paged decode attention for sm_100a: split-KV over KV pages, TMA-staged K/V tiles, GQA packing.
// Replaces flash_attn_with_kvcache on the decode branch of Attention.forward
// (reference nanovllm/layers/attention.py:87-96).
//
// Work item = (sequence b, KV head, split). One CTA = 4 consumer warps + 1 TMA producer warp.
//   * The producer streams 64-token tiles of the (b, kv-head) slice of K and V through a
//     kStages-deep shared-memory ring with cp.async.bulk.tensor (SWIZZLE_128B boxes of
//     64 tokens x 64 dims; one page of 256 tokens = 4 tiles, so tiles never straddle pages) and
//     mbarrier expect-tx completion. Block-table entries are fetched 32 pages at a time by the
//     producer warp so the table lookup is off the critical path.
//   * GQA packing: the G = Hq/Hkv query heads that share this KV head are the N dimension of
//     mma.sync.m16n8k16 (N = 8 >= G), tokens are M. So every K/V byte is read from HBM once per
//     KV head and serves all G query heads:  S^T[tok, head] = K[tok, :] . Q^T[:, head].
//   * Each consumer warp owns 16 tokens of every tile and keeps its own online-softmax state
//     (running max / sum per head, O^T[128 dims x 8 heads] in registers); the max over tokens is
//     a 3-step warp shuffle. P^T is moved from accumulator to B-operand layout with
//     movmatrix.trans, and O^T += V^T . P^T uses ldmatrix.trans on the V tile.
//   * The 4 warps are merged through shared memory at the end; with num_splits > 1 the CTA writes
//     a normalised fp32 partial + log2-sum-exp and a small merge kernel combines the splits.
//
// Roofline: HBM. Algorithmic bytes per (b, kv-head) = ctx * 128 * 2 B * 2 (K and V), read once.
#include "common.h"
#include "ptx.cuh"

#include <cstdlib>

namespace b200 {

namespace {

constexpr int kTile = 64;                   // tokens per stage
constexpr int kConsumerWarps = 4;           // 16 tokens each
constexpr int kDecodeThreads = (kConsumerWarps + 1) * 32;
constexpr int kSubTileBytes = kTile * 128;  // [64 tok][64 dims] bf16 = 8 KiB, 128-byte rows
constexpr int kStageBytes = 4 * kSubTileBytes;  // K lo/hi + V lo/hi = 32 KiB
constexpr int kMergeStride = 132;           // padded fp32 row (bank-conflict-free transposed writes)

// Programmatic dependent launch: block until the primary grid (the cascade prefix pass) has
// completed and its global writes are visible. A no-op when launched without the attribute.
__device__ __forceinline__ void grid_dependency_wait() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ void consumer_bar_sync() {
    asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory");
}

// Combine S normalised partials (fp32 O_s, log2-domain lse_s) of one (b, q-head) row into out:
// one warp per row, 4 dims per lane. lane s (and s + 32, s + 64) holds the weight of slot s; the
// partial loads of 8 slots are issued together so their latencies overlap (S <= 64 + 32 slots).
// Loads go through L2 (ld.global.cg): when the last-arriving split CTA merges in the same launch,
// the partials were written by other SMs a moment ago.
__device__ __forceinline__ void merge_row(const float* __restrict__ ws_o,
                                          const float* __restrict__ ws_lse,
                                          __nv_bfloat16* __restrict__ out, int64_t row, int S,
                                          int St, int lane) {
    const float* lse = ws_lse + row * St;
    float my_lse[3], my_w[3];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        my_lse[i] = lane + 32 * i < S ? __ldcg(lse + lane + 32 * i) : -INFINITY;
        mx = fmaxf(mx, my_lse[i]);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    float wsum = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        my_w[i] = my_lse[i] == -INFINITY ? 0.f : fast_exp2(my_lse[i] - mx);  // empty: weight 0
        wsum += my_w[i];
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, off);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4* po = reinterpret_cast<const float4*>(ws_o + row * St * kHeadDim) + lane;
    for (int s0 = 0; s0 < S; s0 += 8) {
        float4 x[8];
        float w[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int sl = s0 + u;
            const float wa = __shfl_sync(0xffffffffu, my_w[0], sl & 31);
            const float wb = __shfl_sync(0xffffffffu, my_w[1], sl & 31);
            const float wc = __shfl_sync(0xffffffffu, my_w[2], sl & 31);
            w[u] = sl < S ? (sl < 32 ? wa : (sl < 64 ? wb : wc)) : 0.f;
            // an empty slot (weight 0) was never written: do not read it
            x[u] = w[u] != 0.f ? __ldcg(po + sl * (kHeadDim / 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            acc.x += w[u] * x[u].x;
            acc.y += w[u] * x[u].y;
            acc.z += w[u] * x[u].z;
            acc.w += w[u] * x[u].w;
        }
    }
    const float inv = wsum > 0.f ? 1.f / wsum : 0.f;
    uint2 pk;
    pk.x = pack_bf16x2(acc.x * inv, acc.y * inv);
    pk.y = pack_bf16x2(acc.z * inv, acc.w * inv);
    *reinterpret_cast<uint2*>(out + row * kHeadDim + lane * 4) = pk;
}

// In-launch merge (DecodeParams::split_counters): every split CTA of a (sequence, KV head) calls
// this after its partials are written (all `n_threads` threads of barrier `bar_id`). Returns true
// in the CTA that arrived last - that one merges. Classic fence / atomic pattern: writers fence
// before the counter increment, the last arriver fences after it; the counter is left at zero for
// the next launch.
__device__ __forceinline__ bool split_arrive_is_last(int32_t* counter, int n_splits, int bar_id,
                                                     int n_threads, int* flag_smem) {
    __threadfence();
    asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(n_threads) : "memory");
    if (threadIdx.x == 0) {
        const int old = atomicAdd(counter, 1);
        const int last = old == n_splits - 1;
        if (last) *counter = 0;
        *flag_smem = last;
    }
    asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(n_threads) : "memory");
    const bool last = *flag_smem != 0;
    if (last) __threadfence();
    return last;
}

template <int kStages>
__global__ void __launch_bounds__(kDecodeThreads, 2)
decode_split_kernel(const __grid_constant__ CUtensorMap k_map,
                    const __grid_constant__ CUtensorMap v_map, const DecodeParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
    uint64_t* empty_bar = full_bar + kStages;
    float* ml_smem = reinterpret_cast<float*>(empty_bar + kStages);  // [4 warps][8 heads][2]
    __shared__ int last_flag;

    const int kvh = blockIdx.x;
    const int split = blockIdx.y;
    const int b = blockIdx.z;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int G = p.group;
    const int Hq = p.num_q_heads;
    const int S = p.num_splits;

    const int ctx = p.context_lens[b];
    const int tiles_total = ctx > 0 ? (ctx + kTile - 1) / kTile : 0;
    // Shared-prefix cascade: the first P pages were attended by the prefix kernel (its partials
    // sit in workspace slots [S, S + chunks)); this kernel streams the remainder only. P pages
    // lie entirely before token ctx-1 of every active row, so the remainder is never empty.
    const int shared_pages = (p.plan != nullptr && ctx > 0) ? p.plan[0] : 0;
    const int t_skip = shared_pages * (p.block_size / kTile);
    const int per_split = (tiles_total - t_skip + S - 1) / S;
    const int t_begin = t_skip + split * per_split;
    const int t_end = min(t_begin + per_split, tiles_total);
    const int St = S + p.prefix_chunks;  // workspace slots per (b, q-head) row
    const int64_t q_row0 = static_cast<int64_t>(b) * Hq + static_cast<int64_t>(kvh) * G;

    if (t_begin >= t_end) {
        // Nothing to attend in this split (ctx == 0 padding row of a CUDA-graph batch, or a split
        // beyond the sequence). No table entry is read.
        if (S == 1) {
            __nv_bfloat16* out = static_cast<__nv_bfloat16*>(p.out) + q_row0 * kHeadDim;
            for (int i = threadIdx.x; i < G * kHeadDim; i += kDecodeThreads)
                out[i] = __float2bfloat16(0.f);
        } else {
            if (threadIdx.x < G) p.ws_lse[(q_row0 + threadIdx.x) * St + split] = -INFINITY;
            if (p.split_counters != nullptr &&
                split_arrive_is_last(p.split_counters + b * p.num_kv_heads + kvh, S, 0,
                                     kDecodeThreads, &last_flag)) {
                // this empty split arrived last: it merges the row group
                const int n_pre = shared_pages > 0 ? prefix_chunks_of(shared_pages, p.block_size,
                                                                      p.prefix_chunks).valid
                                                   : 0;
                if (n_pre > 0) grid_dependency_wait();
                for (int h = warp; h < G; h += kConsumerWarps + 1)
                    merge_row(p.ws_o, p.ws_lse, static_cast<__nv_bfloat16*>(p.out), q_row0 + h,
                              S + n_pre, St, lane);
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

    if (warp == kConsumerWarps) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            tma_prefetch_desc(&k_map);
            tma_prefetch_desc(&v_map);
        }
        const int tiles_per_page = p.block_size / kTile;
        const int32_t* table = p.block_tables + static_cast<int64_t>(b) * p.table_row_stride;
        const int last_page = (t_end - 1) / tiles_per_page;
        int page_base = t_begin / tiles_per_page;
        int my_block = (page_base + lane <= last_page) ? table[page_base + lane] : 0;
        const uint64_t policy = l2_policy_evict_first();
        // Fused scatter (reference attention.py:70-71): the CTA whose split ends at the newest
        // token writes this step's K/V row of its KV head into the cache slot, then orders the
        // generic-proxy stores before its own TMA reads of that slot (fence.proxy.async). Other
        // CTAs never read the slot: it is token ctx-1 of sequence b, private to this split.
        if (p.new_key != nullptr && t_end == tiles_total) {
            const int64_t slot = p.slot_mapping[b];
            const bool ok = slot >= 0 && slot < p.num_slots;  // warp-uniform
            if (ok) {
                const bool is_v = lane >= 16;
                const int c = lane & 15;  // 16-byte chunk of the 256-byte head row
                const uint8_t* src =
                    static_cast<const uint8_t*>(is_v ? p.new_value : p.new_key) +
                    static_cast<int64_t>(b) * (is_v ? p.value_row_stride : p.key_row_stride) * 2 +
                    kvh * 256 + c * 16;
                uint8_t* dst = static_cast<uint8_t*>(is_v ? p.v_cache : p.k_cache) +
                               (slot * p.num_kv_heads + kvh) * 256 + c * 16;
                int4 row = ld_nc_v4(src);
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
        for (int t = t_begin, it = 0; t < t_end; ++t, ++it) {
            const int stage = it % kStages;
            const uint32_t phase = (it / kStages) & 1u;
            const int page = t / tiles_per_page;
            if (page - page_base >= 32) {
                page_base = page;
                my_block = (page_base + lane <= last_page) ? table[page_base + lane] : 0;
            }
            const int block_id = __shfl_sync(0xffffffffu, my_block, page - page_base);
            if (lane == 0) {
                mbar_wait(&empty_bar[stage], phase ^ 1u);
                mbar_expect_tx(&full_bar[stage], kStageBytes);
                const int slot = block_id * p.block_size + (t % tiles_per_page) * kTile;
                uint8_t* dst = smem + stage * kStageBytes;
                tma_load_3d_hint(dst, &k_map, &full_bar[stage], 0, kvh, slot, policy);
                tma_load_3d_hint(dst + kSubTileBytes, &k_map, &full_bar[stage], 64, kvh, slot,
                                 policy);
                tma_load_3d_hint(dst + 2 * kSubTileBytes, &v_map, &full_bar[stage], 0, kvh, slot,
                                 policy);
                tma_load_3d_hint(dst + 3 * kSubTileBytes, &v_map, &full_bar[stage], 64, kvh, slot,
                                 policy);
            }
            __syncwarp();
        }
        return;
    }

    // ===================== consumers =====================
    // Q^T as the B operand (k = dim, n = head): b0 holds dims 2*(lane%4)+{0,1}, b1 the same +8,
    // for head n = lane/4. Heads >= G are zero.
    uint32_t qf[8][2];
    {
        const int head = lane >> 2;
        const int hl = head < G ? head : 0;  // heads >= G: load row 0 (convergent), zeroed below
        const __nv_bfloat16* qrow = static_cast<const __nv_bfloat16*>(p.q) +
                                    static_cast<int64_t>(b) * p.q_row_stride +
                                    (static_cast<int64_t>(kvh) * G + hl) * kHeadDim +
                                    2 * (lane & 3);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            qf[ks][0] = *reinterpret_cast<const uint32_t*>(qrow + ks * 16);
            qf[ks][1] = *reinterpret_cast<const uint32_t*>(qrow + ks * 16 + 8);
        }
        if (p.nr.positions != nullptr) {
            // Fused q_norm + RoPE (norm_rope.cuh) in the fragment layout: the 4 lanes of a head
            // hold its whole row, lane (lane&3) the dims ks*16 + 2*(lane&3) + {0, 1, 8, 9}; the
            // rotation partner (dim +- 64) is register ks +- 4 of the same thread.
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

    float m_run[2] = {-INFINITY, -INFINITY};  // running max (log2 domain), heads 2*(lane%4)+{0,1}
    float l_run[2] = {0.f, 0.f};              // per-thread partial row sums
    float o[8][4];
#pragma unroll
    for (int mt = 0; mt < 8; ++mt)
#pragma unroll
        for (int c = 0; c < 4; ++c) o[mt][c] = 0.f;

    const uint32_t smem_base = smem_u32(smem);
    // K A-fragment addressing (rows = tokens): lane -> row lane%16, 16-byte chunk +(lane/16)
    const uint32_t k_row = warp * 16 + (lane & 15);
    const uint32_t k_csel = lane >> 4;
    // V^T A-fragment addressing via ldmatrix.trans: matrix j = lane/8 -> tokens (j/2)*8 + lane%8,
    // dim chunk + (j%2)
    const uint32_t v_row = warp * 16 + ((lane >> 4) << 3) + (lane & 7);
    const uint32_t v_csel = (lane >> 3) & 1;

    for (int t = t_begin, it = 0; t < t_end; ++t, ++it) {
        const int stage = it % kStages;
        const uint32_t phase = (it / kStages) & 1u;
        mbar_wait(&full_bar[stage], phase);
        const uint32_t kb = smem_base + stage * kStageBytes;
        const uint32_t vb = kb + 2 * kSubTileBytes;

        // ---- S^T[16 tok x 8 heads] = K . Q^T ----
        float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            uint32_t a[4];
            ldmatrix_x4(a, kb + (ks >> 2) * kSubTileBytes + swz128(k_row, (ks & 3) * 2 + k_csel));
            mma_bf16_16816(s, a, qf[ks][0], qf[ks][1]);
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

        // Tokens beyond ctx inside the last page are whatever the allocator left there; P is 0
        // for them but 0 * NaN would poison O, so their V rows are cleared in shared memory.
        if (t == tiles_total - 1) {
            const int valid_rows = ctx - (t * kTile + warp * 16);  // rows of this warp's 16
            if (valid_rows < 16) {
                const int r_lo = max(valid_rows, 0);
                if (lane < 16) {
                    uint8_t* vsm = smem + stage * kStageBytes + 2 * kSubTileBytes +
                                   (lane >> 3) * kSubTileBytes;
                    for (int r = r_lo; r < 16; ++r)
                        *reinterpret_cast<int4*>(vsm + (warp * 16 + r) * 128 + (lane & 7) * 16) =
                            make_int4(0, 0, 0, 0);
                }
                __syncwarp();
            }
        }

        // ---- O^T[128 dims x 8 heads] = alpha * O^T + V^T . P^T ----
        const uint32_t pb0 = movmatrix_trans(pack_bf16x2(p0, p1));
        const uint32_t pb1 = movmatrix_trans(pack_bf16x2(p2, p3));
#pragma unroll
        for (int mt = 0; mt < 8; ++mt) {
            uint32_t a[4];
            ldmatrix_x4_trans(
                a, vb + (mt >> 2) * kSubTileBytes + swz128(v_row, (mt & 3) * 2 + v_csel));
            o[mt][0] *= alpha0;
            o[mt][1] *= alpha1;
            o[mt][2] *= alpha0;
            o[mt][3] *= alpha1;
            mma_bf16_16816(o[mt], a, pb0, pb1);
        }

        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[stage]);
    }

    // ---- finish the row sums (tokens are spread over lane/4) ----
#pragma unroll
    for (int off = 4; off < 32; off <<= 1) {
        l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], off);
        l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], off);
    }

    // ---- merge the 4 warps through shared memory (ring is drained; reuse stage 0) ----
    if (lane < 4) {
        float* ml = ml_smem + (warp * 8 + 2 * lane) * 2;
        ml[0] = m_run[0];
        ml[1] = l_run[0];
        ml[2] = m_run[1];
        ml[3] = l_run[1];
    }
    consumer_bar_sync();  // all warps done with the ring + ml visible
    float mg[2], fac[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int h = 2 * (lane & 3) + i;
        float m = -INFINITY;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) m = fmaxf(m, ml_smem[(w * 8 + h) * 2]);
        mg[i] = m;
        fac[i] = m == -INFINITY ? 0.f : fast_exp2(m_run[i] - m);
    }
    float* osc = reinterpret_cast<float*>(smem);  // [4 warps][8 heads][kMergeStride]
#pragma unroll
    for (int mt = 0; mt < 8; ++mt)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int h = 2 * (lane & 3) + (c & 1);
            const int d = mt * 16 + (lane >> 2) + (c >> 1) * 8;
            osc[(warp * 8 + h) * kMergeStride + d] = o[mt][c] * fac[c & 1];
        }
    consumer_bar_sync();

    // Final reduction: warp w owns heads w, w+4; lane owns dims 4*lane..4*lane+3 (float4 smem
    // reads, 8-byte bf16 / 16-byte fp32 global stores, a head row is one coalesced line).
    const int n_prefix =
        shared_pages > 0 ? prefix_chunks_of(shared_pages, p.block_size, p.prefix_chunks).valid : 0;
    for (int h = warp; h < G; h += kConsumerWarps) {
        float m = -INFINITY;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) m = fmaxf(m, ml_smem[(w * 8 + h) * 2]);
        float l = 0.f;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) {
            const float mw = ml_smem[(w * 8 + h) * 2];
            l += mw == -INFINITY ? 0.f : ml_smem[(w * 8 + h) * 2 + 1] * fast_exp2(mw - m);
            const float4 x =
                *reinterpret_cast<const float4*>(&osc[(w * 8 + h) * kMergeStride + 4 * lane]);
            acc.x += x.x;
            acc.y += x.y;
            acc.z += x.z;
            acc.w += x.w;
        }
        const int64_t row = q_row0 + h;
        if (S == 1) {
            if (n_prefix > 0) {
                grid_dependency_wait();  // the prefix pass (PDL primary) has written its slots
                // fold in the shared-prefix partials (normalised O_c, lse_c) of this row:
                // out = (acc 2^(m-M) + sum_c O_c 2^(lse_c-M)) / (l 2^(m-M) + sum_c 2^(lse_c-M))
                // lane c holds lse_c; the O_c loads are independent and unrolled so they overlap.
                const float my_lse = lane < n_prefix ? p.ws_lse[row * St + 1 + lane] : -INFINITY;
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
                const float4* o_c =
                    reinterpret_cast<const float4*>(p.ws_o + (row * St + 1) * kHeadDim) + lane;
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
            pk.x = pack_bf16x2(live ? acc.x / l : 0.f, live ? acc.y / l : 0.f);
            pk.y = pack_bf16x2(live ? acc.z / l : 0.f, live ? acc.w / l : 0.f);
            *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(p.out) + row * kHeadDim +
                                      4 * lane) = pk;
        } else {
            const bool live = l > 0.f;
            acc.x = live ? acc.x / l : 0.f;
            acc.y = live ? acc.y / l : 0.f;
            acc.z = live ? acc.z / l : 0.f;
            acc.w = live ? acc.w / l : 0.f;
            *reinterpret_cast<float4*>(p.ws_o + (row * St + split) * kHeadDim + 4 * lane) = acc;
            if (lane == 0) p.ws_lse[row * St + split] = l > 0.f ? m + log2f(l) : -INFINITY;
        }
    }
    (void)mg;
    if (S > 1 && p.split_counters != nullptr &&
        split_arrive_is_last(p.split_counters + b * p.num_kv_heads + kvh, S, 1,
                             kConsumerWarps * 32, &last_flag)) {
        // last split of this (sequence, KV head) to finish: merge its G rows here, no merge launch
        if (n_prefix > 0) grid_dependency_wait();
        for (int h = warp; h < G; h += kConsumerWarps)
            merge_row(p.ws_o, p.ws_lse, static_cast<__nv_bfloat16*>(p.out), q_row0 + h,
                      S + n_prefix, St, lane);
    }
    // With the cascade armed this grid is a programmatic dependent of the prefix pass: make sure
    // it cannot finish before its primary did (a merge kernel that follows reads both).
    if (p.plan != nullptr) grid_dependency_wait();
}

// Combine num_splits normalised partials per (b, q-head) row: one warp per row, 4 dims per lane.
// Slots [0, S_split) hold the split partials; with the cascade, slots [S_split, S_split + valid
// prefix chunks) hold the shared-prefix partials of rows with context_lens > 0 (the other prefix
// slots are never written and never read).
__global__ void __launch_bounds__(128)
decode_merge_kernel(const float* __restrict__ ws_o, const float* __restrict__ ws_lse,
                    __nv_bfloat16* __restrict__ out, int32_t rows, int32_t S_split, int32_t St,
                    int32_t num_q_heads, const int32_t* __restrict__ context_lens,
                    const int32_t* __restrict__ plan, int32_t prefix_chunks, int32_t block_size) {
    const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    int S = S_split;
    if (plan != nullptr && context_lens[row / num_q_heads] > 0)
        S += prefix_chunks_of(plan[0], block_size, prefix_chunks).valid;
    merge_row(ws_o, ws_lse, out, row, S, St, lane);
}

template <int kStages>
constexpr int decode_smem_bytes() {
    return kStages * kStageBytes + 2 * kStages * 8 + kConsumerWarps * 8 * 2 * 4 + 1024;
}

template <int kStages>
cudaError_t launch_split(const DecodeParams& p, const CUtensorMap& k_map, const CUtensorMap& v_map,
                         cudaStream_t stream) {
    static bool configured = false;
    constexpr int smem = decode_smem_bytes<kStages>();
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(decode_split_kernel<kStages>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    dim3 grid(p.num_kv_heads, p.num_splits, p.batch);
    if (p.plan != nullptr) {
        // Programmatic dependent launch behind the prefix pass: this grid may start while the
        // prefix CTAs are still running (they signal launch_dependents at once) and only waits
        // (griddepcontrol.wait) where it consumes their partials. It touches nothing the prefix
        // pass reads or writes before that point: the fused scatter writes token ctx-1, which
        // lies after the shared pages, and the split partials go to other workspace slots.
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = grid;
        cfg.blockDim = dim3(kDecodeThreads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, decode_split_kernel<kStages>, k_map, v_map, p);
        count_launch();
        return e != cudaSuccess ? e : cudaGetLastError();
    }
    decode_split_kernel<kStages><<<grid, kDecodeThreads, smem, stream>>>(k_map, v_map, p);
    count_launch();
    return cudaGetLastError();
}

int decode_stages_from_env() {
    static int stages = [] {
        const char* e = std::getenv("B200ATTN_DECODE_STAGES");
        int s = e ? std::atoi(e) : 3;
        return (s == 2 || s == 3 || s == 4 || s == 6) ? s : 3;
    }();
    return stages;
}

}  // namespace

size_t decode_workspace_bytes(int32_t batch, int32_t num_q_heads, int32_t num_splits) {
    if (num_splits <= 1) return 0;
    const size_t rows = static_cast<size_t>(batch) * num_q_heads;
    return rows * num_splits * (kHeadDim + 1) * sizeof(float);
}

int decode_pick_splits(int32_t batch, int32_t num_kv_heads, int32_t table_width, int32_t block_size,
                       int32_t sm_count) {
    // Host-visible sizes only (graph-capture rule). Calibrated with tools/sweep_splits.py on B200
    // (profiles/r01_sweep_splits.log): one streaming CTA per SM already saturates HBM, so splitting
    // only pays while (sequence, KV head) items are fewer than ~0.7 x SMs; then aim for about one
    // CTA per SM and never cut the longest possible context into pieces shorter than 8 tiles
    // (512 tokens) - below that the split partials + merge cost more than they hide.
    const int64_t base = static_cast<int64_t>(batch) * num_kv_heads;
    if (base <= 0 || base * 10 >= static_cast<int64_t>(sm_count) * 7) return 1;
    const int64_t max_tiles = static_cast<int64_t>(table_width) * block_size / kTile;
    int64_t s = (sm_count + base / 2) / base;  // round(sm_count / base)
    const int64_t cap = max_tiles / 8 > 1 ? max_tiles / 8 : 1;
    if (s > cap) s = cap;
    if (s > 64) s = 64;
    return static_cast<int>(s < 1 ? 1 : s);
}

cudaError_t launch_decode(const DecodeParams& p, const CUtensorMap& k_map, const CUtensorMap& v_map,
                          cudaStream_t stream) {
    if (p.batch == 0) return cudaSuccess;
    cudaError_t e;
    switch (decode_stages_from_env()) {
        case 2: e = launch_split<2>(p, k_map, v_map, stream); break;
        case 4: e = launch_split<4>(p, k_map, v_map, stream); break;
        case 6: e = launch_split<6>(p, k_map, v_map, stream); break;
        default: e = launch_split<3>(p, k_map, v_map, stream); break;
    }
    if (e != cudaSuccess) return e;
    if (p.num_splits > 1 && p.split_counters == nullptr) {
        const int rows = p.batch * p.num_q_heads;
        decode_merge_kernel<<<(rows + 3) / 4, 128, 0, stream>>>(
            p.ws_o, p.ws_lse, static_cast<__nv_bfloat16*>(p.out), rows, p.num_splits,
            p.num_splits + p.prefix_chunks, p.num_q_heads, p.context_lens, p.plan,
            p.prefix_chunks, p.block_size);
        count_launch();
        e = cudaGetLastError();
    }
    return e;
}

}  // namespace b200
