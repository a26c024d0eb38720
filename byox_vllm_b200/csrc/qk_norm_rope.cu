// Fused q/k RMSNorm + RoPE + KV-cache scatter on the raw output of the QKV projection
// (SURVEY 8f-4). Replaces four reference launches between qkv_proj and the attention call,
// nanovllm/models/qwen3.py:75-84:
//     q_norm, k_norm      layers/layernorm.py:16-26   (two torch.compile'd kernels)
//     rotary_emb          layers/rotary_embedding.py:37-48
//     store_kvcache       layers/attention.py:10-49   (Triton)
// with ONE pass over the [n_tokens, (Hq + 2 Hkv) * 128] bf16 rows: every head row (256 B) is read
// once; q rows are normalised + rotated into a contiguous q_out; k rows are normalised + rotated
// and written to their cache slot (and, for dense prefill, back in place so the prefill kernel can
// read them through its strided tensor map); v rows are copied to their cache slot.
//
// HBM bound. A head row is handled by a half-warp (16 lanes x 16 B = one coalesced 256-byte
// line); each half-warp takes kRows consecutive heads of one token with all loads issued up
// front. The q/k heads (arithmetic) and the v heads (pure copy) go to different warps of the grid.
// Algorithmic bytes per token:
//   (Hq + 2 Hkv) * 256 read + Hq * 256 (q_out) + [Hkv * 256 k in place] + 2 Hkv * 256 (cache)
//   + 8 (position) + 4 (slot) + 512 (cos/sin row, L2 resident across heads)
#include "common.h"
#include "norm_rope.cuh"
#include "ptx.cuh"

#include <cstdlib>

namespace b200 {

namespace {

constexpr int kRows = 4;       // head rows per half-warp
constexpr int kThreads = 256;  // 16 half-warps

__device__ __forceinline__ uint4 ld_v4(const void* p) {
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_u4(void* p, const uint4& v) {
    asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}

// Work item of a half-warp = kRows consecutive heads of ONE token (head groups are padded per
// token), so position, cos/sin row, slot and the row base pointers are fetched once per item.
// Blocks [0, qk_blocks) take the q and k heads (norm + RoPE; every half-warp runs the same code,
// so the shuffles are full-mask and convergent); the remaining blocks copy the v heads into the
// cache.
template <bool kHasNorm>
__global__ void __launch_bounds__(kThreads, 4)
qk_norm_rope_store_kernel(const QkNormRopeParams p, uint32_t qk_blocks, uint32_t qk_groups,
                          uint32_t v_groups) {
    const int c = threadIdx.x & 15;
    const uint32_t Hq = p.num_q_heads, Hkv = p.num_kv_heads, Hqk = Hq + Hkv;
    uint4 raw[kRows];

    if (blockIdx.x >= qk_blocks) {  // block-uniform: the shuffles below stay provably convergent
        // ---- v heads: bit-exact copy into the cache slot of the token ----
        const uint32_t item = ((blockIdx.x - qk_blocks) * kThreads + threadIdx.x) >> 4;
        const uint32_t tok = item / v_groups;
        if (tok >= static_cast<uint32_t>(p.n_tokens)) return;
        const uint32_t h0 = (item - tok * v_groups) * kRows;
        const int64_t slot = p.slot_mapping[tok];
        if (slot < 0 || slot >= p.num_slots) return;
        const uint8_t* src = static_cast<const uint8_t*>(p.qkv) +
                             static_cast<int64_t>(tok) * p.qkv_row_stride * 2 +
                             (Hqk + h0) * 256 + c * 16;
        uint8_t* dst = static_cast<uint8_t*>(p.v_cache) + (slot * Hkv + h0) * 256 + c * 16;
#pragma unroll
        for (int u = 0; u < kRows; ++u)
            if (h0 + u < Hkv) raw[u] = ld_v4(src + u * 256);
#pragma unroll
        for (int u = 0; u < kRows; ++u)
            if (h0 + u < Hkv) st_u4(dst + u * 256, raw[u]);
        return;
    }

    // ---- q and k heads ----
    constexpr bool has_norm = kHasNorm;
    const uint32_t item = (blockIdx.x * kThreads + threadIdx.x) >> 4;
    uint32_t tok = item / qk_groups;
    const uint32_t h0 = (item - tok * qk_groups) * kRows;
    const bool live = tok < static_cast<uint32_t>(p.n_tokens);
    if (!live) tok = 0;  // the other half of the warp still needs this half in its shuffles
    // head of the only dependent load chain (position -> cos/sin row): issue it first
    const float* cs_row = cos_sin_row(p.nr, tok);
    const uint8_t* src = static_cast<const uint8_t*>(p.qkv) +
                         static_cast<int64_t>(tok) * p.qkv_row_stride * 2 + c * 16;
#pragma unroll
    for (int u = 0; u < kRows; ++u) {
        const uint32_t h = min(h0 + u, Hqk - 1);  // clamp: loads stay legal
        raw[u] = ld_v4(src + h * 256);
    }
    uint4 wq = make_uint4(0, 0, 0, 0), wk = wq;
    if (has_norm) {
        wq = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(p.nr.q_weight) + c * 16);
        wk = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(p.nr.k_weight) + c * 16);
    }
    const CosSin8 cs = load_cos_sin8(cs_row, c);
    int64_t slot = p.slot_mapping != nullptr ? p.slot_mapping[tok] : -1;
    if (slot >= p.num_slots) slot = -1;
    uint8_t* q_dst = static_cast<uint8_t*>(p.q_out) + static_cast<int64_t>(tok) * Hq * 256 + c * 16;
    uint8_t* k_dst = p.k_out == nullptr
                         ? nullptr
                         : static_cast<uint8_t*>(p.k_out) +
                               static_cast<int64_t>(tok) * p.k_out_row_stride * 2 + c * 16;
    uint8_t* kc_dst =
        slot < 0 ? nullptr : static_cast<uint8_t*>(p.k_cache) + slot * Hkv * 256 + c * 16;
    // all arithmetic (and its shuffles) first, with the warp converged; the predicated stores after
    uint4 rot[kRows];
#pragma unroll
    for (int u = 0; u < kRows; ++u)
        rot[u] = norm_rope_row16(raw[u], h0 + u < Hq ? wq : wk, has_norm, p.nr.eps, cs, c,
                                 0xffffffffu);
#pragma unroll
    for (int u = 0; u < kRows; ++u) {
        const uint32_t h = h0 + u;
        if (live && h < Hqk) {
            if (h < Hq) {
                st_u4(q_dst + h * 256, rot[u]);
            } else {
                if (k_dst != nullptr) st_u4(k_dst + (h - Hq) * 256, rot[u]);
                if (kc_dst != nullptr) st_u4(kc_dst + (h - Hq) * 256, rot[u]);
            }
        }
    }
}


// ---------------------------------------------------------------------------------------------
// Narrow rows (tensor-parallel shards: Hq + 2 Hkv <= kTokenRows heads per token, e.g. 8/1 at
// TP = 8): a half-warp takes ONE WHOLE TOKEN - every q, k and v head row, all loads issued up
// front (2.5 KiB in flight per half-warp at 8/1), position / slot / cos-sin row fetched once per
// token instead of once per group of 4 heads, no idle lanes in a partly filled head group and no
// separate v blocks. Same arithmetic (norm_rope_row16), bit-identical results.
constexpr int kTokenRows = 12;

template <bool kHasNorm>
__global__ void __launch_bounds__(kThreads, 2)
qk_norm_rope_token_kernel(const QkNormRopeParams p) {
    const int c = threadIdx.x & 15;
    const uint32_t Hq = p.num_q_heads, Hkv = p.num_kv_heads, Hqk = Hq + Hkv, R = Hqk + Hkv;
    uint32_t tok = (blockIdx.x * kThreads + threadIdx.x) >> 4;
    const bool live = tok < static_cast<uint32_t>(p.n_tokens);
    if (!live) tok = 0;  // the other half of the warp still needs this half in its shuffles
    const float* cs_row = cos_sin_row(p.nr, tok);  // head of the only dependent load chain
    const uint8_t* src = static_cast<const uint8_t*>(p.qkv) +
                         static_cast<int64_t>(tok) * p.qkv_row_stride * 2 + c * 16;
    uint4 raw[kTokenRows];
#pragma unroll
    for (int u = 0; u < kTokenRows; ++u)
        if (u < R) raw[u] = ld_v4(src + u * 256);
    uint4 wq = make_uint4(0, 0, 0, 0), wk = wq;
    if (kHasNorm) {
        wq = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(p.nr.q_weight) + c * 16);
        wk = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(p.nr.k_weight) + c * 16);
    }
    const CosSin8 cs = load_cos_sin8(cs_row, c);
    int64_t slot = p.slot_mapping != nullptr ? p.slot_mapping[tok] : -1;
    if (slot >= p.num_slots) slot = -1;
    uint8_t* q_dst = static_cast<uint8_t*>(p.q_out) + static_cast<int64_t>(tok) * Hq * 256 + c * 16;
    uint8_t* k_dst = p.k_out == nullptr
                         ? nullptr
                         : static_cast<uint8_t*>(p.k_out) +
                               static_cast<int64_t>(tok) * p.k_out_row_stride * 2 + c * 16;
    uint8_t* kc_dst =
        slot < 0 ? nullptr : static_cast<uint8_t*>(p.k_cache) + slot * Hkv * 256 + c * 16;
    uint8_t* vc_dst =
        slot < 0 ? nullptr : static_cast<uint8_t*>(p.v_cache) + slot * Hkv * 256 + c * 16;
    // arithmetic on the q / k rows with the warp converged (R, Hq, Hkv are launch-uniform)
#pragma unroll
    for (int u = 0; u < kTokenRows; ++u)
        if (u < Hqk)
            raw[u] = norm_rope_row16(raw[u], u < Hq ? wq : wk, kHasNorm, p.nr.eps, cs, c,
                                     0xffffffffu);
    if (!live) return;
#pragma unroll
    for (int u = 0; u < kTokenRows; ++u) {
        if (u < Hq) {
            st_u4(q_dst + u * 256, raw[u]);
        } else if (u < Hqk) {
            if (k_dst != nullptr) st_u4(k_dst + (u - Hq) * 256, raw[u]);
            if (kc_dst != nullptr) st_u4(kc_dst + (u - Hq) * 256, raw[u]);
        } else if (u < R) {
            if (vc_dst != nullptr) st_u4(vc_dst + (u - Hqk) * 256, raw[u]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Bulk-staged variant (default for prefill-size launches): the bytes in flight are bounded by
// shared memory instead of registers. A persistent CTA walks tokens t = blockIdx.x, + gridDim.x,
// ...; a producer thread runs kBulkStages tokens ahead: it reads the token's position and slot,
// then issues two 1-D bulk copies through the TMA engine (cp.async.bulk, mbarrier completion):
// the whole QKV row (Hq + 2 Hkv) * 256 B and the 512-byte cos/sin row of its position. Four
// consumer warps normalise + rotate the q and k heads IN shared memory (a head row per half-warp,
// 16 B per lane, conflict-free), then one thread writes everything back with bulk stores straight
// from shared memory: q heads -> q_out, k heads -> k_out (in place) and the cache slot, v heads ->
// the cache slot (never touched by a compute thread). A stage is handed back to the producer when
// the bulk stores of the PREVIOUS token have finished reading it (cp.async.bulk.wait_group.read
// 1). Measured and not adopted: several tokens per stage for narrow rows (one barrier round per
// ~16 KiB; the extra index arithmetic and producer registers cost more than the rounds save:
// 16/8 rows 61 -> 79 us, 8/1 rows 29 -> 51 us, so narrow rows stay on the direct kernel); 6 / 8
// consumer warps (more registers, fewer CTAs per SM: slower)
// and a separate store thread behind a `done` mbarrier (a second spinning warp per CTA: 16/8 rows
// 62 -> 72 us).
constexpr int kBulkStages = 4;

template <bool kHasNorm, int kBulkConsumerWarps>
__global__ void __launch_bounds__((kBulkConsumerWarps + 1) * 32)
qk_norm_rope_bulk_kernel(const QkNormRopeParams p, uint32_t stage_bytes) {
    extern __shared__ uint8_t bulk_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(bulk_smem_raw) + 127) & ~static_cast<uintptr_t>(127));
    // stage layout: [row: Htot * 256 B][cos/sin: 512 B]; after the stages: barriers, slots
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kBulkStages * stage_bytes);
    uint64_t* empty_bar = full_bar + kBulkStages;
    int64_t* slot_smem = reinterpret_cast<int64_t*>(empty_bar + kBulkStages);

    const uint32_t Hq = p.num_q_heads, Hkv = p.num_kv_heads, Hqk = Hq + Hkv;
    const uint32_t row_bytes = (Hq + 2 * Hkv) * 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tokens = p.n_tokens;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kBulkStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == kBulkConsumerWarps) {
        // ===================== producer (one thread) =====================
        if (lane != 0) return;
        int it = 0;
        for (int tok = blockIdx.x; tok < n_tokens; tok += gridDim.x, ++it) {
            const int stage = it % kBulkStages;
            const uint32_t phase = (it / kBulkStages) & 1u;
            const float* cs_row = cos_sin_row(p.nr, tok);  // dependent load, hidden by the ring
            int64_t slot = p.slot_mapping != nullptr ? p.slot_mapping[tok] : -1;
            if (slot >= p.num_slots) slot = -1;
            mbar_wait(&empty_bar[stage], phase ^ 1u);
            uint8_t* dst = smem + stage * stage_bytes;
            slot_smem[stage] = slot;
            mbar_expect_tx(&full_bar[stage], row_bytes + 512);
            bulk_load_1d(dst, static_cast<const uint8_t*>(p.qkv) +
                                  static_cast<int64_t>(tok) * p.qkv_row_stride * 2,
                         row_bytes, &full_bar[stage]);
            bulk_load_1d(dst + row_bytes, cs_row, 512, &full_bar[stage]);
        }
        return;
    }

    // ===================== consumers =====================
    const int c = lane & 15, half = lane >> 4;
    uint4 wq = make_uint4(0, 0, 0, 0), wk = wq;
    if (kHasNorm) {
        wq = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(p.nr.q_weight) + c * 16);
        wk = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(p.nr.k_weight) + c * 16);
    }
    int it = 0;
    int prev_stage = -1;
    for (int tok = blockIdx.x; tok < n_tokens; tok += gridDim.x, ++it) {
        const int stage = it % kBulkStages;
        const uint32_t phase = (it / kBulkStages) & 1u;
        mbar_wait(&full_bar[stage], phase);
        uint8_t* row = smem + stage * stage_bytes;
        const CosSin8 cs = load_cos_sin8(reinterpret_cast<const float*>(row + row_bytes), c);
        // rows (2*warp + half), + 8, ...: both halves of a warp run the same trip count
        for (uint32_t base = 2 * warp; base < Hqk; base += 2 * kBulkConsumerWarps) {
            const uint32_t h = base + half;
            const bool valid = h < Hqk;
            uint4* cell = reinterpret_cast<uint4*>(row + (valid ? h : base) * 256 + c * 16);
            const uint4 rot = norm_rope_row16(*cell, h < Hq ? wq : wk, kHasNorm, p.nr.eps, cs, c,
                                              0xffffffffu);
            if (valid) *cell = rot;
        }
        fence_proxy_async();  // generic-proxy writes above -> visible to the bulk stores below
        asm volatile("bar.sync 1, %0;" ::"n"(kBulkConsumerWarps * 32) : "memory");
        if (threadIdx.x == 0) {
            const int64_t slot = slot_smem[stage];
            const int64_t t = tok;
            bulk_store_1d(static_cast<uint8_t*>(p.q_out) + t * Hq * 256, row, Hq * 256);
            if (p.k_out != nullptr)
                bulk_store_1d(static_cast<uint8_t*>(p.k_out) + t * p.k_out_row_stride * 2,
                              row + Hq * 256, Hkv * 256);
            if (slot >= 0) {
                bulk_store_1d(static_cast<uint8_t*>(p.k_cache) + slot * Hkv * 256, row + Hq * 256,
                              Hkv * 256);
                bulk_store_1d(static_cast<uint8_t*>(p.v_cache) + slot * Hkv * 256,
                              row + Hqk * 256, Hkv * 256);
            }
            bulk_commit_group();
            // the previous token's stores have finished READING their stage: hand it back
            bulk_wait_group_read<1>();
            if (prev_stage >= 0) mbar_arrive(&empty_bar[prev_stage]);
        }
        prev_stage = stage;
    }
    if (threadIdx.x == 0) bulk_wait_group<0>();  // shared memory must outlive the last stores
}

}  // namespace

// B200ATTN_NORMROPE=bulk|direct forces one variant (A/B runs); default: bulk-staged for
// prefill-size launches, the direct kernel for decode-size ones (latency bound, wide grid).
static int normrope_mode_from_env() {
    static int mode = [] {
        const char* e = std::getenv("B200ATTN_NORMROPE");
        if (!e) return 0;
        return e[0] == 'b' ? 1 : (e[0] == 'd' ? 2 : (e[0] == 'g' ? 3 : 0));
    }();
    return mode;
}

cudaError_t launch_qk_norm_rope_store(const QkNormRopeParams& p, cudaStream_t stream) {
    if (p.n_tokens == 0) return cudaSuccess;
    const int mode = normrope_mode_from_env();
    const uint32_t row_bytes = static_cast<uint32_t>(p.num_q_heads + 2 * p.num_kv_heads) * 256;
    const uint32_t stage_bytes = (row_bytes + 512 + 127) & ~127u;
    const int smem = kBulkStages * stage_bytes + 2 * kBulkStages * 8 + kBulkStages * 8 + 128;
    const bool bulk_ok = smem <= 200 * 1024;
    // rows of >= 8 KiB per token amortise the per-token barrier / issue chain; below that
    // (TP = 8 shards) and for short batches the direct kernel is faster (measured)
    if (bulk_ok && (mode == 1 || (mode == 0 && p.n_tokens >= 2048 && row_bytes >= 8192))) {
        int sm_count = current_device_sm_count();
        if (sm_count <= 0) sm_count = 148;
        static const int warps = [] {
            const char* e = std::getenv("B200ATTN_NORMROPE_WARPS");
            const int w = e ? std::atoi(e) : 4;
            return (w == 4 || w == 6 || w == 8) ? w : 4;
        }();
        const bool norm = p.nr.q_weight != nullptr;
        auto kern = warps == 4   ? (norm ? qk_norm_rope_bulk_kernel<true, 4>
                                         : qk_norm_rope_bulk_kernel<false, 4>)
                    : warps == 6 ? (norm ? qk_norm_rope_bulk_kernel<true, 6>
                                         : qk_norm_rope_bulk_kernel<false, 6>)
                                 : (norm ? qk_norm_rope_bulk_kernel<true, 8>
                                         : qk_norm_rope_bulk_kernel<false, 8>);
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             smem);
        if (e != cudaSuccess) return e;
        int per_sm = (220 * 1024) / (smem + 1024);
        per_sm = per_sm < 1 ? 1 : (per_sm > 6 ? 6 : per_sm);
        const int64_t want = static_cast<int64_t>(sm_count) * per_sm;
        const unsigned grid = static_cast<unsigned>(p.n_tokens < want ? p.n_tokens : want);
        kern<<<grid, (warps + 1) * 32, smem, stream>>>(p, stage_bytes);
        count_launch();
        return cudaGetLastError();
    }
    // B200ATTN_NORMROPE=groups keeps narrow rows on the head-group kernel below (A/B runs)
    if (p.num_q_heads + 2 * p.num_kv_heads <= kTokenRows && p.n_tokens >= 256 && mode != 3) {
        const unsigned grid =
            static_cast<unsigned>((static_cast<int64_t>(p.n_tokens) * 16 + kThreads - 1) / kThreads);
        if (p.nr.q_weight != nullptr)
            qk_norm_rope_token_kernel<true><<<grid, kThreads, 0, stream>>>(p);
        else
            qk_norm_rope_token_kernel<false><<<grid, kThreads, 0, stream>>>(p);
        count_launch();
        return cudaGetLastError();
    }
    const int64_t qk_groups = (p.num_q_heads + p.num_kv_heads + kRows - 1) / kRows;
    const int64_t v_groups = (p.num_kv_heads + kRows - 1) / kRows;
    const int64_t qk_items = static_cast<int64_t>(p.n_tokens) * qk_groups;
    const int64_t v_items = p.slot_mapping != nullptr ? p.n_tokens * v_groups : 0;
    constexpr int kItemsPerBlock = kThreads / 16;
    const int64_t qk_warps = (qk_items + kItemsPerBlock - 1) / kItemsPerBlock;  // blocks
    const int64_t v_blocks = (v_items + kItemsPerBlock - 1) / kItemsPerBlock;
    const unsigned grid = static_cast<unsigned>(qk_warps + v_blocks);
    if (p.nr.q_weight != nullptr)
        qk_norm_rope_store_kernel<true><<<grid, kThreads, 0, stream>>>(
            p, static_cast<uint32_t>(qk_warps), static_cast<uint32_t>(qk_groups),
            static_cast<uint32_t>(v_groups));
    else
        qk_norm_rope_store_kernel<false><<<grid, kThreads, 0, stream>>>(
            p, static_cast<uint32_t>(qk_warps), static_cast<uint32_t>(qk_groups),
            static_cast<uint32_t>(v_groups));
    count_launch();
    return cudaGetLastError();
}

}  // namespace b200
