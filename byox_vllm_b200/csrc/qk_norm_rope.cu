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

}  // namespace

cudaError_t launch_qk_norm_rope_store(const QkNormRopeParams& p, cudaStream_t stream) {
    if (p.n_tokens == 0) return cudaSuccess;
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
