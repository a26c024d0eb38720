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
// HBM bound. Work item = one head row handled by a half-warp (16 lanes x 16 B = one coalesced
// 256-byte line); each half-warp takes kRows consecutive rows with all loads issued up front.
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

__global__ void __launch_bounds__(kThreads)
qk_norm_rope_store_kernel(const QkNormRopeParams p) {
    const int c = threadIdx.x & 15;
    const int64_t hw = (static_cast<int64_t>(blockIdx.x) * kThreads + threadIdx.x) >> 4;
    const int Htot = p.num_q_heads + 2 * p.num_kv_heads;
    const int64_t total = static_cast<int64_t>(p.n_tokens) * Htot;
    const bool has_norm = p.nr.q_weight != nullptr;

    uint4 raw[kRows];
    int64_t tok[kRows];
    int head[kRows];
#pragma unroll
    for (int u = 0; u < kRows; ++u) {
        int64_t r = hw * kRows + u;
        if (r >= total) r = total - 1;  // clamp: loads stay legal, stores are predicated below
        tok[u] = r / Htot;
        head[u] = static_cast<int>(r - tok[u] * Htot);
        raw[u] = ld_v4(static_cast<const uint8_t*>(p.qkv) + tok[u] * p.qkv_row_stride * 2 +
                       head[u] * 256 + c * 16);
    }
#pragma unroll
    for (int u = 0; u < kRows; ++u) {
        const bool live = hw * kRows + u < total;
        const bool is_q = head[u] < p.num_q_heads;
        const bool is_k = !is_q && head[u] < p.num_q_heads + p.num_kv_heads;
        uint4 wv = make_uint4(0, 0, 0, 0);
        if (has_norm)
            wv = *reinterpret_cast<const uint4*>(
                static_cast<const uint8_t*>(is_q ? p.nr.q_weight : p.nr.k_weight) + c * 16);
        // computed for v rows too (result discarded): keeps the warp convergent for the shuffles
        const uint4 rot =
            norm_rope_row16(raw[u], wv, has_norm, p.nr.eps, cos_sin_row(p.nr, tok[u]), c);
        if (live && is_q) {
            st_u4(static_cast<uint8_t*>(p.q_out) +
                      (tok[u] * p.num_q_heads + head[u]) * 256 + c * 16,
                  rot);
        } else if (live) {
            const int kvh = head[u] - p.num_q_heads - (is_k ? 0 : p.num_kv_heads);
            if (is_k && p.k_out != nullptr)
                st_u4(static_cast<uint8_t*>(p.k_out) + tok[u] * p.k_out_row_stride * 2 +
                          kvh * 256 + c * 16,
                      rot);
            if (p.slot_mapping != nullptr) {
                const int64_t slot = p.slot_mapping[tok[u]];
                if (slot >= 0 && slot < p.num_slots)
                    st_u4(static_cast<uint8_t*>(is_k ? p.k_cache : p.v_cache) +
                              (slot * p.num_kv_heads + kvh) * 256 + c * 16,
                          is_k ? rot : raw[u]);
            }
        }
    }
}

}  // namespace

cudaError_t launch_qk_norm_rope_store(const QkNormRopeParams& p, cudaStream_t stream) {
    if (p.n_tokens == 0) return cudaSuccess;
    const int Htot = p.num_q_heads + 2 * p.num_kv_heads;
    const int64_t rows = static_cast<int64_t>(p.n_tokens) * Htot;
    const int64_t half_warps = (rows + kRows - 1) / kRows;
    const unsigned grid = static_cast<unsigned>((half_warps * 16 + kThreads - 1) / kThreads);
    qk_norm_rope_store_kernel<<<grid, kThreads, 0, stream>>>(p);
    count_launch();
    return cudaGetLastError();
}

}  // namespace b200
