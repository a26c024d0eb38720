// Per-head RMSNorm + neox RoPE on one 128-element head row: the step the reference
// runs right before the attention path, nanovllm/models/qwen3.py:80-83:
//     q = q_norm(q); k = k_norm(k)            layers/layernorm.py:16-26  (rms_forward)
//     q, k = rotary_emb(positions, q, k)      layers/rotary_embedding.py:7-16, 44-47
// Rounding points follow the reference's eager arithmetic so results agree to the bf16 bit except
// where the fp32 sum of squares rounds differently (different summation order):
//     x32 = float(x);  x32 *= 1/sqrt(mean(x32^2) + eps);  xb = bf16(x32);  xw = bf16(xb * w)
//     y1 = xw1*cos - xw2*sin;  y2 = xw2*cos + xw1*sin   (fp32 mul, mul, add/sub: no FMA);  bf16(y)
// where (xw1, xw2) = elements (j, j + 64) and cos/sin = cos_sin_cache[pos][j], [pos][64 + j].
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace b200 {

struct NormRopeArgs {
    const int64_t* positions;   // [n_tokens] int64 (models/qwen3.py forward arg); nullptr => off
    const void* q_weight;       // bf16[128] RMSNorm weight of q heads; nullptr => no norm
    const void* k_weight;       // bf16[128] RMSNorm weight of k heads; nullptr => no norm
    const float* cos_sin;       // [max_position, 128] fp32: cos(64) | sin(64)
    int32_t max_position;
    float eps;
};

__device__ __forceinline__ float bf16_round(float x) {
    return __bfloat162float(__float2bfloat16_rn(x));
}
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t bf16_pack(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}

// Row spread over a half-warp: lane c = (lane & 15) holds elements 8c .. 8c+7 as one 16-byte
// vector `raw`; `wv` = weight[8c .. 8c+7] (ignored when !has_norm); cs_row = cos_sin_cache row of
// this token. MUST be called by all 32 lanes of the warp (full-mask shuffles with xor <= 8 stay
// inside each half-warp, so the two halves may work on different rows).
__device__ __forceinline__ uint4 norm_rope_row16(uint4 raw, uint4 wv, bool has_norm, float eps,
                                                 const float* __restrict__ cs_row, int c) {
    float x[8];
    const uint32_t rw[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        x[2 * i] = bf16_lo(rw[i]);
        x[2 * i + 1] = bf16_hi(rw[i]);
    }
    // cos/sin of rotation index j = 8*(c&7) + i; issued early so they overlap the reduction
    const int j0 = 8 * (c & 7);
    const float4 c0 = *reinterpret_cast<const float4*>(cs_row + j0);
    const float4 c1 = *reinterpret_cast<const float4*>(cs_row + j0 + 4);
    const float4 s0 = *reinterpret_cast<const float4*>(cs_row + 64 + j0);
    const float4 s1 = *reinterpret_cast<const float4*>(cs_row + 64 + j0 + 4);
    if (has_norm) {
        float ss = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += x[i] * x[i];
#pragma unroll
        for (int off = 1; off < 16; off <<= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
        const float r = 1.0f / sqrtf(ss * (1.0f / 128.0f) + eps);  // IEEE div/sqrt, as torch CPU
        const uint32_t ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            x[2 * i] = bf16_round(__fmul_rn(bf16_round(__fmul_rn(x[2 * i], r)), bf16_lo(ww[i])));
            x[2 * i + 1] =
                bf16_round(__fmul_rn(bf16_round(__fmul_rn(x[2 * i + 1], r)), bf16_hi(ww[i])));
        }
    }
    // partner half of the rotation pair lives in lane c ^ 8 (elements +-64)
    float y[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t mine = bf16_pack(x[2 * i], x[2 * i + 1]);  // exact: x is bf16-valued
        const uint32_t other = __shfl_xor_sync(0xffffffffu, mine, 8);
        y[2 * i] = bf16_lo(other);
        y[2 * i + 1] = bf16_hi(other);
    }
    const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    const float sn[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    const bool first_half = c < 8;
    float o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float a = __fmul_rn(x[i], cs[i]);
        const float b = __fmul_rn(y[i], sn[i]);
        o[i] = first_half ? __fsub_rn(a, b) : __fadd_rn(a, b);
    }
    uint4 out;
    out.x = bf16_pack(o[0], o[1]);
    out.y = bf16_pack(o[2], o[3]);
    out.z = bf16_pack(o[4], o[5]);
    out.w = bf16_pack(o[6], o[7]);
    return out;
}

// cos_sin_cache row of a token; positions outside [0, max_position) are clamped (the reference
// would raise an index error: rotary_embedding.py:44).
__device__ __forceinline__ const float* cos_sin_row(const NormRopeArgs& a, int64_t token) {
    int64_t pos = a.positions[token];
    pos = pos < 0 ? 0 : (pos >= a.max_position ? a.max_position - 1 : pos);
    return a.cos_sin + pos * 128;
}

}  // namespace b200
