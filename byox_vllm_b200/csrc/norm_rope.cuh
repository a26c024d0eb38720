// Per-head RMSNorm + neox RoPE on one 128-element head row: the step the reference
// runs right before the attention path, nanovllm/models/qwen3.py:80-83:
//     q = q_norm(q); k = k_norm(k)            layers/layernorm.py:16-26  (rms_forward)
//     q, k = rotary_emb(positions, q, k)      layers/rotary_embedding.py:7-16, 44-47
// Rounding points follow the reference's eager arithmetic so results agree to the bf16 bit except
// where the fp32 sum of squares / rsqrt round differently (summation order, MUFU.RSQ):
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

// cos / sin of one token for the 8 rotation indices j = 8*(c&7) .. +7 a lane works on. The sine
// carries the sign of the lane's half (first half: y1 = x1 cos - x2 sin; second: y2 = x2 cos +
// x1 sin), so the rotation is mul, mul, add for every lane: a - b == a + (-b) exactly in IEEE.
struct CosSin8 {
    float c[8], s[8];
};
__device__ __forceinline__ CosSin8 load_cos_sin8(const float* __restrict__ cs_row, int c) {
    const int j0 = 8 * (c & 7);
    const float4 c0 = *reinterpret_cast<const float4*>(cs_row + j0);
    const float4 c1 = *reinterpret_cast<const float4*>(cs_row + j0 + 4);
    const float4 s0 = *reinterpret_cast<const float4*>(cs_row + 64 + j0);
    const float4 s1 = *reinterpret_cast<const float4*>(cs_row + 64 + j0 + 4);
    CosSin8 r;
    r.c[0] = c0.x, r.c[1] = c0.y, r.c[2] = c0.z, r.c[3] = c0.w;
    r.c[4] = c1.x, r.c[5] = c1.y, r.c[6] = c1.z, r.c[7] = c1.w;
    r.s[0] = s0.x, r.s[1] = s0.y, r.s[2] = s0.z, r.s[3] = s0.w;
    r.s[4] = s1.x, r.s[5] = s1.y, r.s[6] = s1.z, r.s[7] = s1.w;
    if (c < 8) {
#pragma unroll
        for (int i = 0; i < 8; ++i) r.s[i] = -r.s[i];
    }
    return r;
}

// bf16(bf16(a) * bf16(b)) for two packed pairs: one HMUL2.BF16 (the product of two bf16 values is
// exact in fp32, so the single rounding of the packed multiply equals torch's float-then-round).
__device__ __forceinline__ uint32_t bf16x2_mul(uint32_t a, uint32_t b) {
    __nv_bfloat162 r = __hmul2(*reinterpret_cast<__nv_bfloat162*>(&a),
                               *reinterpret_cast<__nv_bfloat162*>(&b));
    return *reinterpret_cast<uint32_t*>(&r);
}

// Row spread over a half-warp: lane c = (lane & 15) holds elements 8c .. 8c+7 as one 16-byte
// vector `raw`; `wv` = weight[8c .. 8c+7] (ignored when !has_norm); cs = cos/sin of this token.
// Must be called by all 16 lanes of the half-warp named by `hmask` (0xffff or 0xffff0000): the
// shuffles use xor <= 8, so the two halves of a warp may work on different rows or diverge.
__device__ __forceinline__ uint4 norm_rope_row16(uint4 raw, uint4 wv, bool has_norm, float eps,
                                                 const CosSin8& cs, int c, uint32_t hmask) {
    uint32_t xw[4] = {raw.x, raw.y, raw.z, raw.w};
    if (has_norm) {
        float x[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            x[2 * i] = bf16_lo(xw[i]);
            x[2 * i + 1] = bf16_hi(xw[i]);
        }
        float ss = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += x[i] * x[i];
#pragma unroll
        for (int off = 1; off < 16; off <<= 1) ss += __shfl_xor_sync(hmask, ss, off);
        // rsqrtf (MUFU.RSQ, <= 2 ulp) is what torch.rsqrt executes on a GPU, eager or compiled;
        // an IEEE 1/sqrt (what torch does on the CPU) costs ~40 instructions per row and this pass is
        // instruction-issue bound before it is HBM bound
        const float r = rsqrtf(ss * (1.0f / 128.0f) + eps);
        const uint32_t ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
            xw[i] = bf16x2_mul(bf16_pack(__fmul_rn(x[2 * i], r), __fmul_rn(x[2 * i + 1], r)),
                               ww[i]);
    }
    // partner half of the rotation pair lives in lane c ^ 8 (elements +-64)
    uint32_t ow[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t other = __shfl_xor_sync(hmask, xw[i], 8);
        const float a0 = __fmul_rn(bf16_lo(xw[i]), cs.c[2 * i]);
        const float a1 = __fmul_rn(bf16_hi(xw[i]), cs.c[2 * i + 1]);
        const float b0 = __fmul_rn(bf16_lo(other), cs.s[2 * i]);
        const float b1 = __fmul_rn(bf16_hi(other), cs.s[2 * i + 1]);
        ow[i] = bf16_pack(__fadd_rn(a0, b0), __fadd_rn(a1, b1));
    }
    return make_uint4(ow[0], ow[1], ow[2], ow[3]);
}

// cos_sin_cache row of a token; positions outside [0, max_position) are clamped (the reference
// would raise an index error: rotary_embedding.py:44).
__device__ __forceinline__ const float* cos_sin_row(const NormRopeArgs& a, int64_t token) {
    int64_t pos = a.positions[token];
    pos = pos < 0 ? 0 : (pos >= a.max_position ? a.max_position - 1 : pos);
    return a.cos_sin + pos * 128;
}

}  // namespace b200
