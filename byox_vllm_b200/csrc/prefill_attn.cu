// Varlen causal prefill attention (bottom-right aligned mask), dense or paged-prefix K/V.
// Replaces flash_attn_varlen_func on the prefill branch of Attention.forward
// (reference nanovllm/layers/attention.py:72-86).
//
// Round-1 baseline kernel: one CTA = 64 query rows of one (sequence, q-head), 4 warps x 16 rows,
// K/V streamed in 64-token tiles through a 2-stage cp.async ring (zero-filled beyond Lk so
// unwritten cache rows can never inject NaNs), S = Q.K^T and O += P.V on mma.sync.m16n8k16 with
// fp32 online softmax in registers. Query tiles are scheduled longest-first (reverse order) to
// balance the causal triangle. The tcgen05/TMEM kernel that replaces this one for long sequences
// lives in prefill_tc.cu once it is parity-green (DESIGN.md "prefill").
//
// Roofline: bf16 tensor pipe. Algorithmic flops = 4 * 128 * (visible q-k pairs) per q-head.
#include "common.h"
#include "ptx.cuh"

namespace b200 {

namespace {

constexpr int kBlockM = 64;
constexpr int kBlockN = 64;
constexpr int kPrefillThreads = 128;
constexpr int kHalfBytes = 64 * 128;  // [64 rows][64 dims] bf16, 128-byte rows, swz128
constexpr int kTileBytes = 2 * kHalfBytes;
constexpr int kPrefillSmem = kTileBytes /*Q*/ + 2 * 2 * kTileBytes /*K,V x 2 stages*/ + 1024;

// 64 rows x 256 B -> two swizzled [64][64] halves. `row_ptr(r)` returns the global address of
// row r or nullptr for rows that must read as zero.
template <class RowPtr>
__device__ __forceinline__ void load_tile_async(uint8_t* smem_tile, RowPtr row_ptr) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int idx = threadIdx.x + i * kPrefillThreads;  // 0..1023
        const int r = idx >> 4;
        const int c = idx & 15;  // 16-byte chunk of the 256-byte row
        const uint8_t* src = row_ptr(r);
        uint8_t* dst = smem_tile + (c >> 3) * kHalfBytes + swz128(r, c & 7);
        cp_async_16(dst, src ? src + c * 16 : static_cast<const uint8_t*>(nullptr) , src != nullptr);
    }
}

template <bool kF16>  // fp16 storage / operands (dtype code 1) instead of bf16
__global__ void __launch_bounds__(kPrefillThreads, 2)
prefill_varlen_kernel(const PrefillParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* q_smem = smem;
    uint8_t* kv_smem = smem + kTileBytes;  // stage s: K at s*2*kTileBytes, V at +kTileBytes

    const int seq = blockIdx.z;
    const int h = blockIdx.y;
    const int kvh = h / p.group;
    const int q_begin = p.cu_seqlens_q[seq];
    const int Lq = p.cu_seqlens_q[seq + 1] - q_begin;
    const int k_begin = p.cu_seqlens_k[seq];
    const int Lk = p.cu_seqlens_k[seq + 1] - k_begin;
    const int n_qtiles = (Lq + kBlockM - 1) / kBlockM;
    if (static_cast<int>(blockIdx.x) >= n_qtiles) return;
    const int qt = n_qtiles - 1 - blockIdx.x;  // heaviest (last) tiles first
    const int offset = Lk - Lq;                // row i sees keys j <= i + offset
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int Hq = p.num_q_heads;
    const int Hkv = p.num_kv_heads;

    const int j_end = min(Lk, qt * kBlockM + kBlockM + offset);  // exclusive bound on visible keys
    const int n_ktiles = j_end > 0 ? (j_end + kBlockN - 1) / kBlockN : 0;

    const uint8_t* qg = static_cast<const uint8_t*>(p.q);
    auto q_row_ptr = [&](int r) -> const uint8_t* {
        const int i = qt * kBlockM + r;
        if (i >= Lq) return nullptr;
        return qg + ((static_cast<int64_t>(q_begin) + i) * Hq + h) * 256;
    };
    const bool paged = p.block_tables != nullptr;
    const int32_t* table = paged ? p.block_tables + static_cast<int64_t>(seq) * p.table_row_stride
                                 : nullptr;
    auto load_kv = [&](int kt, int stage) {
        uint8_t* ks = kv_smem + stage * 2 * kTileBytes;
        uint8_t* vs = ks + kTileBytes;
        const int j0 = kt * kBlockN;
        if (paged) {
            // a 64-token tile never straddles a page (block_size % 64 == 0)
            const int64_t block = table[j0 / p.block_size];
            const int64_t slot0 = block * p.block_size + (j0 % p.block_size);
            const uint8_t* kc = static_cast<const uint8_t*>(p.k_cache);
            const uint8_t* vc = static_cast<const uint8_t*>(p.v_cache);
            load_tile_async(ks, [&](int r) -> const uint8_t* {
                return j0 + r < Lk ? kc + ((slot0 + r) * Hkv + kvh) * 256 : nullptr;
            });
            load_tile_async(vs, [&](int r) -> const uint8_t* {
                return j0 + r < Lk ? vc + ((slot0 + r) * Hkv + kvh) * 256 : nullptr;
            });
        } else {
            const uint8_t* kg = static_cast<const uint8_t*>(p.key);
            const uint8_t* vg = static_cast<const uint8_t*>(p.value);
            load_tile_async(ks, [&](int r) -> const uint8_t* {
                return j0 + r < Lk
                           ? kg + (static_cast<int64_t>(k_begin) + j0 + r) * p.key_row_stride * 2 +
                                 kvh * 256
                           : nullptr;
            });
            load_tile_async(vs, [&](int r) -> const uint8_t* {
                return j0 + r < Lk
                           ? vg + (static_cast<int64_t>(k_begin) + j0 + r) * p.value_row_stride * 2 +
                                 kvh * 256
                           : nullptr;
            });
        }
    };

    load_tile_async(q_smem, q_row_ptr);
    if (n_ktiles > 0) load_kv(0, 0);
    cp_async_commit();

    float m_run[2] = {-INFINITY, -INFINITY};
    float l_run[2] = {0.f, 0.f};
    float o[16][4];
#pragma unroll
    for (int n = 0; n < 16; ++n)
#pragma unroll
        for (int c = 0; c < 4; ++c) o[n][c] = 0.f;
    uint32_t qf[8][4];

    const uint32_t q_base = smem_u32(q_smem);
    const uint32_t kv_base = smem_u32(kv_smem);
    const int row_in_tile = warp * 16 + (lane >> 2);  // this thread's rows: +0 and +8
    const int qi0 = qt * kBlockM + row_in_tile;

    for (int kt = 0; kt < n_ktiles; ++kt) {
        const int stage = kt & 1;
        if (kt + 1 < n_ktiles) {
            load_kv(kt + 1, stage ^ 1);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (kt == 0) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
                ldmatrix_x4(qf[ks], q_base + (ks >> 2) * kHalfBytes +
                                        swz128(warp * 16 + (lane & 15), (ks & 3) * 2 + (lane >> 4)));
        }
        const uint32_t kb = kv_base + stage * 2 * kTileBytes;
        const uint32_t vb = kb + kTileBytes;

        // ---- S[16 x 64] = Q . K^T ----
        float s[8][4];
#pragma unroll
        for (int n = 0; n < 8; ++n)
#pragma unroll
            for (int c = 0; c < 4; ++c) s[n][c] = 0.f;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
#pragma unroll
            for (int np = 0; np < 4; ++np) {  // pairs of 8-token n-tiles
                uint32_t bfr[4];
                const uint32_t tok = np * 16 + ((lane >> 4) << 3) + (lane & 7);
                const uint32_t chunk = (ks & 3) * 2 + ((lane >> 3) & 1);
                ldmatrix_x4(bfr, kb + (ks >> 2) * kHalfBytes + swz128(tok, chunk));
                mma_16816<kF16>(s[2 * np], qf[ks], bfr[0], bfr[1]);
                mma_16816<kF16>(s[2 * np + 1], qf[ks], bfr[2], bfr[3]);
            }
        }

        // ---- scale, causal / length mask, online softmax ----
        const int j0 = kt * kBlockN + 2 * (lane & 3);
        const bool need_mask = (kt * kBlockN + kBlockN - 1 > qt * kBlockM + offset) ||
                               (kt * kBlockN + kBlockN > Lk);
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int n = 0; n < 8; ++n) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float v = s[n][c] * p.scale_log2;
                if (need_mask) {
                    const int j = j0 + n * 8 + (c & 1);
                    const int i = qi0 + (c >> 1) * 8;
                    if (j > i + offset || j >= Lk) v = -INFINITY;
                }
                s[n][c] = v;
                mx[c >> 1] = fmaxf(mx[c >> 1], v);
            }
        }
        float alpha[2], mu[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
            const float mn = fmaxf(m_run[r], mx[r]);
            mu[r] = mn == -INFINITY ? 0.f : mn;
            alpha[r] = fast_exp2(m_run[r] - mu[r]);
            m_run[r] = mn;
            l_run[r] *= alpha[r];
        }
#pragma unroll
        for (int n = 0; n < 8; ++n)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float e = fast_exp2(s[n][c] - mu[c >> 1]);
                s[n][c] = e;
                l_run[c >> 1] += e;
            }
#pragma unroll
        for (int n = 0; n < 16; ++n) {
            o[n][0] *= alpha[0];
            o[n][1] *= alpha[0];
            o[n][2] *= alpha[1];
            o[n][3] *= alpha[1];
        }

        // ---- O[16 x 128] += P . V ----
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {  // 16 tokens per k-step
            uint32_t a[4];
            a[0] = pack_16x2<kF16>(s[2 * kk][0], s[2 * kk][1]);
            a[1] = pack_16x2<kF16>(s[2 * kk][2], s[2 * kk][3]);
            a[2] = pack_16x2<kF16>(s[2 * kk + 1][0], s[2 * kk + 1][1]);
            a[3] = pack_16x2<kF16>(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
            for (int np = 0; np < 8; ++np) {  // pairs of 8-dim n-tiles
                uint32_t bfr[4];
                const uint32_t tok = kk * 16 + (((lane >> 3) & 1) << 3) + (lane & 7);
                const uint32_t ntile = np * 2 + (lane >> 4);
                ldmatrix_x4_trans(bfr, vb + (ntile >> 3) * kHalfBytes + swz128(tok, ntile & 7));
                mma_16816<kF16>(o[2 * np], a, bfr[0], bfr[1]);
                mma_16816<kF16>(o[2 * np + 1], a, bfr[2], bfr[3]);
            }
        }
        __syncthreads();  // stage is free for the prefetch issued by the next iteration
    }
    if (n_ktiles == 0) {
        cp_async_wait<0>();
        __syncthreads();
    }

    // ---- normalise, stage through shared memory (Q tile is dead), coalesced 16-byte stores ----
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }
    const float inv0 = l_run[0] > 0.f ? 1.f / l_run[0] : 0.f;
    const float inv1 = l_run[1] > 0.f ? 1.f / l_run[1] : 0.f;
    // q_smem is only read by ldmatrix at kt == 0 (before a __syncthreads) or never; safe to reuse
#pragma unroll
    for (int n = 0; n < 16; ++n) {
        const int chunk = n;  // 8 dims = one 16-byte chunk; this thread holds 2 of its 8 dims
        uint8_t* base = q_smem + (chunk >> 3) * kHalfBytes;
        *reinterpret_cast<uint32_t*>(base + swz128(row_in_tile, chunk & 7) + (lane & 3) * 4) =
            pack_16x2<kF16>(o[n][0] * inv0, o[n][1] * inv0);
        *reinterpret_cast<uint32_t*>(base + swz128(row_in_tile + 8, chunk & 7) + (lane & 3) * 4) =
            pack_16x2<kF16>(o[n][2] * inv1, o[n][3] * inv1);
    }
    __syncthreads();
    uint8_t* og = static_cast<uint8_t*>(p.out);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int idx = threadIdx.x + i * kPrefillThreads;
        const int r = idx >> 4;
        const int c = idx & 15;
        const int qi = qt * kBlockM + r;
        if (qi < Lq) {
            const int4 v =
                *reinterpret_cast<const int4*>(q_smem + (c >> 3) * kHalfBytes + swz128(r, c & 7));
            st_v4(og + ((static_cast<int64_t>(q_begin) + qi) * Hq + h) * 256 + c * 16, v);
        }
    }
}

}  // namespace

namespace {
template <bool kF16>
cudaError_t launch_prefill_t(const PrefillParams& p, cudaStream_t stream) {
    // the opt-in shared-memory size is a per-device function attribute
    static bool configured[64] = {};
    int dev = 0;
    if (cudaError_t e = cudaGetDevice(&dev); e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        cudaError_t e = cudaFuncSetAttribute(prefill_varlen_kernel<kF16>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kPrefillSmem);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    const int q_tiles = (p.max_seqlen_q + kBlockM - 1) / kBlockM;
    dim3 grid(q_tiles, p.num_q_heads, p.num_seqs);
    prefill_varlen_kernel<kF16><<<grid, kPrefillThreads, kPrefillSmem, stream>>>(p);
    count_launch();
    return cudaGetLastError();
}
}  // namespace

cudaError_t launch_prefill(const PrefillParams& p, cudaStream_t stream) {
    return p.f16 ? launch_prefill_t<true>(p, stream) : launch_prefill_t<false>(p, stream);
}

}  // namespace b200
