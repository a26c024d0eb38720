// store_kvcache: slot_mapping scatter of this step's K/V rows into the paged cache.
// Replaces the Triton store_kvcache_kernel (reference nanovllm/layers/attention.py:10-31).
//
// Pure data movement, HBM/latency bound: every thread moves one 16-byte chunk of a K row and the
// matching chunk of the V row (two independent 128-bit loads in flight, then two 128-bit stores),
// consecutive threads take consecutive chunks of a row, so a row of Hkv*128 bf16 (256 B .. 2 KiB)
// is read and written as fully coalesced 128-byte lines. Offsets are 64-bit.
#include "common.h"
#include "ptx.cuh"

namespace b200 {

template <int kTokensPerThread>
__global__ void __launch_bounds__(256)
store_kvcache_kernel(const uint8_t* __restrict__ key, int64_t key_row_bytes,
                     const uint8_t* __restrict__ value, int64_t value_row_bytes,
                     uint8_t* __restrict__ k_cache, uint8_t* __restrict__ v_cache,
                     const int32_t* __restrict__ slot_mapping, int32_t n_tokens,
                     int32_t chunks_per_row, int64_t num_slots) {
    // work item = (token group, chunk); chunk is the fast index so a warp covers contiguous bytes
    const int64_t gid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int32_t chunk = static_cast<int32_t>(gid % chunks_per_row);
    const int64_t tok0 = (gid / chunks_per_row) * kTokensPerThread;
    const int64_t row_bytes = static_cast<int64_t>(chunks_per_row) * 16;

    int4 kv[kTokensPerThread], vv[kTokensPerThread];
    int64_t slot[kTokensPerThread];
#pragma unroll
    for (int i = 0; i < kTokensPerThread; ++i) {
        const int64_t tok = tok0 + i;
        slot[i] = -1;
        if (tok < n_tokens) {
            const int64_t s = slot_mapping[tok];
            if (s >= 0 && s < num_slots) {
                slot[i] = s;
                kv[i] = ld_nc_v4(key + tok * key_row_bytes + chunk * 16);
                vv[i] = ld_nc_v4(value + tok * value_row_bytes + chunk * 16);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < kTokensPerThread; ++i) {
        if (slot[i] >= 0) {
            st_v4(k_cache + slot[i] * row_bytes + chunk * 16, kv[i]);
            st_v4(v_cache + slot[i] * row_bytes + chunk * 16, vv[i]);
        }
    }
}

cudaError_t launch_store_kvcache(const void* key, int64_t key_row_stride, const void* value,
                                 int64_t value_row_stride, void* k_cache, void* v_cache,
                                 const int32_t* slot_mapping, int32_t n_tokens, int32_t row_elems,
                                 int64_t num_slots, cudaStream_t stream) {
    if (n_tokens == 0) return cudaSuccess;
    const int32_t chunks_per_row = row_elems / 8;  // 8 x 16-bit = 16 bytes
    constexpr int kThreads = 256;
    // Small (decode) launches are latency bound: one token per thread keeps the grid wide.
    // Large (prefill) launches move up to 128 MiB: four tokens per thread = 8 loads in flight.
    if (n_tokens >= 2048) {
        constexpr int kTok = 4;
        const int64_t items = (static_cast<int64_t>(n_tokens) + kTok - 1) / kTok * chunks_per_row;
        const unsigned grid = static_cast<unsigned>((items + kThreads - 1) / kThreads);
        store_kvcache_kernel<kTok><<<grid, kThreads, 0, stream>>>(
            static_cast<const uint8_t*>(key), key_row_stride * 2,
            static_cast<const uint8_t*>(value), value_row_stride * 2,
            static_cast<uint8_t*>(k_cache), static_cast<uint8_t*>(v_cache), slot_mapping, n_tokens,
            chunks_per_row, num_slots);
    } else {
        const int64_t items = static_cast<int64_t>(n_tokens) * chunks_per_row;
        const unsigned grid = static_cast<unsigned>((items + kThreads - 1) / kThreads);
        store_kvcache_kernel<1><<<grid, kThreads, 0, stream>>>(
            static_cast<const uint8_t*>(key), key_row_stride * 2,
            static_cast<const uint8_t*>(value), value_row_stride * 2,
            static_cast<uint8_t*>(k_cache), static_cast<uint8_t*>(v_cache), slot_mapping, n_tokens,
            chunks_per_row, num_slots);
    }
    count_launch();
    return cudaGetLastError();
}

}  // namespace b200
