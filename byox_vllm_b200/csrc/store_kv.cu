// store_kvcache: slot_mapping scatter of this step's K/V rows into the paged cache.
// Replaces the Triton store_kvcache_kernel (reference nanovllm/layers/attention.py:10-31).
//
// Pure data movement, HBM/latency bound: every thread moves ONE 16-byte chunk of a K row and the
// matching chunk of the V row; consecutive threads take consecutive chunks of a row, so a row of
// Hkv*128 bf16 (256 B .. 2 KiB) is read and written as fully coalesced 128-byte lines. 22
// registers -> 8 CTAs x 8 warps per SM: the copy is kept in flight by occupancy, not by unrolling.
// Offsets are 64-bit; slots outside [0, num_slots) are skipped.
//
// What was tried and removed (numbers: profiles/r03_store_sweep.txt, 16384 tokens x 8 KV heads,
// L2 flushed clean before each call): 1 / 2 / 4 / 8 tokens per thread x 16- / 32-byte vectors.
// One token x 16 bytes is the fastest (21.5-22.5 us = 0.91-0.95 of the measured copy peak; the
// reference's Triton kernel 22.5); more bytes per thread only cost registers and occupancy (4 x 32 B: 82
// registers, 16 warps per SM, 24.6 us; 8 x 32 B: 154 registers, 26.6 us). Rounds 1-2 could not
// see this: their timing flushed L2 by WRITING a buffer, and every variant (and Triton) then
// measured 28.7 us = the write-back of that buffer's dirty lines. A TMA bulk-copy variant (one
// thread per persistent CTA driving cp.async.bulk through a shared-memory ring) was built in
// round 1, measured no faster at any size and removed in round 2 (profiles/r01_microbench_v3.log).
#include "common.h"
#include "ptx.cuh"

namespace b200 {

namespace {

constexpr int kVec = 16;

__global__ void __launch_bounds__(256)
store_kvcache_kernel(const uint8_t* __restrict__ key, int64_t key_row_bytes,
                     const uint8_t* __restrict__ value, int64_t value_row_bytes,
                     uint8_t* __restrict__ k_cache, uint8_t* __restrict__ v_cache,
                     const int32_t* __restrict__ slot_mapping, int32_t n_tokens,
                     int32_t chunks_per_row, int64_t num_slots) {
    // work item = (token, chunk); chunk is the fast index so a warp covers contiguous bytes
    const int64_t gid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int32_t chunk = static_cast<int32_t>(gid % chunks_per_row);
    const int64_t tok = gid / chunks_per_row;
    if (tok >= n_tokens) return;
    const int64_t slot = slot_mapping[tok];
    if (slot < 0 || slot >= num_slots) return;
    const int64_t row_bytes = static_cast<int64_t>(chunks_per_row) * kVec;
    const int4 k = ld_nc_v4(key + tok * key_row_bytes + chunk * kVec);
    const int4 v = ld_nc_v4(value + tok * value_row_bytes + chunk * kVec);
    st_v4(k_cache + slot * row_bytes + chunk * kVec, k);
    st_v4(v_cache + slot * row_bytes + chunk * kVec, v);
}

}  // namespace

cudaError_t launch_store_kvcache(const void* key, int64_t key_row_stride, const void* value,
                                 int64_t value_row_stride, void* k_cache, void* v_cache,
                                 const int32_t* slot_mapping, int32_t n_tokens, int32_t row_elems,
                                 int64_t num_slots, cudaStream_t stream) {
    if (n_tokens == 0) return cudaSuccess;
    constexpr int kThreads = 256;
    const int32_t chunks_per_row = row_elems * 2 / kVec;
    const int64_t items = static_cast<int64_t>(n_tokens) * chunks_per_row;
    const unsigned grid = static_cast<unsigned>((items + kThreads - 1) / kThreads);
    store_kvcache_kernel<<<grid, kThreads, 0, stream>>>(
        static_cast<const uint8_t*>(key), key_row_stride * 2, static_cast<const uint8_t*>(value),
        value_row_stride * 2, static_cast<uint8_t*>(k_cache), static_cast<uint8_t*>(v_cache),
        slot_mapping, n_tokens, chunks_per_row, num_slots);
    count_launch();
    return cudaGetLastError();
}

}  // namespace b200
