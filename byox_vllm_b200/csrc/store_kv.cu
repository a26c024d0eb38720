// store_kvcache: slot_mapping scatter of this step's K/V rows into the paged cache.
// Replaces the Triton store_kvcache_kernel (reference nanovllm/layers/attention.py:10-31).
//
// Pure data movement, HBM/latency bound: every thread moves one kVec-byte chunk (16, or 32 with
// the 256-bit global accesses of sm_100) of a K row and the matching chunk of the V row, for
// kTokensPerThread tokens: all loads are issued before the first store, consecutive threads take
// consecutive chunks of a row, so a row of Hkv*128 bf16 (256 B .. 2 KiB) is read and written as
// fully coalesced 128-byte lines. Offsets are 64-bit; slots outside [0, num_slots) are skipped.
// (A TMA bulk-copy variant — one thread per persistent CTA driving cp.async.bulk through a
// shared-memory ring — was built in round 1, measured no faster at any size and removed in
// round 2; profiles/r01_microbench_v3.log has its numbers.)
#include "common.h"
#include "ptx.cuh"

#include <cstdlib>

namespace b200 {

namespace {

template <int kVec>
struct Chunk;
template <>
struct Chunk<16> {
    int4 v;
    __device__ __forceinline__ void load(const uint8_t* p) { v = ld_nc_v4(p); }
    __device__ __forceinline__ void store(uint8_t* p) const { st_v4(p, v); }
};
template <>
struct Chunk<32> {
    uint32_t v[8];
    __device__ __forceinline__ void load(const uint8_t* p) { ld_nc_v8(p, v); }
    __device__ __forceinline__ void store(uint8_t* p) const { st_v8(p, v); }
};

template <int kTokensPerThread, int kVec>
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
    const int64_t row_bytes = static_cast<int64_t>(chunks_per_row) * kVec;

    Chunk<kVec> kv[kTokensPerThread], vv[kTokensPerThread];
    int64_t slot[kTokensPerThread];
#pragma unroll
    for (int i = 0; i < kTokensPerThread; ++i) {
        const int64_t tok = tok0 + i;
        slot[i] = -1;
        if (tok < n_tokens) {
            const int64_t s = slot_mapping[tok];
            if (s >= 0 && s < num_slots) {
                slot[i] = s;
                kv[i].load(key + tok * key_row_bytes + chunk * kVec);
                vv[i].load(value + tok * value_row_bytes + chunk * kVec);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < kTokensPerThread; ++i) {
        if (slot[i] >= 0) {
            kv[i].store(k_cache + slot[i] * row_bytes + chunk * kVec);
            vv[i].store(v_cache + slot[i] * row_bytes + chunk * kVec);
        }
    }
}

template <int kTok, int kVec>
void launch(const void* key, int64_t key_row_stride, const void* value, int64_t value_row_stride,
            void* k_cache, void* v_cache, const int32_t* slot_mapping, int32_t n_tokens,
            int32_t row_elems, int64_t num_slots, cudaStream_t stream) {
    constexpr int kThreads = 256;
    const int32_t chunks_per_row = row_elems * 2 / kVec;
    const int64_t items = (static_cast<int64_t>(n_tokens) + kTok - 1) / kTok * chunks_per_row;
    const unsigned grid = static_cast<unsigned>((items + kThreads - 1) / kThreads);
    store_kvcache_kernel<kTok, kVec><<<grid, kThreads, 0, stream>>>(
        static_cast<const uint8_t*>(key), key_row_stride * 2, static_cast<const uint8_t*>(value),
        value_row_stride * 2, static_cast<uint8_t*>(k_cache), static_cast<uint8_t*>(v_cache),
        slot_mapping, n_tokens, chunks_per_row, num_slots);
}

bool aligned32(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31u) == 0; }

}  // namespace

cudaError_t launch_store_kvcache(const void* key, int64_t key_row_stride, const void* value,
                                 int64_t value_row_stride, void* k_cache, void* v_cache,
                                 const int32_t* slot_mapping, int32_t n_tokens, int32_t row_elems,
                                 int64_t num_slots, cudaStream_t stream) {
    if (n_tokens == 0) return cudaSuccess;
    // B200ATTN_STORE_VEC=16|32, B200ATTN_STORE_TOK=4|8: experiment knobs for the prefill-size path
    static const int want_vec = [] {
        const char* e = std::getenv("B200ATTN_STORE_VEC");
        return e && std::atoi(e) == 16 ? 16 : 32;
    }();
    static const int want_tok = [] {
        const char* e = std::getenv("B200ATTN_STORE_TOK");
        const int t = e ? std::atoi(e) : 4;
        return (t == 1 || t == 2 || t == 8) ? t : 4;
    }();
    const bool vec32 = want_vec == 32 && row_elems % 16 == 0 && key_row_stride % 16 == 0 &&
                       value_row_stride % 16 == 0 && aligned32(key) && aligned32(value) &&
                       aligned32(k_cache) && aligned32(v_cache);
#define B200_STORE_ARGS                                                                        \
    key, key_row_stride, value, value_row_stride, k_cache, v_cache, slot_mapping, n_tokens,    \
        row_elems, num_slots, stream
    // Small (decode) launches are latency bound: one token per thread keeps the grid wide.
    // Large (prefill) launches move up to 128 MiB: several tokens per thread, all loads in flight
    // before the first store.
    if (n_tokens >= 2048) {
        if (vec32) {
            if (want_tok == 8) launch<8, 32>(B200_STORE_ARGS);
            else if (want_tok == 2) launch<2, 32>(B200_STORE_ARGS);
            else if (want_tok == 1) launch<1, 32>(B200_STORE_ARGS);
            else launch<4, 32>(B200_STORE_ARGS);
        } else {
            if (want_tok == 8) launch<8, 16>(B200_STORE_ARGS);
            else if (want_tok == 2) launch<2, 16>(B200_STORE_ARGS);
            else if (want_tok == 1) launch<1, 16>(B200_STORE_ARGS);
            else launch<4, 16>(B200_STORE_ARGS);
        }
    } else {
        launch<1, 16>(B200_STORE_ARGS);
    }
#undef B200_STORE_ARGS
    count_launch();
    return cudaGetLastError();
}

}  // namespace b200
