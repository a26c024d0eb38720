// store_kvcache: slot_mapping scatter of this step's K/V rows into the paged cache.
// Replaces the Triton store_kvcache_kernel (reference nanovllm/layers/attention.py:10-31).
//
// Pure data movement, HBM/latency bound: every thread moves one 16-byte chunk of a K row and the
// matching chunk of the V row (two independent 128-bit loads in flight, then two 128-bit stores),
// consecutive threads take consecutive chunks of a row, so a row of Hkv*128 bf16 (256 B .. 2 KiB)
// is read and written as fully coalesced 128-byte lines. Offsets are 64-bit.
#include "common.h"
#include "ptx.cuh"

#include <cstdlib>

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


// ---------------------------------------------------------------------------------------------
// Bulk-copy variant for prefill-size launches: no thread ever touches the payload. One thread per
// persistent CTA drives the TMA engine: for token t it bulk-loads the K row and the (strided) V
// row into a shared-memory stage (cp.async.bulk, mbarrier completion), and as soon as they land
// bulk-stores both rows from shared memory to the token's cache slot. Loads run kCopyStages - 1
// tokens ahead of the stores; a stage is reloaded once the stores of the token that used it have
// finished reading it (cp.async.bulk.wait_group.read). Tokens with a negative / out-of-range slot
// move no bytes. Bytes in flight per SM = CTAs/SM x stages x 2 rows, bounded by shared memory.
constexpr int kCopyStages = 6;

__global__ void __launch_bounds__(32)
store_kvcache_bulk_kernel(const uint8_t* __restrict__ key, int64_t key_row_bytes,
                          const uint8_t* __restrict__ value, int64_t value_row_bytes,
                          uint8_t* __restrict__ k_cache, uint8_t* __restrict__ v_cache,
                          const int32_t* __restrict__ slot_mapping, int32_t n_tokens,
                          uint32_t row_bytes, int64_t num_slots) {
    extern __shared__ uint8_t copy_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(copy_smem_raw) + 127) & ~static_cast<uintptr_t>(127));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kCopyStages * 2 * row_bytes);
    if (threadIdx.x != 0) return;
#pragma unroll
    for (int s = 0; s < kCopyStages; ++s) mbar_init(&full_bar[s], 1);
    fence_mbar_init();

    const int stride = gridDim.x;
    const int my_count = (n_tokens - static_cast<int>(blockIdx.x) + stride - 1) / stride;
    int64_t slots[kCopyStages];  // slot of the token held by each stage (-1: nothing loaded)

    auto issue_load = [&](int it) {
        const int stage = it % kCopyStages;
        const int64_t tok = blockIdx.x + static_cast<int64_t>(it) * stride;
        int64_t slot = slot_mapping[tok];
        if (slot < 0 || slot >= num_slots) slot = -1;
        slots[stage] = slot;
        if (slot < 0) return;  // nothing to move: the consumer side skips the barrier too
        uint8_t* dst = smem + stage * 2 * row_bytes;
        mbar_expect_tx(&full_bar[stage], 2 * row_bytes);
        bulk_load_1d(dst, key + tok * key_row_bytes, row_bytes, &full_bar[stage]);
        bulk_load_1d(dst + row_bytes, value + tok * value_row_bytes, row_bytes, &full_bar[stage]);
    };

    const int ahead = kCopyStages - 1;
    for (int it = 0; it < ahead && it < my_count; ++it) issue_load(it);
    uint32_t phase_bits = 0;  // bit s = parity the next wait on stage s expects
    for (int it = 0; it < my_count; ++it) {
        const int stage = it % kCopyStages;
        const int64_t slot = slots[stage];
        if (slot >= 0) {
            mbar_wait(&full_bar[stage], (phase_bits >> stage) & 1u);
            phase_bits ^= 1u << stage;
            const uint8_t* src = smem + stage * 2 * row_bytes;
            bulk_store_1d(k_cache + slot * row_bytes, src, row_bytes);
            bulk_store_1d(v_cache + slot * row_bytes, src + row_bytes, row_bytes);
        }
        bulk_commit_group();
        // the stores of token it - 1 have read their stage (it - 1) % stages == (it + ahead) %
        // stages (never used yet when it == 0): load token it + ahead into it
        bulk_wait_group_read<1>();
        if (it + ahead < my_count) issue_load(it + ahead);
    }
    bulk_wait_group<0>();
}

cudaError_t launch_store_kvcache(const void* key, int64_t key_row_stride, const void* value,
                                 int64_t value_row_stride, void* k_cache, void* v_cache,
                                 const int32_t* slot_mapping, int32_t n_tokens, int32_t row_elems,
                                 int64_t num_slots, cudaStream_t stream) {
    if (n_tokens == 0) return cudaSuccess;
    const int32_t chunks_per_row = row_elems / 8;  // 8 x 16-bit = 16 bytes
    constexpr int kThreads = 256;
    // B200ATTN_STORE=bulk selects the TMA bulk-copy variant. Measured on B200 (N = 16384, Hkv 8,
    // 134 MB moved): 29.7 us vs 27.7-29.2 us for the direct kernel below - both sit on the same
    // floor (about 5 us of launch + a 0.86-of-peak copy), so the direct kernel stays the default.
    static const int mode = [] {
        const char* e = std::getenv("B200ATTN_STORE");
        return !e ? 0 : (e[0] == 'b' ? 1 : (e[0] == 'd' ? 2 : 0));
    }();
    const uint32_t row_bytes = static_cast<uint32_t>(row_elems) * 2;
    const int smem = kCopyStages * 2 * row_bytes + kCopyStages * 8 + 128;
    if (smem <= 64 * 1024 &&
        mode == 1 && row_bytes >= 256) {
        int sm_count = current_device_sm_count();
        if (sm_count <= 0) sm_count = 148;
        cudaError_t e = cudaFuncSetAttribute(store_kvcache_bulk_kernel,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        int per_sm = (200 * 1024) / (smem + 1024);
        per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
        const int64_t want = static_cast<int64_t>(sm_count) * per_sm;
        const unsigned grid = static_cast<unsigned>(n_tokens < want ? n_tokens : want);
        store_kvcache_bulk_kernel<<<grid, 32, smem, stream>>>(
            static_cast<const uint8_t*>(key), key_row_stride * 2,
            static_cast<const uint8_t*>(value), value_row_stride * 2,
            static_cast<uint8_t*>(k_cache), static_cast<uint8_t*>(v_cache), slot_mapping, n_tokens,
            row_bytes, num_slots);
        count_launch();
        return cudaGetLastError();
    }
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
