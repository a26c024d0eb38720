// Decode-step metadata expansion on the device. Replaces, per decode step,
//   ModelRunner._prepare_decode        reference nanovllm/engine/model_runner.py:206-226
//   ModelRunner._prepare_block_tables  :152-158
//   the graph-variable refresh of ModelRunner._run_model  :242-254
// i.e. five Python lists -> five pinned tensors -> five H2D copies -> seven small device ops,
// with ONE H2D copy of a compact record and ONE launch that writes the CUDA-graph input buffers
// directly. Pure int32/int64 index work: results must be bit-identical to the reference's.
//
// Compact record (int32, host-packed): last_token[B] | num_tokens[B] | table_len[B] | blocks[sum]
// where blocks is the concatenation of every sequence's block_table (table_len[b] entries).
//
// Grid: every CTA re-derives the exclusive scan of table_len (B <= 1024 rows, one per thread,
// Hillis-Steele in shared memory) and then expands its slice of rows. Latency bound (a few KB).
#include "common.h"

namespace b200 {

namespace {

constexpr int kMetaThreads = 1024;  // == max batch
constexpr int kRowsPerCta = 32;

__global__ void __launch_bounds__(kMetaThreads)
prepare_decode_kernel(const PrepareDecodeParams p) {
    __shared__ int32_t scan[kMetaThreads];
    const int tid = threadIdx.x;
    const int B = p.batch;
    const int32_t* last_token = p.packed;
    const int32_t* num_tokens = p.packed + B;
    const int32_t* table_len = p.packed + 2 * B;
    const int32_t* blocks = p.packed + 3 * B;

    // inclusive scan of table_len over the batch
    const int32_t my_len = tid < B ? table_len[tid] : 0;
    scan[tid] = my_len;
    __syncthreads();
    for (int off = 1; off < kMetaThreads; off <<= 1) {
        const int32_t add = tid >= off ? scan[tid - off] : 0;
        __syncthreads();
        scan[tid] += add;
        __syncthreads();
    }

    // per-row scalars: handled by the CTA that owns the row (and the padding rows by the last
    // CTAs' spare threads: rows >= B up to capacity get slot -1 / context_len 0, which is what
    // `slot_mapping.fill_(-1)` / `context_lens.zero_()` leave there, model_runner.py:248-251)
    const int row0 = blockIdx.x * kRowsPerCta;
    if (tid < kRowsPerCta) {
        const int b = row0 + tid;
        if (b < B) {
            const int32_t n = num_tokens[b];
            const int32_t len = table_len[b];
            const int32_t off = scan[b] - len;
            // sequence.py:62-68: num_blocks, last_block_num_tokens
            const int32_t nblk = (n + p.block_size - 1) / p.block_size;
            const int32_t last_n = n - (nblk - 1) * p.block_size;
            // model_runner.py:215: seq.block_table[-1] * block_size + last_block_num_tokens - 1
            const int32_t last_block = len > 0 ? blocks[off + len - 1] : 0;
            p.input_ids[b] = last_token[b];
            p.positions[b] = n - 1;
            p.context_lens[b] = n;
            p.slot_mapping[b] = last_block * p.block_size + last_n - 1;
        }
    }
    for (int b = B + blockIdx.x * kMetaThreads + tid; b < p.capacity;
         b += gridDim.x * kMetaThreads) {
        p.slot_mapping[b] = -1;
        p.context_lens[b] = 0;
    }
    // block table rows [row0, row0 + kRowsPerCta) x [0, table_width): ids, then -1 padding
    // (model_runner.py:153-154); columns >= table_width keep their old content, as the
    // reference's `block_tables[:bs, :W] = ...` does.
    const int W = p.table_width;
    const int rows = min(kRowsPerCta, B - row0);
    for (int i = tid; i < rows * W; i += kMetaThreads) {
        const int r = i / W, j = i - r * W;
        const int b = row0 + r;
        const int32_t len = table_len[b];
        const int32_t off = scan[b] - len;
        p.block_tables[static_cast<int64_t>(b) * p.table_row_stride + j] =
            j < len ? blocks[off + j] : -1;
    }
}

}  // namespace

cudaError_t launch_prepare_decode(const PrepareDecodeParams& p, cudaStream_t stream) {
    if (p.batch == 0 && p.capacity == 0) return cudaSuccess;
    const int grid = p.batch > 0 ? (p.batch + kRowsPerCta - 1) / kRowsPerCta : 1;
    prepare_decode_kernel<<<grid, kMetaThreads, 0, stream>>>(p);
    count_launch();
    return cudaGetLastError();
}

}  // namespace b200

// ---------------------------------------------------------------------------------------------
// Prefill-step metadata expansion (second half). Replaces the per-token Python list
// building of ModelRunner._prepare_prefill (reference engine/model_runner.py:160-204): for every
// sequence s with n_s tokens, c_s cached tokens and block table bt_s,
//     positions     += range(c_s, n_s)                                      (int64)
//     cu_seqlens_q  += n_s - c_s,   cu_seqlens_k += n_s                      (int32 prefix sums)
//     slot_mapping  += bt_s[i] * bs + j   for blocks i in [c_s / bs, ceil(n_s / bs)), j in
//                      [0, bs) or [0, last_block_num_tokens) for the last block - only for
//                      sequences that have a block table (warm-up sequences have none, :179-180)
// Compact record (int32): num_tokens[S] | num_cached[S] | table_len[S] | blocks[sum(table_len)].
// Every CTA re-derives the three inclusive scans in shared memory (S <= 1024) and expands its
// slice of the output indices with a binary search over the sequences. Bit-exact integer work.
namespace b200 {

namespace {

__device__ __forceinline__ void block_inclusive_scan(int32_t* a, int tid) {
    for (int off = 1; off < kMetaThreads; off <<= 1) {
        const int32_t add = tid >= off ? a[tid - off] : 0;
        __syncthreads();
        a[tid] += add;
        __syncthreads();
    }
}

// first s with scan[s] > i (scan inclusive, non-decreasing)
__device__ __forceinline__ int upper_seq(const int32_t* scan, int S, int32_t i) {
    int lo = 0, hi = S - 1;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (scan[mid] > i) hi = mid; else lo = mid + 1;
    }
    return lo;
}

__global__ void __launch_bounds__(kMetaThreads)
prepare_prefill_kernel(const PreparePrefillParams p) {
    __shared__ int32_t q_scan[kMetaThreads];     // inclusive scan of n - cached
    __shared__ int32_t slot_scan[kMetaThreads];  // inclusive scan of slots per sequence
    __shared__ int32_t tab_scan[kMetaThreads];   // inclusive scan of table_len
    const int tid = threadIdx.x;
    const int S = p.num_seqs;
    const int32_t* num_tokens = p.packed;
    const int32_t* num_cached = p.packed + S;
    const int32_t* table_len = p.packed + 2 * S;
    const int32_t* blocks = p.packed + 3 * S;
    const int32_t bs = p.block_size;

    int32_t n = 0, c = 0, len = 0;
    if (tid < S) {
        n = num_tokens[tid];
        c = num_cached[tid];
        len = table_len[tid];
    }
    q_scan[tid] = n - c;
    // sequence.py:58-60: num_cached_blocks = num_cached_tokens // block_size
    slot_scan[tid] = len > 0 ? n - (c / bs) * bs : 0;
    tab_scan[tid] = len;
    __syncthreads();
    block_inclusive_scan(q_scan, tid);
    block_inclusive_scan(slot_scan, tid);
    block_inclusive_scan(tab_scan, tid);

    if (blockIdx.x == 0) {
        // cu_seqlens_q / cu_seqlens_k: [0, prefix sums]
        if (tid == 0) {
            p.cu_seqlens_q[0] = 0;
            p.cu_seqlens_k[0] = 0;
        }
        if (tid < S) p.cu_seqlens_q[tid + 1] = q_scan[tid];
    }
    __shared__ int32_t k_scan[kMetaThreads];  // inclusive scan of n (cu_seqlens_k)
    k_scan[tid] = n;
    __syncthreads();
    block_inclusive_scan(k_scan, tid);
    if (blockIdx.x == 0 && tid < S) p.cu_seqlens_k[tid + 1] = k_scan[tid];
    if (S == 0) return;

    const int32_t total_q = q_scan[S - 1];
    const int32_t total_slots = slot_scan[S - 1];
    const int32_t stride = gridDim.x * kMetaThreads;
    for (int32_t i = blockIdx.x * kMetaThreads + tid; i < total_q; i += stride) {
        const int s = upper_seq(q_scan, S, i);
        const int32_t local = i - (q_scan[s] - (num_tokens[s] - num_cached[s]));
        p.positions[i] = num_cached[s] + local;
    }
    for (int32_t i = blockIdx.x * kMetaThreads + tid; i < total_slots; i += stride) {
        const int s = upper_seq(slot_scan, S, i);
        const int32_t first_block = num_cached[s] / bs;
        const int32_t seq_slots = num_tokens[s] - first_block * bs;
        const int32_t local = i - (slot_scan[s] - seq_slots);  // token offset from first_block
        const int32_t blk = first_block + local / bs;
        const int32_t off = tab_scan[s] - table_len[s];
        p.slot_mapping[i] = blocks[off + blk] * bs + local % bs;
    }
}

}  // namespace

cudaError_t launch_prepare_prefill(const PreparePrefillParams& p, int32_t total_hint,
                                   cudaStream_t stream) {
    int grid = (total_hint + kMetaThreads - 1) / kMetaThreads;
    grid = grid < 1 ? 1 : (grid > 64 ? 64 : grid);
    prepare_prefill_kernel<<<grid, kMetaThreads, 0, stream>>>(p);
    count_launch();
    return cudaGetLastError();
}

}  // namespace b200
