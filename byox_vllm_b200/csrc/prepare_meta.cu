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
