// Shared-prefix ("cascade") decode: plan kernel + host heuristic.
//
// nanovllm's BlockManager gives sequences that share a prompt prefix the SAME physical block ids
// at the front of their block tables (reference nanovllm/engine/block_manager.py:41-71, prefix
// cache hits), and prepare_decode ships those tables unchanged (engine/model_runner.py:205-226).
// A decode batch of B rows over a common P-page prefix therefore re-reads the same K/V pages B
// times. The cascade computes the shared part ONCE per KV head as a dense problem
// (B*G query rows x P*block_size keys, tcgen05 kernel in prefill_tc.cu), the per-row remainder with
// the streaming split-KV kernel (decode_attn.cu) and merges the two by log-sum-exp.
//
// Everything is decided on the device so the sequence stays CUDA-graph capturable
// (model_runner.py:256-297: under replay block_tables / context_lens are device-only):
//   cascade_plan_kernel  one CTA; plan = {P, b0, active_rows, 0} with
//       P  = number of leading pages whose block id is identical in every active row
//            (context_lens > 0) and that lie entirely before every row's newest token,
//       b0 = first active row (its table names the shared pages).
// The later kernels read P from `plan` and size their work from it; P == 0 degrades to the plain
// split-KV path (the prefix kernel exits at once).
#include "common.h"

#include <climits>
#include <cstdlib>

namespace b200 {

namespace {

constexpr int kPlanThreads = 1024;

__global__ void __launch_bounds__(kPlanThreads)
cascade_plan_kernel(const int32_t* __restrict__ block_tables, int64_t table_row_stride,
                    int32_t table_width, const int32_t* __restrict__ context_lens, int32_t batch,
                    int32_t block_size, int32_t* __restrict__ plan) {
    __shared__ int s_b0, s_pmax, s_active, s_mismatch;
    if (threadIdx.x == 0) {
        s_b0 = INT_MAX;
        s_pmax = INT_MAX;
        s_active = 0;
        s_mismatch = INT_MAX;
    }
    __syncthreads();
    for (int b = threadIdx.x; b < batch; b += kPlanThreads) {
        const int ctx = context_lens[b];
        if (ctx > 0) {
            atomicMin(&s_b0, b);
            atomicMin(&s_pmax, (ctx - 1) / block_size);  // pages entirely before token ctx-1
            atomicAdd(&s_active, 1);
        }
    }
    __syncthreads();
    const int active = s_active;
    const int pmax = min(s_pmax, table_width);
    if (active < 2 || pmax <= 0) {  // nothing to share (uniform for the CTA)
        if (threadIdx.x == 0) {
            plan[0] = 0;
            plan[1] = 0;
            plan[2] = active;
            plan[3] = 0;
        }
        return;
    }
    const int32_t* t0 = block_tables + static_cast<int64_t>(s_b0) * table_row_stride;
    const int64_t total = static_cast<int64_t>(batch) * pmax;
    for (int64_t idx = threadIdx.x; idx < total; idx += kPlanThreads) {
        const int b = static_cast<int>(idx / pmax);
        const int j = static_cast<int>(idx % pmax);
        if (j >= *reinterpret_cast<volatile int*>(&s_mismatch)) continue;
        if (context_lens[b] <= 0) continue;  // padding row: its table is stale
        if (block_tables[static_cast<int64_t>(b) * table_row_stride + j] != t0[j])
            atomicMin(&s_mismatch, j);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        plan[0] = min(pmax, s_mismatch);
        plan[1] = s_b0;
        plan[2] = active;
        plan[3] = 0;
    }
}

// B200ATTN_CASCADE=0 never, =1 whenever the shape allows (tests, A/B runs), unset: by size.
int cascade_mode_from_env() {
    static int mode = [] {
        const char* e = std::getenv("B200ATTN_CASCADE");
        if (!e || !e[0]) return -1;
        return e[0] == '0' ? 0 : 1;
    }();
    return mode;
}

}  // namespace

cudaError_t launch_cascade_plan(const int32_t* block_tables, int64_t table_row_stride,
                                int32_t table_width, const int32_t* context_lens, int32_t batch,
                                int32_t block_size, int32_t* plan, cudaStream_t stream) {
    cascade_plan_kernel<<<1, kPlanThreads, 0, stream>>>(block_tables, table_row_stride, table_width,
                                                        context_lens, batch, block_size, plan);
    count_launch();
    return cudaGetLastError();
}

int cascade_pick_chunks(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                        int32_t table_width, int32_t block_size, int32_t sm_count) {
    // Host-visible sizes only. The prefix pass costs one extra (possibly empty) launch per layer,
    // so by default it is only armed where a shared prefix can pay for it: a batch that fills
    // 128-row tensor-core tiles and tables that can hold long contexts.
    const int mode = cascade_mode_from_env();
    if (mode == 0 || batch < 2 || batch > kDecodeMaxBatch || block_size % 128 != 0 ||
        num_kv_heads <= 0)
        return 0;
    const int group = num_q_heads / num_kv_heads;
    if (group < 1 || 128 % group != 0) return 0;
    const int64_t capacity = static_cast<int64_t>(table_width) * block_size;
    if (mode < 0 && (batch < 64 || capacity < 8192)) return 0;
    if (table_width < 2) return 0;  // one page per row: nothing can lie before the newest token
    const int64_t row_tiles = (static_cast<int64_t>(batch) * group + 255) / 256;
    const int64_t ctas = row_tiles * num_kv_heads;
    // The prefix CTAs run one per SM: the largest chunk count that still fits ONE wave (a second,
    // partly filled wave costs more than shorter chunks save; profiles/r01_cascade_chunks.log).
    int64_t chunks = sm_count / ctas;
    const int64_t max_tiles = (capacity - 1) / 128;
    if (chunks > max_tiles) chunks = max_tiles;
    if (chunks > 16) chunks = 16;
    if (chunks < 1) chunks = 1;
    return static_cast<int>(chunks);
}

}  // namespace b200
