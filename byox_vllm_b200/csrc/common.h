// Internal declarations shared by the kernel translation units and the C ABI (api.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>

#include "decode_plan.cuh"
#include "norm_rope.cuh"

namespace b200 {

constexpr int kHeadDim = 128;  // every BASELINE model (Qwen3 0.6B / 8B / 32B) has head_dim 128

// Launch accounting (b200attn_launch_count): every kernel launch of the library calls this once.
void count_launch(int n = 1);

// SM count of the CURRENT device (cached per device ordinal: a process may drive several GPUs).
int current_device_sm_count();

// ---- store_kv.cu -----------------------------------------------------------------------------
cudaError_t launch_store_kvcache(const void* key, int64_t key_row_stride, const void* value,
                                 int64_t value_row_stride, void* k_cache, void* v_cache,
                                 const int32_t* slot_mapping, int32_t n_tokens, int32_t row_elems,
                                 int64_t num_slots, cudaStream_t stream);

// ---- qk_norm_rope.cu -------------------------------------------------------------------------
struct QkNormRopeParams {
    const void* qkv;         // [n_tokens, (Hq + 2 Hkv) * 128] bf16 rows: q heads | k heads | v heads
    int64_t qkv_row_stride;  // elements
    NormRopeArgs nr;
    void* q_out;             // [n_tokens, Hq, 128] contiguous
    void* k_out;             // nullable; rows [n_tokens, Hkv, 128], may alias the k part of qkv
    int64_t k_out_row_stride;
    void* k_cache;           // [num_slots, Hkv, 128]
    void* v_cache;
    const int32_t* slot_mapping;  // nullable => no cache scatter
    int64_t num_slots;
    int32_t n_tokens;
    int32_t num_q_heads;
    int32_t num_kv_heads;
};
cudaError_t launch_qk_norm_rope_store(const QkNormRopeParams& p, cudaStream_t stream);

// ---- prepare_meta.cu -------------------------------------------------------------------------
struct PrepareDecodeParams {
    const int32_t* packed;  // device: last_token[B] | num_tokens[B] | table_len[B] | blocks[sum]
    int64_t* input_ids;     // [capacity]
    int64_t* positions;     // [capacity]
    int32_t* slot_mapping;  // [capacity]
    int32_t* context_lens;  // [capacity]
    int32_t* block_tables;  // [capacity, table_row_stride]
    int64_t table_row_stride;
    int32_t batch;
    int32_t capacity;
    int32_t table_width;
    int32_t block_size;
};
cudaError_t launch_prepare_decode(const PrepareDecodeParams& p, cudaStream_t stream);

struct PreparePrefillParams {
    const int32_t* packed;  // device: num_tokens[S] | num_cached[S] | table_len[S] | blocks[sum]
    int64_t* positions;     // [sum(n - cached)]
    int32_t* slot_mapping;  // [sum over sequences with a table of n - (cached / bs) * bs]
    int32_t* cu_seqlens_q;  // [S + 1]
    int32_t* cu_seqlens_k;  // [S + 1]
    int32_t num_seqs;
    int32_t block_size;
};
cudaError_t launch_prepare_prefill(const PreparePrefillParams& p, int32_t total_hint,
                                   cudaStream_t stream);

// ---- decode_attn.cu --------------------------------------------------------------------------
struct DecodeParams {
    const void* q;                // [B, Hq, 128] bf16
    void* out;                    // [B, Hq, 128] bf16
    const int32_t* block_tables;  // [B, table_row_stride]
    const int32_t* context_lens;  // [B]
    // workspace (decode_workspace_layout): partial pieces of rows cut by a CTA-range boundary,
    // two slots per CTA (2c = tail of a row the previous CTA started, 2c+1 = head of a row the
    // next CTA continues), each [8 heads][128] normalised fp32 O + [8] log2-domain log-sum-exp
    float* piece_o;               // [2 * num_ctas, 8, 128]   (num_ctas > 1)
    float* piece_lse;             // [2 * num_ctas, 8]
    // shared-prefix partials written by the cascade prefix pass
    float* prefix_o;              // [B * Hq, prefix_chunks, 128]   (prefix_chunks > 0)
    float* prefix_lse;            // [B * Hq, prefix_chunks]
    int64_t table_row_stride;
    int32_t table_width;
    int32_t batch;
    int32_t num_q_heads;
    int32_t num_kv_heads;
    int32_t group;  // Hq / Hkv, 1..8
    int32_t block_size;
    int32_t num_ctas;   // persistent CTAs
    StreamTunables tun; // partition tunables (decode_plan.cuh)
    float scale_log2;   // softmax scale * log2(e)
    // optional fused scatter of this step's K/V rows (nullptr => no store)
    const void* new_key;
    const void* new_value;
    const int32_t* slot_mapping;
    void* k_cache;
    void* v_cache;
    int64_t key_row_stride;
    int64_t value_row_stride;
    int64_t num_slots;
    // optional shared-prefix cascade (cascade.cu): plan[0] = leading pages shared by every active
    // row (device-resident, written by cascade_plan_kernel); their attention was computed by the
    // prefix kernel into the prefix region of the workspace.
    const int32_t* plan;
    int32_t prefix_chunks;  // 0 => no cascade
    // optional fused q/k RMSNorm + RoPE (norm_rope.cuh): with nr.positions set, q
    // and new_key are raw QKV-projection outputs; q rows are normalised + rotated while they are
    // loaded into registers, the new K row while it is scattered (needs new_key / slot_mapping).
    int64_t q_row_stride;  // elements between q rows of consecutive sequences
    NormRopeArgs nr;
    // in-launch merge (num_ctas > 1): zeroed int32 arrival counters, one per (sequence, KV
    // head); the piece of a cut row that arrives last merges the partials itself and resets the
    // counter, so no merge kernel is launched. Required when num_ctas > 1.
    int32_t* split_counters;
    // rows engine: pieces per (sequence, KV head) row, grid = (KV heads, num_splits, batch)
    int32_t num_splits;
    // dynamic work distribution (num_ctas > 1): sched[0] = next work item beyond the first wave,
    // sched[1] = CTAs finished; both zero between launches (the last CTA resets them).
    int32_t* sched;
    // Overlap mode (b200attn_set_decode_overlap): the launch is a programmatic dependent of the
    // kernel before it in the stream; scan, first work item and its block-table window run while
    // that kernel drains, griddepcontrol.wait comes before anything else is read or written.
    // Requires that context_lens / block_tables are not written by that kernel. Ignored with the
    // cascade armed (the launch is then a dependent of the prefix pass).
    int32_t overlap;
    int32_t f16;  // 1: fp16 storage / operands (dtype code 1), 0: bf16
    // debug: 16 uint64 stamps per CTA (b200attn_debug_decode_trace); nullptr = the product kernel
    unsigned long long* trace;
};
cudaError_t launch_decode(const DecodeParams& p, const CUtensorMap& k_map, const CUtensorMap& v_map,
                          cudaStream_t stream);
constexpr int kDecodeMaxBatch = 1024;  // rows of one launch (the scan lives in shared memory)
size_t decode_workspace_bytes(int32_t batch, int32_t num_q_heads, int32_t num_ctas,
                              int32_t prefix_chunks);
void decode_workspace_layout(void* workspace, int32_t batch, int32_t num_q_heads, int32_t num_ctas,
                             int32_t prefix_chunks, DecodeParams* p);
int decode_pick_ctas(int32_t batch, int32_t num_kv_heads, int32_t table_width, int32_t block_size,
                     int32_t sm_count);
int decode_partition_host(const int32_t* context_lens, int32_t batch, int32_t num_kv_heads,
                          int32_t num_ctas, int32_t shared_pages, int32_t block_size,
                          int32_t* pieces, int32_t max_pieces);
// Two engines behind the same entry points (B200ATTN_DECODE_ENGINE = rows | stream, default rows):
//   rows   - one CTA per (KV head, split, sequence), hardware block scheduler, num_splits = equal
//            pieces per row picked on the host from sizes;
//   stream - num_splits persistent CTAs, device-built partition of the batch's tile stream.
bool decode_engine_rows();
int decode_min_units();
int decode_row_cost();
int decode_snap();
int decode_dyn_ok();
int decode_dyn_pct();
int decode_static_pct();
StreamTunables decode_tunables();

// ---- prefill_attn.cu -------------------------------------------------------------------------
struct PrefillParams {
    const void* q;  // [total_q, Hq, 128]
    void* out;      // [total_q, Hq, 128]
    // dense source
    const void* key;
    const void* value;
    int64_t key_row_stride;    // elements
    int64_t value_row_stride;  // elements
    // paged source
    const void* k_cache;
    const void* v_cache;
    const int32_t* block_tables;  // nullptr => dense
    int64_t table_row_stride;
    int32_t table_width;
    const int32_t* cu_seqlens_q;
    const int32_t* cu_seqlens_k;
    int32_t num_seqs;
    int32_t total_q;
    int32_t total_k;  // rows of key/value (dense source)
    int32_t max_seqlen_q;
    int32_t max_seqlen_k;
    int32_t num_q_heads;
    int32_t num_kv_heads;
    int32_t group;
    int32_t block_size;
    float scale_log2;
    // shared-prefix cascade mode (launch_cascade_prefix): q is the decode batch [batch, Hq, 128]
    const int32_t* plan;   // device int32[4]: {shared pages P, first active row b0, active rows, 0}
    float* ws_o;           // [batch * Hq, ws_slots, 128] fp32 normalised partials
    float* ws_lse;         // [batch * Hq, ws_slots] log2-domain log-sum-exp
    int32_t batch;
    int32_t prefix_chunks; // split-KV chunks over the shared prefix (grid.y)
    int32_t ws_slot0;      // chunk c writes slot ws_slot0 + c
    int32_t ws_slots;
    int32_t f16;           // 1: fp16 storage / operands (dtype code 1), 0: bf16
    int32_t sched_mode;    // prefill_tc.cu: 0 clc (work stealing), 1 static zig-zag, 2 one item per CTA
    // debug: [kTraceCtas][kTraceStreams][kTraceSteps][4] uint64 stamps (b200attn_debug_prefill_trace); nullptr =
    // the product kernel
    unsigned long long* trace;
};
cudaError_t launch_prefill(const PrefillParams& p, cudaStream_t stream);

// ---- prefill_tc.cu (tcgen05 / TMEM / TMA) ----------------------------------------------------
// q_map / k_map / v_map: 3-D tensor maps (128 dims, heads, rows), box 64 x 1 x 128, SWIZZLE_128B.
// Persistent over a cost-ordered item list every CTA builds in shared memory: at most kTcMaxSeqs
// sequences per call (more -> the caller takes the mma.sync shape fallback); batches with more
// than kTcMaxItems 256-row pairs per KV head run in descending index order instead of sorted.
constexpr int kTraceCtas = 8;
constexpr int kTraceStreams = 6;
constexpr int kTraceSteps = 128;
constexpr int kTcMaxSeqs = 1023;
constexpr int kTcMaxItems = 4096;
cudaError_t launch_prefill_tc(const PrefillParams& p, const CUtensorMap& q_map,
                              const CUtensorMap& k_map, const CUtensorMap& v_map,
                              cudaStream_t stream);

// Host mirror of the kernel's item schedule (prefill_plan.cuh), see prefill_tc.cu.
int prefill_schedule_host(const int32_t* cu_q, const int32_t* cu_k, int32_t num_seqs,
                          int32_t group, int32_t* items, int32_t max_items);

// Shared-prefix pass of the cascade decode: q_map is the decode batch q viewed as
// (128 dims, Hq, batch) with box 64 x G x 128/G; k_map / v_map as for launch_prefill_tc.
cudaError_t launch_cascade_prefix(const PrefillParams& p, const CUtensorMap& q_map,
                                  const CUtensorMap& k_map, const CUtensorMap& v_map,
                                  cudaStream_t stream);

// ---- cascade.cu ------------------------------------------------------------------------------
// Chunk geometry of the shared prefix, used identically by the prefix kernel (writer) and by the
// decode epilogue / merge kernel (readers): 128-token tiles of the P shared pages are cut into at
// most `chunks` contiguous ranges of `per` tiles; `valid` of them are non-empty.
struct PrefixChunks {
    int tiles;  // 128-token tiles in the shared prefix
    int per;    // tiles per chunk
    int valid;  // non-empty chunks
};
__host__ __device__ inline PrefixChunks prefix_chunks_of(int shared_pages, int block_size,
                                                         int chunks) {
    PrefixChunks c;
    c.tiles = shared_pages * (block_size / 128);
    c.per = chunks > 0 ? (c.tiles + chunks - 1) / chunks : 0;
    c.valid = c.per > 0 ? (c.tiles + c.per - 1) / c.per : 0;
    return c;
}
cudaError_t launch_cascade_plan(const int32_t* block_tables, int64_t table_row_stride,
                                int32_t table_width, const int32_t* context_lens, int32_t batch,
                                int32_t block_size, int32_t* plan, cudaStream_t stream);
int cascade_pick_chunks(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                        int32_t table_width, int32_t block_size, int32_t sm_count);

}  // namespace b200
