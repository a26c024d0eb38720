/*
 * b200attn.h — C ABI of the B200-native attention hot path (libb200attn.so).
 *
 * This is the drop-in boundary for nanovllm's attention path. Every entry point names the
 * reference interface it replaces (paths relative to the reference tree):
 *
 *   b200attn_store_kvcache     nanovllm/layers/attention.py:10-49   (Triton store_kvcache_kernel
 *                                                                     + its Python launcher)
 *   b200attn_decode            nanovllm/layers/attention.py:87-96   (flash_attn_with_kvcache:
 *                                                                     paged single-query attention)
 *   b200attn_store_decode      nanovllm/layers/attention.py:70-71 + :87-96 fused in one call
 *   b200attn_prefill_varlen    nanovllm/layers/attention.py:72-86   (flash_attn_varlen_func:
 *                                                                     dense and paged-prefix)
 *   b200attn_cascade_*,        same decode call site (:87-96) for batches whose block tables
 *   b200attn_store_decode_cascade   start with identical block ids — the prefix-cache hits
 *                              nanovllm/engine/block_manager.py:41-71 produces
 *   b200attn_qk_norm_rope_store  nanovllm/models/qwen3.py:80-83 + layers/attention.py:70-71:
 *                              q_norm / k_norm (layers/layernorm.py:16-26), rotary_emb
 *                              (layers/rotary_embedding.py:37-48) and store_kvcache in one pass
 *                              over the QKV projection output
 *   b200attn_qkv_decode        the same four steps + the paged decode (:87-96) in ONE launch
 *   b200attn_prepare_decode    nanovllm/engine/model_runner.py:206-226 (_prepare_decode),
 *                              :152-158 (_prepare_block_tables) and the graph-variable refresh of
 *                              _run_model :242-254, expanded on the device
 *
 * Conventions
 *   - Plain pointers and sizes only; all buffers are DEVICE pointers owned by the caller (torch
 *     tensors on the Python side). The library never allocates or frees device memory, never
 *     synchronises the device and never reads device data on the host, so every call can be
 *     captured into a CUDA graph (nanovllm/engine/model_runner.py:256-297).
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream).
 *   - Return value: 0 on success; negative = argument / unsupported-shape error
 *     (B200ATTN_E_*); positive = cudaError_t of the failed launch. b200attn_last_error()
 *     returns a thread-local human-readable message for the last non-zero return.
 *   - dtype: 16-bit storage; `dtype` 0 = bfloat16, 1 = float16 (store, decode, cascade, prefill;
 *     fp32 accumulation and softmax in both). The fused q/k norm + RoPE entry points
 *     (b200attn_qk_norm_rope_store, b200attn_qkv_decode) restate the reference's bf16 rounding
 *     points and are bf16 only.
 *   - There is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ATTN_ABI_VERSION 10

#define B200ATTN_DTYPE_BF16 0
#define B200ATTN_DTYPE_F16 1

#define B200ATTN_E_BADARG (-1)      /* null pointer / negative size / misalignment           */
#define B200ATTN_E_UNSUPPORTED (-2) /* head_dim != 128, dtype, block_size % 64, G > 8 ...     */
#define B200ATTN_E_WORKSPACE (-3)   /* workspace too small for the requested split count      */
#define B200ATTN_E_DRIVER (-4)      /* cuTensorMapEncodeTiled unavailable / failed            */

/* ABI version of the loaded library (== B200ATTN_ABI_VERSION it was built with). */
int b200attn_abi_version(void);

/* Thread-local message for the last failing call on this thread ("" if none). */
const char* b200attn_last_error(void);

/* Number of SMs of the current device (cached); <= 0 when no CUDA device is usable. */
int b200attn_sm_count(void);

/* Total number of kernel launches this library has issued in this process (all threads).
 * Instrumentation for bench.py's `gpu_launches`; has no reference counterpart. */
int64_t b200attn_launch_count(void);

/*
 * KV scatter. For every token i in [0, n_tokens) with slot_mapping[i] >= 0:
 *     k_cache[slot_mapping[i] * row_elems + j] = key  [i * key_row_stride   + j]
 *     v_cache[slot_mapping[i] * row_elems + j] = value[i * value_row_stride + j]   j < row_elems
 * Negative slots are skipped (the reference skips -1: attention.py:23-24). Bit-exact copy of
 * 16-bit elements; offsets are 64-bit (the reference's int32 `slot * D` overflow is not kept).
 *   row_elems        = num_kv_heads * head_dim (elements per token row, multiple of 8)
 *   *_row_stride     in elements (value is a strided view in the caller: models/qwen3.py:76-79)
 *   num_slots        = num_blocks * block_size, bound for slot_mapping (slots >= num_slots are
 *                      skipped rather than written out of bounds)
 */
int b200attn_store_kvcache(const void* key, int64_t key_row_stride, const void* value,
                           int64_t value_row_stride, void* k_cache, void* v_cache,
                           const int32_t* slot_mapping, int32_t n_tokens, int32_t row_elems,
                           int64_t num_slots, int32_t dtype, void* stream);

/*
 * Work partition of the decode kernel. The KV tiles (64 tokens) of the sequences of the batch,
 * plus a small fixed cost per sequence, form one unit stream per KV head. The kernel launches
 * `num_splits` persistent CTAs and cuts the stream into contiguous ranges ON THE DEVICE from
 * context_lens; a work item is (range, KV head), items with adjacent ids stream the same tokens of
 * adjacent KV heads side by side. Many rows (the longest at most B200ATTN_DECODE_DYN_PCT % of a
 * CTA group's share): a range is one whole sequence (no row is cut, nothing is merged) and CTAs
 * take items beyond their first from an atomic counter, so the faster SMs take more of them.
 * Few or long rows: one equal range per CTA group (num_splits /
 * num_kv_heads groups); a range boundary cuts at most one row and cut rows are finished by the
 * piece that arrives last (in-launch merge through the workspace). Both regimes need the counters
 * of b200attn_set_split_counters when num_splits > 1. Any num_splits in [1, 2048] gives the same
 * result within rounding (1 = one CTA walks the whole batch).
 *
 * b200attn_decode_pick_splits: host-only choice of num_splits from host-visible sizes (graph-
 * capture rule: context_lens is device-only under replay, model_runner.py:242-254): two CTAs per
 * SM, fewer when even full tables could not give every CTA its minimum amount of work.
 */
int b200attn_decode_pick_splits(int32_t batch, int32_t num_kv_heads, int32_t table_width,
                                int32_t block_size, int32_t sm_count);

/*
 * Host-only introspection of that partition (no GPU needed; runs the kernel's own partition code on
 * the host): for HOST arrays context_lens[batch] and `shared_pages` leading pages skipped by the
 * cascade, writes up to max_pieces records of 12 int32
 *     {item, sequence, kv_head, tile_begin, tile_end, tiles_total, first, last, range_first,
 *      range_last, range, units_per_range}
 * in work-item order and returns the number of pieces (which may exceed max_pieces), or a
 * negative error. Field 0 is the work item (= range * num_kv_heads + kv_head; items below
 * num_splits are the CTAs' first items, later ones are taken dynamically), field 10 its range,
 * field 11 the units per range with bit 30 set in the dynamic regime. Tiles are 64 tokens. Uses
 * the same B200ATTN_DECODE_MIN_TILES / _ROW_COST / _SNAP / _DYNAMIC / _DYN_PCT tunables.
 */
int b200attn_decode_partition_host(const int32_t* context_lens, int32_t batch,
                                   int32_t num_kv_heads, int32_t num_splits, int32_t shared_pages,
                                   int32_t block_size, int32_t* pieces, int32_t max_pieces);

/*
 * Overlap mode of the decode entry points (process-wide switch, off by default; returns the
 * previous setting). When on, a decode launch is a programmatic dependent (PDL) of the kernel that
 * precedes it in the stream: the kernel's prologue - the scan of context_lens, its first work item
 * and that item's block-table window - runs while the preceding kernel drains, and
 * griddepcontrol.wait comes before q, the new K/V rows and the cache are read and before anything
 * is written. One exception: the K/V tiles of a row that lie entirely before its newest token
 * (position context_lens[b] - 1) are requested before the wait, so that the ring is full when the
 * preceding kernel has drained. THE CALLER PROMISES that the kernel immediately before the decode
 * launch in the stream writes neither context_lens / block_tables (in nanovllm they are written
 * once per step, model_runner.py:206-226, many kernels earlier) nor any cache row other than the
 * rows' newest tokens (in nanovllm that kernel is the qkv GEMM or store_kvcache of the same
 * layer, attention.py:67-71, which writes exactly those); byox_vllm_b200.install() switches
 * overlap on.
 * Ignored for launches with a cascade plan (those are dependents of the shared-prefix pass).
 * Every decode launch also signals launch_dependents at its start, so consecutive decode launches
 * (one per layer) overlap tail and prologue.
 */
int b200attn_set_decode_overlap(int on);

/* Debug only: with a device buffer of 16 uint64 per CTA registered, the decode entry points launch
 * an instrumented build of the kernel that stamps its phases there (tools/trace_decode.py);
 * NULL (the default) restores the product kernel. Not thread-safe, not for production. */
void b200attn_debug_decode_trace(void* stamps);

/* Debug only: with a device buffer of 8 CTAs x 6 streams x 128 steps x 4 uint64 registered,
 * b200attn_prefill_varlen launches an instrumented build of the tcgen05 kernel whose first 8 CTAs
 * stamp their SM clock at the phase boundaries of every tile step (tools/trace_prefill.py); NULL
 * (the default) restores the product kernel. Not thread-safe, not for production. */
void b200attn_debug_prefill_trace(void* stamps);

/* Bytes of fp32 workspace the decode entry points need: two partial-piece slots per work item that
 * can hold a cut row when num_splits > 1 (at most num_splits + num_q_heads items), plus batch * num_q_heads * prefix_chunks shared-prefix partials for the cascade
 * (prefix_chunks = 0 otherwise). 0 when num_splits <= 1 and prefix_chunks == 0. */
size_t b200attn_decode_workspace_bytes(int32_t batch, int32_t num_q_heads, int32_t head_dim,
                                       int32_t num_splits, int32_t prefix_chunks);

/*
 * Paged decode attention (one query token per sequence):
 *   out[b, h, :] = softmax(scale * q[b, h, :] . K[b, :ctx_b, h / G, :]^T) . V[b, :ctx_b, h / G, :]
 * K/V rows of sequence b, token t live at cache slot
 *   block_tables[b * table_row_stride + t / block_size] * block_size + t % block_size,
 * cache layout [num_blocks, block_size, num_kv_heads, head_dim] contiguous. ctx_b =
 * context_lens[b] (includes the token being decoded). Rows with ctx_b <= 0 produce zeros and
 * read no table entry; table entries at or beyond ceil(ctx_b / block_size) are never read.
 *   q   [batch, num_q_heads, head_dim] contiguous;  out same shape/dtype
 *   workspace: >= b200attn_decode_workspace_bytes(...) bytes when num_splits > 1, else may be NULL
 *   num_splits > 1 needs the arrival counters of b200attn_set_split_counters (else
 *   B200ATTN_E_WORKSPACE). batch may exceed 1024: the call then issues one launch per 1024 rows.
 */
int b200attn_decode(const void* q, const void* k_cache, const void* v_cache,
                    const int32_t* block_tables, int64_t table_row_stride, int32_t table_width,
                    const int32_t* context_lens, void* out, void* workspace, size_t workspace_bytes,
                    int32_t batch, int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                    int32_t block_size, int64_t num_blocks, float scale, int32_t num_splits,
                    int32_t dtype, void* stream);

/*
 * Split paged decode for callers that have no device memory to give. No counterpart in the
 * reference (flash_attn_with_kvcache allocates its split-KV accumulators through torch itself).
 *
 * b200attn_decode with num_splits > 1 needs two pieces of scratch from its caller: the fp32
 * workspace for the partial results of cut rows, and the zeroed counters registered with
 * b200attn_set_split_counters. Some callers cannot supply them. An engine that has already handed
 * every free byte of HBM to its KV cache (nanovllm sizes the cache from what is left after warm-up,
 * engine/model_runner.py:114-140) gets an out-of-memory error from any later allocation. A plugin
 * called inside another framework's stream capture must not allocate at all. A binding from a
 * language without a device allocator (cgo, JNI, N-API) has only the buffers it was handed. For
 * them this entry point computes exactly what b200attn_decode computes (same arguments, minus the
 * workspace pair; same kernel) with the scratch taken from an array built into the library image.
 * It calls no cudaMalloc / cudaFree and no allocating library routine.
 *
 * Limits (checked; beyond them the call fails with B200ATTN_E_WORKSPACE and a message that names
 * the sizes - use b200attn_decode with your own buffers then):
 *     b200attn_decode_workspace_bytes(batch, num_q_heads, head_dim, num_splits, 0)
 *                                        <= B200ATTN_STATIC_SCRATCH_PARTIAL_BYTES   (8 MiB)
 *     batch * num_kv_heads + 2           <= B200ATTN_STATIC_SCRATCH_COUNTERS        (4096;
 *                                        implied by the first limit today, checked all the same)
 * num_splits == 1 uses no scratch and has no limit. The price of the convenience: every GPU on
 * which the library is loaded carries the 8 MiB + 16 KiB array until the process exits, used or not.
 *
 * First call on a device: the library looks up the device address of the array (with that device
 * current; one driver query, result kept per device). Make that first call outside any stream
 * capture. Later calls only enqueue a 16 KiB memset of the counters and the kernel on `stream`.
 *
 * Sharing: there is ONE array per device. Every call on that device uses it, the counters are
 * cleared at the start of each call and the partials keep whatever the previous call left (the
 * kernel writes each slot before it reads it, so stale partials are harmless). Two calls whose GPU
 * work can overlap - different streams without an event between them, or two host threads - race on
 * it and produce wrong rows without any error. Keep calls on one device in stream order. The array
 * is separate from the buffer of b200attn_set_split_counters: this entry point neither needs nor
 * disturbs that registration.
 */
#define B200ATTN_STATIC_SCRATCH_PARTIAL_BYTES (8u << 20)
#define B200ATTN_STATIC_SCRATCH_COUNTERS 4096
int b200attn_decode_static_scratch(const void* q, const void* k_cache, const void* v_cache,
                                   const int32_t* block_tables, int64_t table_row_stride,
                                   int32_t table_width, const int32_t* context_lens, void* out,
                                   int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                                   int32_t head_dim, int32_t block_size, int64_t num_blocks,
                                   float scale, int32_t num_splits, int32_t dtype, void* stream);

/*
 * Fused decode step of one layer in ONE launch: scatter this step's K/V rows (as
 * b200attn_store_kvcache with n_tokens == batch) and run b200attn_decode over the updated cache.
 * Same results as the two calls in sequence under the precondition the reference's decode step
 * guarantees (engine/model_runner.py:211-215): slot_mapping[b] is negative (row skipped) or the
 * cache slot of token context_lens[b] - 1 of row b, i.e.
 *     block_tables[b][(ctx_b-1)/block_size] * block_size + (ctx_b-1) % block_size.
 * The CTA that attends that token writes it first (generic stores, fence.proxy.async, then its
 * own TMA reads), so no other work item ever reads a slot being written. A caller that breaks the
 * precondition (a slot that belongs to another token of the batch) gets the row stored, but rows
 * that attend that slot may read it before or after the store: use b200attn_store_kvcache followed
 * by b200attn_decode when slot_mapping is not the decode step's own.
 */
int b200attn_store_decode(const void* q, const void* key, int64_t key_row_stride, const void* value,
                          int64_t value_row_stride, const int32_t* slot_mapping, void* k_cache,
                          void* v_cache, const int32_t* block_tables, int64_t table_row_stride,
                          int32_t table_width, const int32_t* context_lens, void* out,
                          void* workspace, size_t workspace_bytes, int32_t batch,
                          int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                          int32_t block_size, int64_t num_blocks, float scale, int32_t num_splits,
                          int32_t dtype, void* stream);

/*
 * Shared-prefix ("cascade") decode. When many rows of a decode batch start with the same
 * physical pages (prefix-cache hits), the shared pages are attended once per KV head as a dense
 * tensor-core problem (all batch * G query rows against P * block_size keys) instead of once per
 * row, the remainder of every row is streamed as in b200attn_decode, and the two are merged by
 * log-sum-exp. Results equal b200attn_store_decode within the same tolerance.
 *
 *   b200attn_cascade_pick_chunks  host-only: number of split-KV chunks for the shared-prefix pass,
 *       or 0 when the cascade should not be armed for this shape (small batch / short tables /
 *       block_size % 128 != 0; B200ATTN_CASCADE=0|1 forces it off / on).
 *   b200attn_cascade_plan  one tiny launch per decode STEP (the plan depends only on the
 *       metadata, so every layer of the step reuses it): writes plan[0] = P, the number of leading
 *       pages whose block id is identical in every row with context_lens > 0 and that lie
 *       entirely before every such row's newest token (P = 0 with fewer than 2 active rows);
 *       plan[1] = first active row; plan[2] = active rows; plan[3] = 0. `plan` is a device int32[4].
 *   b200attn_store_decode_cascade  = b200attn_store_decode + (plan, prefix_chunks). The split
 *       counts and launch shapes depend on host-visible sizes only; P is read on the device, so the
 *       call is graph-capturable and P == 0 degrades to the plain path. The workspace must hold
 *       b200attn_decode_workspace_bytes(batch, num_q_heads, head_dim, num_splits, prefix_chunks).
 *       plan == NULL and prefix_chunks == 0 is exactly b200attn_store_decode. batch <= 1024.
 */
int b200attn_cascade_pick_chunks(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                                 int32_t table_width, int32_t block_size, int32_t sm_count);

int b200attn_cascade_plan(const int32_t* block_tables, int64_t table_row_stride,
                          int32_t table_width, const int32_t* context_lens, int32_t batch,
                          int32_t block_size, int32_t* plan, void* stream);

int b200attn_store_decode_cascade(const void* q, const void* key, int64_t key_row_stride,
                                  const void* value, int64_t value_row_stride,
                                  const int32_t* slot_mapping, void* k_cache, void* v_cache,
                                  const int32_t* block_tables, int64_t table_row_stride,
                                  int32_t table_width, const int32_t* context_lens, void* out,
                                  void* workspace, size_t workspace_bytes, int32_t batch,
                                  int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                                  int32_t block_size, int64_t num_blocks, float scale,
                                  int32_t num_splits, const int32_t* plan, int32_t prefix_chunks,
                                  int32_t dtype, void* stream);

/*
 * Varlen causal prefill (bottom-right aligned causal mask, as flash_attn_varlen_func):
 * query row i of sequence s (i < Lq_s = cu_seqlens_q[s+1] - cu_seqlens_q[s]) attends keys
 * 0 .. Lk_s - Lq_s + i of that sequence (Lk_s = cu_seqlens_k[s+1] - cu_seqlens_k[s]).
 *   dense  (block_tables == NULL): K/V rows are key/value[cu_seqlens_k[s] + t] with the given
 *          row strides (elements); k_cache/v_cache ignored.
 *   paged  (block_tables != NULL): K/V rows come from the cache through block_tables as in
 *          b200attn_decode; key/value ignored.
 *   q [total_q, num_q_heads, head_dim] contiguous; out same. Rows with no visible key get zeros.
 *   total_k = number of rows of key/value (dense; bounds the TMA tensor map), ignored when paged.
 * Kernel choice depends on host-visible sizes only: the tcgen05/TMEM kernel (rows = query token x
 * GQA group head) whenever pages are multiples of 128 tokens, the group divides 128 and the call
 * has at most 1023 sequences, else the mma.sync kernel (B200ATTN_PREFILL=tc|mma forces one).
 * The tcgen05 kernel is persistent: its grid is sized from total_q / num_seqs (an upper bound of
 * the 256-row work items), the items are ordered by cost on the device from cu_seqlens and the
 * running CTAs take them through the hardware work queue (clusterlaunchcontrol.try_cancel):
 * nothing to allocate, no counter to reset, CUDA-graph capturable like every other entry point.
 */
int b200attn_prefill_varlen(const void* q, const void* key, int64_t key_row_stride,
                            const void* value, int64_t value_row_stride, const void* k_cache,
                            const void* v_cache, const int32_t* block_tables,
                            int64_t table_row_stride, int32_t table_width,
                            const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                            int32_t num_seqs, int32_t total_q, int32_t total_k,
                            int32_t max_seqlen_q, int32_t max_seqlen_k, void* out,
                            int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                            int64_t num_blocks, float scale, int32_t dtype, void* stream);

/*
 * The item schedule of the persistent tcgen05 prefill kernel, computed on the HOST from host copies
 * of cu_seqlens_q / cu_seqlens_k with the kernel's own code (csrc/prefill_plan.cuh): no GPU
 * needed (tests/test_prefill_schedule.py). items: 8 int32 per item
 * {sequence, 256-row pair, first query token of sub-tile 0, of sub-tile 1, KV tiles of sub-tile 0,
 * of sub-tile 1, cost, 1 if cost-sorted} in the order the CTAs take them, per KV head. Returns
 * the number of items per KV head (which may exceed max_items: only max_items are written), or a
 * negative B200ATTN_E_* code.
 */
int b200attn_prefill_schedule_host(const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                                   int32_t num_seqs, int32_t num_q_heads, int32_t num_kv_heads,
                                   int32_t* items, int32_t max_items);

/*
 * Fused q/k RMSNorm + neox RoPE + KV scatter over the raw QKV projection output (the steps of
 * nanovllm/models/qwen3.py:75-84 that precede the attention call). Row i of `qkv` holds
 * num_q_heads q heads, then num_kv_heads k heads, then num_kv_heads v heads of head_dim bf16.
 * For every token i and head row x of 128 elements:
 *     q/k rows:  x32 = float(x) / sqrt(mean(x32^2) + eps); xw = bf16(bf16(x32) * weight)
 *                (skipped when both weights are NULL: models with qkv_bias, qwen3.py:66-68,80-82)
 *                y[j]      = xw[j] * cos[j] - xw[j + 64] * sin[j]
 *                y[j + 64] = xw[j + 64] * cos[j] + xw[j] * sin[j]      j < 64, fp32, then bf16
 *                with cos[j] = cos_sin_cache[positions[i] * 128 + j], sin[j] = ...[... + 64 + j]
 *                (the layout RotaryEmbedding builds: rotary_embedding.py:27-35)
 *     q -> q_out[i, h, :]  (contiguous [n_tokens, num_q_heads, 128])
 *     k -> k_out[i * k_out_row_stride + h * 128 ...] when k_out != NULL (may alias the k part of
 *          qkv: dense prefill reads K/V through strided views of qkv), and, when slot_mapping !=
 *          NULL and slot_mapping[i] in [0, num_slots), -> k_cache[slot, h, :]
 *     v -> v_cache[slot, h, :] (bit-exact copy) under the same slot condition
 * positions are clamped to [0, max_position). Strides are in elements.
 */
int b200attn_qk_norm_rope_store(const void* qkv, int64_t qkv_row_stride, const int64_t* positions,
                                const void* q_norm_weight, const void* k_norm_weight, float eps,
                                const float* cos_sin_cache, int32_t max_position, void* q_out,
                                void* k_out, int64_t k_out_row_stride, void* k_cache,
                                void* v_cache, const int32_t* slot_mapping, int32_t n_tokens,
                                int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                                int64_t num_slots, int32_t dtype, void* stream);

/*
 * Whole decode step of one layer between qkv_proj and o_proj in ONE launch: the decode kernel of
 * b200attn_store_decode reads q and this step's K/V rows straight from `qkv` (batch rows, layout
 * as above), applies q_norm + RoPE to the q heads while loading them into registers and k_norm +
 * RoPE to the new K row while scattering it, then attends over the updated cache. Equal to
 * b200attn_qk_norm_rope_store followed by b200attn_decode. Same precondition on slot_mapping as
 * b200attn_store_decode. Not combinable with the cascade (run the two-call sequence instead).
 */
int b200attn_qkv_decode(const void* qkv, int64_t qkv_row_stride, const int64_t* positions,
                        const void* q_norm_weight, const void* k_norm_weight, float eps,
                        const float* cos_sin_cache, int32_t max_position,
                        const int32_t* slot_mapping, void* k_cache, void* v_cache,
                        const int32_t* block_tables, int64_t table_row_stride, int32_t table_width,
                        const int32_t* context_lens, void* out, void* workspace,
                        size_t workspace_bytes, int32_t batch, int32_t num_q_heads,
                        int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                        int64_t num_blocks, float scale, int32_t num_splits, int32_t dtype,
                        void* stream);

/*
 * Decode-step metadata expansion. `packed` is a DEVICE int32 array the caller
 * filled with one host->device copy:
 *     last_token[batch] | num_tokens[batch] | table_len[batch] | blocks[sum(table_len)]
 * (blocks = every sequence's block_table back to back). One launch writes, bit-identically to
 * what ModelRunner._prepare_decode + _run_model leave in the CUDA-graph input buffers:
 *     input_ids[b]    = last_token[b]                       (int64)      b < batch
 *     positions[b]    = num_tokens[b] - 1                   (int64)
 *     context_lens[b] = num_tokens[b]
 *     slot_mapping[b] = block_table_b[-1] * block_size + last_block_num_tokens_b - 1
 *                       with last_block_num_tokens = num_tokens - (ceil(num_tokens / block_size)
 *                       - 1) * block_size          (engine/sequence.py:62-68, model_runner.py:215)
 *     block_tables[b, j] = block_table_b[j] for j < table_len[b], -1 for table_len[b] <= j <
 *                       table_width; columns >= table_width are left untouched
 *     slot_mapping[b] = -1, context_lens[b] = 0             batch <= b < capacity
 * batch <= 1024, table_len[b] <= table_width. All output buffers hold `capacity` rows.
 */
int b200attn_prepare_decode(const int32_t* packed, int32_t batch, int32_t table_width,
                            int32_t block_size, int64_t* input_ids, int64_t* positions,
                            int32_t* slot_mapping, int32_t* context_lens, int32_t* block_tables,
                            int64_t table_row_stride, int32_t capacity, void* stream);

/*
 * Counters of the decode kernel. A (sequence, KV head) row whose tiles were cut across work items
 * is finished by the piece that arrives last (acq_rel atomic arrival counter, reset to zero before
 * the launch ends); the dynamic work distribution uses two more counters (next item, CTAs done;
 * the CTA that finishes last resets them). The caller registers a ZEROED device buffer of `count`
 * int32 for the current device once; every decode call with num_splits > 1 needs
 * min(batch, 1024) * num_kv_heads + 2 <= count (the last two entries of the buffer are the
 * scheduler's). The buffer stays owned by the caller and must outlive every later decode call on
 * that device (the library keeps the pointer: the one exception, besides cached TMA descriptors, to
 * "keeps no pointers"). Decode calls on one device must be stream-ordered with respect to each
 * other (one engine per device, as nanovllm runs): two decode launches in flight at once on
 * different streams would share the counters. A launch that is aborted mid-flight may leave
 * counters non-zero: re-zero and re-register the buffer. counters == NULL unregisters.
 */
int b200attn_set_split_counters(void* counters, int64_t count);

/*
 * Prefill-step metadata expansion: the per-token part of
 * ModelRunner._prepare_prefill (engine/model_runner.py:160-204) on the device. `packed` is a
 * DEVICE int32 array  num_tokens[S] | num_cached_tokens[S] | table_len[S] | blocks[sum(table_len)].
 * One launch writes, bit-identically to the reference's lists:
 *     positions    (int64)  concat over s of range(num_cached[s], num_tokens[s])
 *     cu_seqlens_q (int32)  [0, prefix sums of num_tokens - num_cached]      S + 1 entries
 *     cu_seqlens_k (int32)  [0, prefix sums of num_tokens]                   S + 1 entries
 *     slot_mapping (int32)  for every s with table_len[s] > 0 (warm-up sequences have no table,
 *                           :179-180): block_table[i] * block_size + j over blocks i from
 *                           num_cached[s] / block_size to the last one, j < block_size (or <
 *                           last_block_num_tokens in the last block, engine/sequence.py:58-68)
 * The caller sizes the outputs from host-visible sums: total_q = sum(num_tokens - num_cached),
 * total_slots = sum over s with a table of num_tokens[s] - (num_cached[s] / block_size) *
 * block_size. num_seqs <= 1024. max_seqlen_q / max_seqlen_k and input_ids stay on the host side.
 */
int b200attn_prepare_prefill(const int32_t* packed, int32_t num_seqs, int32_t block_size,
                             int64_t* positions, int32_t total_q, int32_t* slot_mapping,
                             int32_t total_slots, int32_t* cu_seqlens_q, int32_t* cu_seqlens_k,
                             void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
