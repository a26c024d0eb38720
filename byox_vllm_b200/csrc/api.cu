// C ABI of libb200attn.so (declared in include/b200attn.h): argument checking, TMA tensor-map
// construction and kernel dispatch. No device allocation, no synchronisation, no host reads of
// device memory => every entry point is CUDA-graph capturable.
#include "../../include/b200attn.h"
#include "common.h"

#include <cuda_bf16.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <atomic>
#include <mutex>
#include <unordered_map>

namespace b200 {
static std::atomic<int64_t> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int current_device_sm_count() {
    static std::atomic<int> sms[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        (void)cudaGetLastError();
        return 0;
    }
    int n = sms[dev].load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
            (void)cudaGetLastError();
            return 0;
        }
        sms[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}
}  // namespace b200

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(cudaError_t e, const char* what) {
    snprintf(g_err, sizeof(g_err), "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
    return static_cast<int>(e);
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda link) ----------
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault,
                                                &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            return static_cast<EncodeTiledFn>(nullptr);
        }
        return reinterpret_cast<EncodeTiledFn>(sym);
    }();
    return fn;
}

struct MapKey {
    const void* base;
    int64_t rows;
    int64_t row_stride;
    int32_t heads;
    int32_t box_rows;
    bool operator==(const MapKey& o) const {
        return base == o.base && rows == o.rows && row_stride == o.row_stride &&
               heads == o.heads && box_rows == o.box_rows;
    }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        size_t h = std::hash<const void*>()(k.base);
        h ^= std::hash<int64_t>()(k.rows) * 1315423911u;
        h ^= std::hash<int64_t>()(k.row_stride) * 2654435761u;
        h ^= (static_cast<size_t>(k.heads) << 20) ^ (static_cast<size_t>(k.box_rows) << 40);
        return h;
    }
};

// Rows of `heads` x 128 bf16 ([rows, heads, 128], row stride in elements) as a 3-D tensor map
// (dims fastest-first: 128, heads, rows), box = 64 dims x 1 head x box_rows rows, SWIZZLE_128B,
// out-of-bounds rows read as zero. Encoding is pure host work (no driver call touches the
// device), so it is legal during stream capture.
int encode_rows_map(const void* base, int32_t heads, int64_t rows, int64_t row_stride,
                    int32_t box_rows, CUtensorMap* out, int32_t box_heads = 1) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return fail(B200ATTN_E_DRIVER, "cuTensorMapEncodeTiled not available from the driver");
    const cuuint64_t dims[3] = {static_cast<cuuint64_t>(b200::kHeadDim),
                                static_cast<cuuint64_t>(heads), static_cast<cuuint64_t>(rows)};
    const cuuint64_t strides[2] = {static_cast<cuuint64_t>(b200::kHeadDim) * 2,
                                   static_cast<cuuint64_t>(row_stride) * 2};
    const cuuint32_t box[3] = {64, static_cast<cuuint32_t>(box_heads),
                               static_cast<cuuint32_t>(box_rows)};
    const cuuint32_t elem_strides[3] = {1, 1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims,
                     strides, box, elem_strides, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(B200ATTN_E_DRIVER, "cuTensorMapEncodeTiled failed with CUresult %d",
                    static_cast<int>(r));
    return 0;
}

// The paged cache [num_slots, Hkv, 128] keeps its address for the life of the engine (nanovllm
// allocates it once, model_runner.py:137), so its descriptors are cached per (view, box).
int get_cache_map(const void* base, int32_t num_kv_heads, int64_t num_slots, int32_t box_rows,
                  CUtensorMap* out) {
    static std::mutex mu;
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    const int64_t row_stride = static_cast<int64_t>(num_kv_heads) * b200::kHeadDim;
    const MapKey key{base, num_slots, row_stride, num_kv_heads, box_rows};
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
        *out = it->second;
        return 0;
    }
    CUtensorMap map;
    if (int rc = encode_rows_map(base, num_kv_heads, num_slots, row_stride, box_rows, &map))
        return rc;
    if (cache.size() > 4096) cache.clear();
    cache.emplace(key, map);
    *out = map;
    return 0;
}

// B200ATTN_PREFILL=tc|mma forces one prefill kernel (tests, microbench); default: by shape.
int prefill_mode_from_env() {
    static int mode = [] {
        const char* e = std::getenv("B200ATTN_PREFILL");
        if (!e) return 0;
        if (e[0] == 't') return 1;
        if (e[0] == 'm') return 2;
        return 0;
    }();
    return mode;
}

int check_attn_shape(int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                     int32_t block_size, int32_t dtype) {
    if (dtype != B200ATTN_DTYPE_BF16 && dtype != B200ATTN_DTYPE_F16)
        return fail(B200ATTN_E_UNSUPPORTED, "dtype %d unsupported (0 = bf16, 1 = fp16)", dtype);
    if (head_dim != b200::kHeadDim)
        return fail(B200ATTN_E_UNSUPPORTED, "head_dim %d unsupported (128 only)", head_dim);
    if (num_q_heads <= 0 || num_kv_heads <= 0 || num_q_heads % num_kv_heads != 0)
        return fail(B200ATTN_E_BADARG, "bad head counts Hq=%d Hkv=%d", num_q_heads, num_kv_heads);
    if (num_q_heads / num_kv_heads > 8)
        return fail(B200ATTN_E_UNSUPPORTED, "GQA group %d > 8 unsupported",
                    num_q_heads / num_kv_heads);
    if (block_size <= 0 || block_size % 64 != 0)
        return fail(B200ATTN_E_UNSUPPORTED, "block_size %d must be a positive multiple of 64",
                    block_size);
    return 0;
}

}  // namespace

// Caller-registered zeroed counters, per device: arrival counters of the in-launch merge of cut
// rows + the two counters of the dynamic work distribution (the last two of the buffer).
struct SplitCounters {
    int32_t* ptr = nullptr;
    int64_t count = 0;
};
static SplitCounters g_split_counters[64];
static unsigned long long* g_decode_trace = nullptr;  // b200attn_debug_decode_trace
static unsigned long long* g_prefill_trace = nullptr; // b200attn_debug_prefill_trace
static std::atomic<int> g_decode_overlap{0};          // b200attn_set_decode_overlap
static std::mutex g_split_counters_mu;

static int32_t* split_counters_for(int64_t need, int32_t** sched) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        (void)cudaGetLastError();
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_split_counters_mu);
    const SplitCounters& c = g_split_counters[dev];
    if (c.ptr == nullptr || c.count < need) return nullptr;
    *sched = c.ptr + c.count - 2;  // the last two counters drive the dynamic work distribution
    return c.ptr;
}

// Shared body of every decode entry point. `nr` != nullptr: q / key are raw QKV-projection rows
// (q rows q_row_stride elements apart) and the kernel applies q/k RMSNorm + RoPE itself.
static int store_decode_impl(const void* q, int64_t q_row_stride, const void* key,
                             int64_t key_row_stride, const void* value, int64_t value_row_stride,
                             const int32_t* slot_mapping, void* k_cache, void* v_cache,
                             const int32_t* block_tables, int64_t table_row_stride,
                             int32_t table_width, const int32_t* context_lens, void* out,
                             void* workspace, size_t workspace_bytes, int32_t batch,
                             int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                             int32_t block_size, int64_t num_blocks, float scale,
                             int32_t num_splits, const int32_t* plan, int32_t prefix_chunks,
                             int32_t dtype, void* stream, const b200::NormRopeArgs* nr,
                             int32_t* own_counters = nullptr, int64_t own_count = 0) {
    if (int rc = check_attn_shape(num_q_heads, num_kv_heads, head_dim, block_size, dtype)) return rc;
    if (nr && dtype != B200ATTN_DTYPE_BF16)
        return fail(B200ATTN_E_UNSUPPORTED, "the fused q/k norm + RoPE decode is bf16 only");
    if (prefix_chunks < 0 || prefix_chunks > 32 || (prefix_chunks > 0) != (plan != nullptr))
        return fail(B200ATTN_E_BADARG, "plan and prefix_chunks (0..32) must be given together");
    if (nr && prefix_chunks > 0)
        return fail(B200ATTN_E_UNSUPPORTED,
                    "qkv_decode cannot be combined with the cascade (run qk_norm_rope_store first)");
    if (prefix_chunks > 0 && (block_size % 128 != 0 || 128 % (num_q_heads / num_kv_heads) != 0))
        return fail(B200ATTN_E_UNSUPPORTED,
                    "cascade needs block_size %% 128 == 0 and a GQA group dividing 128");
    if (batch < 0 || num_blocks <= 0 || table_width < 0 || num_splits < 1)
        return fail(B200ATTN_E_BADARG, "negative or zero size");
    if (batch == 0) return 0;
    const bool rows_engine = b200::decode_engine_rows();
    if (num_splits > (rows_engine ? 1024 : 2048))
        return fail(B200ATTN_E_UNSUPPORTED, "num_splits %d too large", num_splits);
    if (rows_engine && batch > 65535)
        return fail(B200ATTN_E_UNSUPPORTED, "batch %d > 65535 rows per launch", batch);
    if (prefix_chunks > 0 && batch > b200::kDecodeMaxBatch)
        return fail(B200ATTN_E_UNSUPPORTED, "cascade decode handles at most %d rows, got %d",
                    b200::kDecodeMaxBatch, batch);
    if (!q || !k_cache || !v_cache || !block_tables || !context_lens || !out)
        return fail(B200ATTN_E_BADARG, "null pointer");
    if (!std::isfinite(scale)) return fail(B200ATTN_E_BADARG, "scale is not finite");
    const int64_t num_slots = num_blocks * block_size;
    if (num_slots >= (1LL << 31))
        return fail(B200ATTN_E_UNSUPPORTED, "cache of %lld slots exceeds int32 TMA coordinates",
                    static_cast<long long>(num_slots));
    if (!aligned16(q) || !aligned16(out) || !aligned16(k_cache) || !aligned16(v_cache))
        return fail(B200ATTN_E_BADARG, "pointers must be 16-byte aligned");
    if (q_row_stride % 8 != 0) return fail(B200ATTN_E_BADARG, "q rows must be 16-byte aligned");
    if (num_splits > 1 || prefix_chunks > 0) {
        const size_t need =
            b200::decode_workspace_bytes(batch, num_q_heads, num_splits, prefix_chunks);
        if (!workspace || workspace_bytes < need || !aligned16(workspace))
            return fail(B200ATTN_E_WORKSPACE,
                        "workspace %zu B < %zu B needed for %d CTA ranges + %d prefix chunks",
                        workspace_bytes, need, num_splits, prefix_chunks);
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool with_store = key || value || slot_mapping;
    // One launch: the decode kernel scatters the new rows itself (B200ATTN_FUSED_STORE=0 keeps the
    // two-launch sequence for A/B runs).
    static const bool fused_store = [] {
        const char* e = std::getenv("B200ATTN_FUSED_STORE");
        return !(e && e[0] == '0');
    }();
    if (with_store) {
        if (!key || !value || !slot_mapping)
            return fail(B200ATTN_E_BADARG, "store_decode needs key, value and slot_mapping");
        if (key_row_stride % 8 != 0 || value_row_stride % 8 != 0 || !aligned16(key) ||
            !aligned16(value))
            return fail(B200ATTN_E_UNSUPPORTED, "key/value rows must be 16-byte aligned");
        if (key_row_stride < num_kv_heads * head_dim || value_row_stride < num_kv_heads * head_dim)
            return fail(B200ATTN_E_BADARG, "row stride smaller than the row");
        if (nr && !fused_store)
            return fail(B200ATTN_E_UNSUPPORTED, "qkv_decode needs the fused store");
        if (!fused_store) {
            if (int rc = b200attn_store_kvcache(key, key_row_stride, value, value_row_stride,
                                                k_cache, v_cache, slot_mapping, batch,
                                                num_kv_heads * head_dim, num_slots, dtype, stream))
                return rc;
        }
    }
    CUtensorMap k_map, v_map;
    if (int rc = get_cache_map(k_cache, num_kv_heads, num_slots, 64, &k_map)) return rc;
    if (int rc = get_cache_map(v_cache, num_kv_heads, num_slots, 64, &v_map)) return rc;

    b200::DecodeParams p{};
    p.q = q;
    p.q_row_stride = q_row_stride;
    if (nr) p.nr = *nr;
    p.out = out;
    p.block_tables = block_tables;
    p.context_lens = context_lens;
    p.table_row_stride = table_row_stride;
    p.table_width = table_width;
    p.batch = batch;
    p.num_q_heads = num_q_heads;
    p.num_kv_heads = num_kv_heads;
    p.group = num_q_heads / num_kv_heads;
    p.block_size = block_size;
    p.num_ctas = num_splits;
    p.num_splits = num_splits;
    p.tun = b200::decode_tunables();
    if (num_splits == 1) p.tun.dyn_ok = 0;  // no counters: static regime (one CTA walks every item)
    p.trace = g_decode_trace;
    p.overlap = (g_decode_overlap.load(std::memory_order_relaxed) != 0 && plan == nullptr) ? 1 : 0;
    p.f16 = dtype == B200ATTN_DTYPE_F16 ? 1 : 0;
    p.scale_log2 = scale * 1.4426950408889634f;
    if (with_store && fused_store) {
        p.new_key = key;
        p.new_value = value;
        p.slot_mapping = slot_mapping;
        p.k_cache = k_cache;
        p.v_cache = v_cache;
        p.key_row_stride = key_row_stride;
        p.value_row_stride = value_row_stride;
        p.num_slots = num_slots;
    }
    b200::decode_workspace_layout(workspace, batch, num_q_heads, num_splits, prefix_chunks, &p);
    p.plan = plan;
    p.prefix_chunks = prefix_chunks;
    if (num_splits > 1) {
        const int64_t rows =
            (rows_engine || batch < b200::kDecodeMaxBatch) ? batch : b200::kDecodeMaxBatch;
        // layout of the registered buffer: [rows * Hkv arrival counters | ... | sched[2] at the end]
        if (own_counters != nullptr) {  // b200attn_decode_static_scratch: sizes already checked
            p.split_counters = own_counters;
            p.sched = own_counters + own_count - 2;
        } else {
            p.split_counters = split_counters_for(rows * num_kv_heads + 2, &p.sched);
        }
        if (p.split_counters == nullptr)
            return fail(B200ATTN_E_WORKSPACE,
                        "num_splits %d > 1 needs %lld zeroed counters registered with "
                        "b200attn_set_split_counters on this device",
                        num_splits, static_cast<long long>(rows * num_kv_heads + 2));
    }
    if (prefix_chunks > 0) {
        // Shared-prefix pass first (tcgen05): B*G query rows of every KV head against the P
        // shared pages, partials into the prefix region of the workspace. It never touches the
        // slot the fused scatter writes (that token lies after the prefix).
        const int group = num_q_heads / num_kv_heads;
        CUtensorMap q_map, k_map128, v_map128;
        if (int rc = encode_rows_map(q, num_q_heads, batch,
                                     static_cast<int64_t>(num_q_heads) * head_dim, 128 / group,
                                     &q_map, group))
            return rc;
        if (int rc = get_cache_map(k_cache, num_kv_heads, num_slots, 128, &k_map128)) return rc;
        if (int rc = get_cache_map(v_cache, num_kv_heads, num_slots, 128, &v_map128)) return rc;
        b200::PrefillParams pp{};
        pp.q = q;
        pp.k_cache = k_cache;
        pp.v_cache = v_cache;
        pp.block_tables = block_tables;
        pp.table_row_stride = table_row_stride;
        pp.table_width = table_width;
        pp.num_q_heads = num_q_heads;
        pp.num_kv_heads = num_kv_heads;
        pp.group = group;
        pp.block_size = block_size;
        pp.scale_log2 = p.scale_log2;
        pp.plan = plan;
        pp.f16 = dtype == B200ATTN_DTYPE_F16 ? 1 : 0;
        pp.ws_o = p.prefix_o;
        pp.ws_lse = p.prefix_lse;
        pp.batch = batch;
        pp.prefix_chunks = prefix_chunks;
        pp.ws_slot0 = 0;
        pp.ws_slots = prefix_chunks;
        cudaError_t e0 = b200::launch_cascade_prefix(pp, q_map, k_map128, v_map128, st);
        if (e0 != cudaSuccess) return cuda_fail(e0, "cascade prefix launch");
    }
    // The kernel keeps the scan of the batch in shared memory: larger batches (eager decode above
    // the graph sizes, model_runner.py:230) run as consecutive launches over row blocks.
    const int32_t rows_per_launch = rows_engine ? batch : b200::kDecodeMaxBatch;
    for (int32_t row0 = 0; row0 < batch; row0 += rows_per_launch) {
        b200::DecodeParams pc = p;
        const int32_t rows = batch - row0 < rows_per_launch ? batch - row0 : rows_per_launch;
        pc.batch = rows;
        pc.q = static_cast<const __nv_bfloat16*>(q) + static_cast<int64_t>(row0) * q_row_stride;
        pc.out = static_cast<__nv_bfloat16*>(out) +
                 static_cast<int64_t>(row0) * num_q_heads * head_dim;
        pc.block_tables = block_tables + static_cast<int64_t>(row0) * table_row_stride;
        pc.context_lens = context_lens + row0;
        if (pc.new_key) {
            pc.new_key = static_cast<const __nv_bfloat16*>(key) +
                         static_cast<int64_t>(row0) * key_row_stride;
            pc.new_value = static_cast<const __nv_bfloat16*>(value) +
                           static_cast<int64_t>(row0) * value_row_stride;
            pc.slot_mapping = slot_mapping + row0;
        }
        if (pc.nr.positions) pc.nr.positions += row0;
        cudaError_t e = b200::launch_decode(pc, k_map, v_map, st);
        if (e != cudaSuccess) return cuda_fail(e, "decode launch");
    }
    return 0;
}

// ---- b200attn_decode_static_scratch: split decode on library-owned scratch ------------------------
// The partial results of cut rows and the arrival / scheduler counters of a split decode launch
// live in this file-scope array instead of caller-provided buffers. It is part of the module image:
// every device that loads libb200attn.so holds these 8 MiB + 16 KiB from module load to process
// exit, whether or not the entry point is ever called.
namespace {
constexpr size_t kStaticPartialBytes = B200ATTN_STATIC_SCRATCH_PARTIAL_BYTES;
constexpr int64_t kStaticCounters = B200ATTN_STATIC_SCRATCH_COUNTERS;
constexpr size_t kStaticScratchBytes = kStaticPartialBytes + kStaticCounters * sizeof(int32_t);
}  // namespace
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// `g_static_scratch` names a host-side shadow here, not device memory: the address a kernel can
// use has to be asked for, per device (each device has its own copy of the array), with that
// device current. Resolved on the first call on a device and kept.
static unsigned char* static_scratch_for_current_device() {
    static std::mutex mu;
    static unsigned char* ptrs[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        cuda_fail(e, "cudaGetDevice");
        return nullptr;
    }
    if (dev < 0 || dev >= 64) {
        fail(B200ATTN_E_UNSUPPORTED, "device %d out of range", dev);
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (ptrs[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || p == nullptr) {
            cuda_fail(e, "cudaGetSymbolAddress(static decode scratch)");
            return nullptr;
        }
        ptrs[dev] = static_cast<unsigned char*>(p);
    }
    return ptrs[dev];
}

extern "C" int b200attn_decode_static_scratch(
    const void* q, const void* k_cache, const void* v_cache, const int32_t* block_tables,
    int64_t table_row_stride, int32_t table_width, const int32_t* context_lens, void* out,
    int32_t batch, int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
    int64_t num_blocks, float scale, int32_t num_splits, int32_t dtype, void* stream) {
    if (int rc = check_attn_shape(num_q_heads, num_kv_heads, head_dim, block_size, dtype)) return rc;
    if (batch < 0 || num_splits < 1) return fail(B200ATTN_E_BADARG, "negative or zero size");
    if (batch == 0) return 0;
    unsigned char* scratch = nullptr;
    if (num_splits > 1) {
        const size_t need = b200::decode_workspace_bytes(batch, num_q_heads, num_splits, 0);
        if (need > kStaticPartialBytes)
            return fail(B200ATTN_E_WORKSPACE,
                        "batch %d x %d q heads x %d splits needs %zu B of partials; the built-in "
                        "scratch holds %zu B (use b200attn_decode with a caller workspace)",
                        batch, num_q_heads, num_splits, need, kStaticPartialBytes);
        const int64_t counters = static_cast<int64_t>(batch) * num_kv_heads + 2;
        if (counters > kStaticCounters)
            return fail(B200ATTN_E_WORKSPACE,
                        "batch %d x %d KV heads needs %lld counters; the built-in scratch holds %lld "
                        "(use b200attn_decode with b200attn_set_split_counters)",
                        batch, num_kv_heads, static_cast<long long>(counters),
                        static_cast<long long>(kStaticCounters));
        scratch = static_scratch_for_current_device();
        if (scratch == nullptr) return B200ATTN_E_DRIVER;  // message already set
        // The counters must read zero when the kernel starts. The kernel leaves them at zero, but
        // an aborted launch (or a racing one) may not have: clear them in stream order each call.
        // The partials need no clearing (every slot is written before the merge reads it).
        cudaError_t e = cudaMemsetAsync(scratch + kStaticPartialBytes, 0,
                                        kStaticCounters * sizeof(int32_t),
                                        static_cast<cudaStream_t>(stream));
        if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(static decode counters)");
    }
    return store_decode_impl(q, static_cast<int64_t>(num_q_heads) * head_dim, nullptr, 0, nullptr,
                             0, nullptr, const_cast<void*>(k_cache), const_cast<void*>(v_cache),
                             block_tables, table_row_stride, table_width, context_lens, out,
                             scratch, scratch ? kStaticPartialBytes : 0, batch, num_q_heads,
                             num_kv_heads, head_dim, block_size, num_blocks, scale, num_splits,
                             nullptr, 0, dtype, stream, nullptr,
                             scratch ? reinterpret_cast<int32_t*>(scratch + kStaticPartialBytes)
                                     : nullptr,
                             kStaticCounters);
}


extern "C" {

int b200attn_abi_version(void) { return B200ATTN_ABI_VERSION; }

const char* b200attn_last_error(void) { return g_err; }

int64_t b200attn_launch_count(void) { return b200::g_launches.load(std::memory_order_relaxed); }

int b200attn_sm_count(void) { return b200::current_device_sm_count(); }

int b200attn_store_kvcache(const void* key, int64_t key_row_stride, const void* value,
                           int64_t value_row_stride, void* k_cache, void* v_cache,
                           const int32_t* slot_mapping, int32_t n_tokens, int32_t row_elems,
                           int64_t num_slots, int32_t dtype, void* stream) {
    if (dtype != B200ATTN_DTYPE_BF16 && dtype != B200ATTN_DTYPE_F16)  // a byte copy of 16-bit rows
        return fail(B200ATTN_E_UNSUPPORTED, "dtype %d unsupported (0 = bf16, 1 = fp16)", dtype);
    if (n_tokens < 0 || row_elems <= 0 || num_slots < 0)
        return fail(B200ATTN_E_BADARG, "negative size");
    if (n_tokens == 0) return 0;
    if (!key || !value || !k_cache || !v_cache || !slot_mapping)
        return fail(B200ATTN_E_BADARG, "null pointer");
    if (row_elems % 8 != 0 || key_row_stride % 8 != 0 || value_row_stride % 8 != 0)
        return fail(B200ATTN_E_UNSUPPORTED,
                    "row_elems/strides must be multiples of 8 elements (16-byte vectors)");
    if (!aligned16(key) || !aligned16(value) || !aligned16(k_cache) || !aligned16(v_cache))
        return fail(B200ATTN_E_BADARG, "pointers must be 16-byte aligned");
    if (key_row_stride < row_elems || value_row_stride < row_elems)
        return fail(B200ATTN_E_BADARG, "row stride smaller than the row");
    cudaError_t e = b200::launch_store_kvcache(key, key_row_stride, value, value_row_stride, k_cache,
                                               v_cache, slot_mapping, n_tokens, row_elems, num_slots,
                                               static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : cuda_fail(e, "store_kvcache launch");
}

int b200attn_decode_pick_splits(int32_t batch, int32_t num_kv_heads, int32_t table_width,
                                int32_t block_size, int32_t sm_count) {
    if (sm_count <= 0) sm_count = 148;
    return b200::decode_pick_ctas(batch, num_kv_heads, table_width, block_size, sm_count);
}

int b200attn_decode_partition_host(const int32_t* context_lens, int32_t batch,
                                   int32_t num_kv_heads, int32_t num_splits, int32_t shared_pages,
                                   int32_t block_size, int32_t* pieces, int32_t max_pieces) {
    if (batch < 0 || num_kv_heads <= 0 || num_splits < 1 || shared_pages < 0 ||
        block_size <= 0 || block_size % 64 != 0 || max_pieces < 0)
        return fail(B200ATTN_E_BADARG, "bad sizes");
    if (batch > 0 && !context_lens) return fail(B200ATTN_E_BADARG, "null context_lens");
    if (max_pieces > 0 && !pieces) return fail(B200ATTN_E_BADARG, "null pieces");
    return b200::decode_partition_host(context_lens, batch, num_kv_heads, num_splits, shared_pages,
                                       block_size, pieces, max_pieces);
}

int b200attn_prefill_schedule_host(const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                                   int32_t num_seqs, int32_t num_q_heads, int32_t num_kv_heads,
                                   int32_t* items, int32_t max_items) {
    if (num_seqs < 0 || num_q_heads <= 0 || num_kv_heads <= 0 || num_q_heads % num_kv_heads != 0 ||
        max_items < 0)
        return fail(B200ATTN_E_BADARG, "bad sizes");
    const int group = num_q_heads / num_kv_heads;
    if (group > 8 || 128 % group != 0)
        return fail(B200ATTN_E_UNSUPPORTED, "GQA group %d does not divide 128", group);
    if (num_seqs > 0 && (!cu_seqlens_q || !cu_seqlens_k))
        return fail(B200ATTN_E_BADARG, "null cu_seqlens");
    if (max_items > 0 && !items) return fail(B200ATTN_E_BADARG, "null items");
    return b200::prefill_schedule_host(cu_seqlens_q, cu_seqlens_k, num_seqs, group, items,
                                       max_items);
}

size_t b200attn_decode_workspace_bytes(int32_t batch, int32_t num_q_heads, int32_t head_dim,
                                       int32_t num_splits, int32_t prefix_chunks) {
    if (batch <= 0 || num_q_heads <= 0 || head_dim != b200::kHeadDim) return 0;
    return b200::decode_workspace_bytes(batch, num_q_heads, num_splits, prefix_chunks);
}

int b200attn_cascade_pick_chunks(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                                 int32_t table_width, int32_t block_size, int32_t sm_count) {
    if (sm_count <= 0) sm_count = 148;
    if (batch <= 0 || num_q_heads <= 0 || num_kv_heads <= 0 || table_width <= 0 || block_size <= 0)
        return 0;
    return b200::cascade_pick_chunks(batch, num_q_heads, num_kv_heads, table_width, block_size,
                                     sm_count);
}

int b200attn_cascade_plan(const int32_t* block_tables, int64_t table_row_stride,
                          int32_t table_width, const int32_t* context_lens, int32_t batch,
                          int32_t block_size, int32_t* plan, void* stream) {
    if (batch < 0 || table_width < 0 || block_size <= 0)
        return fail(B200ATTN_E_BADARG, "negative or zero size");
    if (!plan) return fail(B200ATTN_E_BADARG, "null plan");
    if (batch > 0 && (!block_tables || !context_lens)) return fail(B200ATTN_E_BADARG, "null pointer");
    cudaError_t e = b200::launch_cascade_plan(block_tables, table_row_stride, table_width,
                                              context_lens, batch, block_size, plan,
                                              static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : cuda_fail(e, "cascade_plan launch");
}

int b200attn_store_decode(const void* q, const void* key, int64_t key_row_stride, const void* value,
                          int64_t value_row_stride, const int32_t* slot_mapping, void* k_cache,
                          void* v_cache, const int32_t* block_tables, int64_t table_row_stride,
                          int32_t table_width, const int32_t* context_lens, void* out,
                          void* workspace, size_t workspace_bytes, int32_t batch,
                          int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                          int32_t block_size, int64_t num_blocks, float scale, int32_t num_splits,
                          int32_t dtype, void* stream) {
    return b200attn_store_decode_cascade(q, key, key_row_stride, value, value_row_stride,
                                         slot_mapping, k_cache, v_cache, block_tables,
                                         table_row_stride, table_width, context_lens, out,
                                         workspace, workspace_bytes, batch, num_q_heads,
                                         num_kv_heads, head_dim, block_size, num_blocks, scale,
                                         num_splits, nullptr, 0, dtype, stream);
}

int b200attn_store_decode_cascade(const void* q, const void* key, int64_t key_row_stride,
                                  const void* value, int64_t value_row_stride,
                                  const int32_t* slot_mapping, void* k_cache, void* v_cache,
                                  const int32_t* block_tables, int64_t table_row_stride,
                                  int32_t table_width, const int32_t* context_lens, void* out,
                                  void* workspace, size_t workspace_bytes, int32_t batch,
                                  int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                                  int32_t block_size, int64_t num_blocks, float scale,
                                  int32_t num_splits, const int32_t* plan, int32_t prefix_chunks,
                                  int32_t dtype, void* stream) {
    return store_decode_impl(q, static_cast<int64_t>(num_q_heads) * head_dim, key, key_row_stride,
                             value, value_row_stride, slot_mapping, k_cache, v_cache, block_tables,
                             table_row_stride, table_width, context_lens, out, workspace,
                             workspace_bytes, batch, num_q_heads, num_kv_heads, head_dim,
                             block_size, num_blocks, scale, num_splits, plan, prefix_chunks, dtype,
                             stream, nullptr);
}

int b200attn_decode(const void* q, const void* k_cache, const void* v_cache,
                    const int32_t* block_tables, int64_t table_row_stride, int32_t table_width,
                    const int32_t* context_lens, void* out, void* workspace, size_t workspace_bytes,
                    int32_t batch, int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                    int32_t block_size, int64_t num_blocks, float scale, int32_t num_splits,
                    int32_t dtype, void* stream) {
    return b200attn_store_decode(q, nullptr, 0, nullptr, 0, nullptr, const_cast<void*>(k_cache),
                                 const_cast<void*>(v_cache), block_tables, table_row_stride,
                                 table_width, context_lens, out, workspace, workspace_bytes, batch,
                                 num_q_heads, num_kv_heads, head_dim, block_size, num_blocks, scale,
                                 num_splits, dtype, stream);
}

int b200attn_prefill_varlen(const void* q, const void* key, int64_t key_row_stride,
                            const void* value, int64_t value_row_stride, const void* k_cache,
                            const void* v_cache, const int32_t* block_tables,
                            int64_t table_row_stride, int32_t table_width,
                            const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                            int32_t num_seqs, int32_t total_q, int32_t total_k,
                            int32_t max_seqlen_q, int32_t max_seqlen_k, void* out,
                            int32_t num_q_heads,
                            int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                            int64_t num_blocks, float scale, int32_t dtype, void* stream) {
    const bool paged = block_tables != nullptr;
    if (int rc = check_attn_shape(num_q_heads, num_kv_heads, head_dim, paged ? block_size : 64,
                                  dtype))
        return rc;
    if (num_seqs < 0 || total_q < 0 || total_k < 0 || max_seqlen_q < 0 || max_seqlen_k < 0)
        return fail(B200ATTN_E_BADARG, "negative size");
    if (num_seqs == 0 || total_q == 0 || max_seqlen_q == 0) return 0;
    if (!q || !out || !cu_seqlens_q || !cu_seqlens_k) return fail(B200ATTN_E_BADARG, "null pointer");
    if (!std::isfinite(scale)) return fail(B200ATTN_E_BADARG, "scale is not finite");
    if (paged) {
        if (!k_cache || !v_cache) return fail(B200ATTN_E_BADARG, "paged prefill needs the caches");
        if (num_blocks <= 0 || table_width <= 0)
            return fail(B200ATTN_E_BADARG, "paged prefill needs num_blocks/table_width > 0");
        if (!aligned16(k_cache) || !aligned16(v_cache))
            return fail(B200ATTN_E_BADARG, "caches must be 16-byte aligned");
    } else {
        if (!key || !value) return fail(B200ATTN_E_BADARG, "dense prefill needs key/value");
        if (key_row_stride % 8 != 0 || value_row_stride % 8 != 0 || !aligned16(key) ||
            !aligned16(value))
            return fail(B200ATTN_E_UNSUPPORTED, "key/value rows must be 16-byte aligned");
        if (key_row_stride < num_kv_heads * head_dim || value_row_stride < num_kv_heads * head_dim)
            return fail(B200ATTN_E_BADARG, "row stride smaller than the row");
    }
    if (!aligned16(q) || !aligned16(out))
        return fail(B200ATTN_E_BADARG, "q/out must be 16-byte aligned");

    b200::PrefillParams p{};
    p.q = q;
    p.out = out;
    p.key = key;
    p.value = value;
    p.key_row_stride = key_row_stride;
    p.value_row_stride = value_row_stride;
    p.k_cache = k_cache;
    p.v_cache = v_cache;
    p.block_tables = block_tables;
    p.table_row_stride = table_row_stride;
    p.table_width = table_width;
    p.cu_seqlens_q = cu_seqlens_q;
    p.cu_seqlens_k = cu_seqlens_k;
    p.num_seqs = num_seqs;
    p.total_q = total_q;
    p.total_k = total_k;
    p.max_seqlen_q = max_seqlen_q;
    p.max_seqlen_k = max_seqlen_k;
    p.num_q_heads = num_q_heads;
    p.num_kv_heads = num_kv_heads;
    p.group = num_q_heads / num_kv_heads;
    p.block_size = block_size;
    p.scale_log2 = scale * 1.4426950408889634f;
    p.f16 = dtype == B200ATTN_DTYPE_F16 ? 1 : 0;
    p.trace = g_prefill_trace;
    // Kernel choice from host-visible sizes only. The tcgen05 kernel works on pairs of 128-row
    // tiles of (token, group head) rows and 128-token KV tiles: it needs pages that are multiples
    // of 128 tokens and a GQA group that divides 128.
    const int mode = prefill_mode_from_env();
    const int group = num_q_heads / num_kv_heads;
    const bool tc_ok = scale > 0.f && (!paged || block_size % 128 == 0) && 128 % group == 0 &&
                       static_cast<int64_t>(num_blocks) * block_size < (1LL << 31) &&
                       num_seqs <= b200::kTcMaxSeqs;
    // With GQA-packed rows the tcgen05 kernel wins at every measured shape, short query blocks
    // included (profiles/r01_microbench_prefill_v4.log); mma.sync is the shape fallback only.
    bool use_tc = tc_ok;
    if (mode == 1) {
        if (!tc_ok) return fail(B200ATTN_E_UNSUPPORTED, "B200ATTN_PREFILL=tc: shape unsupported");
        use_tc = true;
    } else if (mode == 2) {
        use_tc = false;
    }
    cudaError_t e;
    if (use_tc) {
        CUtensorMap q_map, k_map, v_map;
        // GQA-packed query tiles: box = 64 dims x G heads x 128/G tokens
        if (int rc = encode_rows_map(q, num_q_heads, total_q,
                                     static_cast<int64_t>(num_q_heads) * head_dim, 128 / group,
                                     &q_map, group))
            return rc;
        if (paged) {
            const int64_t num_slots = num_blocks * block_size;
            if (int rc = get_cache_map(k_cache, num_kv_heads, num_slots, 128, &k_map)) return rc;
            if (int rc = get_cache_map(v_cache, num_kv_heads, num_slots, 128, &v_map)) return rc;
        } else {
            if (int rc = encode_rows_map(key, num_kv_heads, total_k, key_row_stride, 128, &k_map))
                return rc;
            if (int rc = encode_rows_map(value, num_kv_heads, total_k, value_row_stride, 128,
                                         &v_map))
                return rc;
        }
        e = b200::launch_prefill_tc(p, q_map, k_map, v_map, static_cast<cudaStream_t>(stream));
    } else {
        e = b200::launch_prefill(p, static_cast<cudaStream_t>(stream));
    }
    return e == cudaSuccess ? 0 : cuda_fail(e, "prefill launch");
}

// Validates the q/k-norm + RoPE arguments shared by the two fused entry points.
static int check_norm_rope(const int64_t* positions, const void* q_weight, const void* k_weight,
                           float eps, const float* cos_sin_cache, int32_t max_position,
                           b200::NormRopeArgs* out) {
    if (!positions || !cos_sin_cache) return fail(B200ATTN_E_BADARG, "null positions/cos_sin_cache");
    if ((q_weight == nullptr) != (k_weight == nullptr))
        return fail(B200ATTN_E_BADARG, "q_norm_weight and k_norm_weight must be given together");
    if (max_position <= 0) return fail(B200ATTN_E_BADARG, "max_position must be positive");
    if (!(eps >= 0.f) || !std::isfinite(eps)) return fail(B200ATTN_E_BADARG, "bad eps");
    if (!aligned16(cos_sin_cache) || (q_weight && (!aligned16(q_weight) || !aligned16(k_weight))))
        return fail(B200ATTN_E_BADARG, "weights / cos_sin_cache must be 16-byte aligned");
    out->positions = positions;
    out->q_weight = q_weight;
    out->k_weight = k_weight;
    out->cos_sin = cos_sin_cache;
    out->max_position = max_position;
    out->eps = eps;
    return 0;
}

int b200attn_qk_norm_rope_store(const void* qkv, int64_t qkv_row_stride, const int64_t* positions,
                                const void* q_norm_weight, const void* k_norm_weight, float eps,
                                const float* cos_sin_cache, int32_t max_position, void* q_out,
                                void* k_out, int64_t k_out_row_stride, void* k_cache,
                                void* v_cache, const int32_t* slot_mapping, int32_t n_tokens,
                                int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                                int64_t num_slots, int32_t dtype, void* stream) {
    if (dtype != B200ATTN_DTYPE_BF16)
        return fail(B200ATTN_E_UNSUPPORTED, "dtype %d unsupported (bf16 only)", dtype);
    if (head_dim != b200::kHeadDim)
        return fail(B200ATTN_E_UNSUPPORTED, "head_dim %d unsupported (128 only)", head_dim);
    if (n_tokens < 0 || num_q_heads <= 0 || num_kv_heads <= 0 || num_slots < 0)
        return fail(B200ATTN_E_BADARG, "negative or zero size");
    if (n_tokens == 0) return 0;
    if (static_cast<int64_t>(n_tokens) * (num_q_heads + 2 * num_kv_heads) >= (1LL << 31))
        return fail(B200ATTN_E_UNSUPPORTED, "n_tokens * heads exceeds 32-bit row indexing");
    b200::QkNormRopeParams p{};
    if (int rc = check_norm_rope(positions, q_norm_weight, k_norm_weight, eps, cos_sin_cache,
                                 max_position, &p.nr))
        return rc;
    if (!qkv || !q_out) return fail(B200ATTN_E_BADARG, "null qkv/q_out");
    const int64_t row = static_cast<int64_t>(num_q_heads + 2 * num_kv_heads) * head_dim;
    if (qkv_row_stride < row || qkv_row_stride % 8 != 0)
        return fail(B200ATTN_E_BADARG, "qkv_row_stride %lld < row of %lld elements or not 16-byte",
                    static_cast<long long>(qkv_row_stride), static_cast<long long>(row));
    if (k_out && (k_out_row_stride < num_kv_heads * head_dim || k_out_row_stride % 8 != 0))
        return fail(B200ATTN_E_BADARG, "bad k_out_row_stride");
    if (slot_mapping && (!k_cache || !v_cache))
        return fail(B200ATTN_E_BADARG, "slot_mapping given without caches");
    if (!aligned16(qkv) || !aligned16(q_out) || !aligned16(k_out) || !aligned16(k_cache) ||
        !aligned16(v_cache))
        return fail(B200ATTN_E_BADARG, "pointers must be 16-byte aligned");
    p.qkv = qkv;
    p.qkv_row_stride = qkv_row_stride;
    p.q_out = q_out;
    p.k_out = k_out;
    p.k_out_row_stride = k_out_row_stride;
    p.k_cache = k_cache;
    p.v_cache = v_cache;
    p.slot_mapping = slot_mapping;
    p.num_slots = num_slots;
    p.n_tokens = n_tokens;
    p.num_q_heads = num_q_heads;
    p.num_kv_heads = num_kv_heads;
    cudaError_t e = b200::launch_qk_norm_rope_store(p, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : cuda_fail(e, "qk_norm_rope_store launch");
}

int b200attn_qkv_decode(const void* qkv, int64_t qkv_row_stride, const int64_t* positions,
                        const void* q_norm_weight, const void* k_norm_weight, float eps,
                        const float* cos_sin_cache, int32_t max_position,
                        const int32_t* slot_mapping, void* k_cache, void* v_cache,
                        const int32_t* block_tables, int64_t table_row_stride, int32_t table_width,
                        const int32_t* context_lens, void* out, void* workspace,
                        size_t workspace_bytes, int32_t batch, int32_t num_q_heads,
                        int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                        int64_t num_blocks, float scale, int32_t num_splits, int32_t dtype,
                        void* stream) {
    if (batch == 0) return 0;
    b200::NormRopeArgs nr{};
    if (int rc = check_norm_rope(positions, q_norm_weight, k_norm_weight, eps, cos_sin_cache,
                                 max_position, &nr))
        return rc;
    if (!qkv || !slot_mapping) return fail(B200ATTN_E_BADARG, "null qkv/slot_mapping");
    if (num_q_heads <= 0 || num_kv_heads <= 0 || head_dim <= 0)
        return fail(B200ATTN_E_BADARG, "bad head counts");
    const int64_t row = static_cast<int64_t>(num_q_heads + 2 * num_kv_heads) * head_dim;
    if (qkv_row_stride < row || qkv_row_stride % 8 != 0)
        return fail(B200ATTN_E_BADARG, "qkv_row_stride %lld < row of %lld elements or not 16-byte",
                    static_cast<long long>(qkv_row_stride), static_cast<long long>(row));
    const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(qkv);
    const void* key = base + static_cast<int64_t>(num_q_heads) * head_dim;
    const void* value = base + static_cast<int64_t>(num_q_heads + num_kv_heads) * head_dim;
    return store_decode_impl(qkv, qkv_row_stride, key, qkv_row_stride, value, qkv_row_stride,
                             slot_mapping, k_cache, v_cache, block_tables, table_row_stride,
                             table_width, context_lens, out, workspace, workspace_bytes, batch,
                             num_q_heads, num_kv_heads, head_dim, block_size, num_blocks, scale,
                             num_splits, nullptr, 0, dtype, stream, &nr);
}

int b200attn_prepare_decode(const int32_t* packed, int32_t batch, int32_t table_width,
                            int32_t block_size, int64_t* input_ids, int64_t* positions,
                            int32_t* slot_mapping, int32_t* context_lens, int32_t* block_tables,
                            int64_t table_row_stride, int32_t capacity, void* stream) {
    if (batch < 0 || capacity < batch || table_width < 0 || block_size <= 0)
        return fail(B200ATTN_E_BADARG, "bad sizes (batch %d, capacity %d, table_width %d)", batch,
                    capacity, table_width);
    if (batch > 1024) return fail(B200ATTN_E_UNSUPPORTED, "batch %d > 1024", batch);
    if (capacity == 0) return 0;
    if (!slot_mapping || !context_lens) return fail(B200ATTN_E_BADARG, "null pointer");
    if (batch > 0 && (!packed || !input_ids || !positions || !block_tables))
        return fail(B200ATTN_E_BADARG, "null pointer");
    if (table_row_stride < table_width)
        return fail(B200ATTN_E_BADARG, "table_row_stride smaller than table_width");
    b200::PrepareDecodeParams p{};
    p.packed = packed;
    p.input_ids = input_ids;
    p.positions = positions;
    p.slot_mapping = slot_mapping;
    p.context_lens = context_lens;
    p.block_tables = block_tables;
    p.table_row_stride = table_row_stride;
    p.batch = batch;
    p.capacity = capacity;
    p.table_width = table_width;
    p.block_size = block_size;
    cudaError_t e = b200::launch_prepare_decode(p, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : cuda_fail(e, "prepare_decode launch");
}

int b200attn_set_decode_overlap(int on) {
    return g_decode_overlap.exchange(on ? 1 : 0, std::memory_order_relaxed);
}

void b200attn_debug_decode_trace(void* stamps) {
    g_decode_trace = static_cast<unsigned long long*>(stamps);
}

void b200attn_debug_prefill_trace(void* stamps) {
    g_prefill_trace = static_cast<unsigned long long*>(stamps);
}

int b200attn_set_split_counters(void* counters, int64_t count) {
    if (count < 0 || (counters == nullptr) != (count == 0))
        return fail(B200ATTN_E_BADARG, "counters and count must be given together");
    if (counters && (reinterpret_cast<uintptr_t>(counters) & 3u))
        return fail(B200ATTN_E_BADARG, "counters must be 4-byte aligned");
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    if (dev < 0 || dev >= 64) return fail(B200ATTN_E_UNSUPPORTED, "device %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_split_counters_mu);
    g_split_counters[dev].ptr = static_cast<int32_t*>(counters);
    g_split_counters[dev].count = count;
    return 0;
}

int b200attn_prepare_prefill(const int32_t* packed, int32_t num_seqs, int32_t block_size,
                             int64_t* positions, int32_t total_q, int32_t* slot_mapping,
                             int32_t total_slots, int32_t* cu_seqlens_q, int32_t* cu_seqlens_k,
                             void* stream) {
    if (num_seqs < 0 || block_size <= 0 || total_q < 0 || total_slots < 0)
        return fail(B200ATTN_E_BADARG, "negative or zero size");
    if (num_seqs > 1024) return fail(B200ATTN_E_UNSUPPORTED, "num_seqs %d > 1024", num_seqs);
    if (!cu_seqlens_q || !cu_seqlens_k) return fail(B200ATTN_E_BADARG, "null cu_seqlens");
    if (num_seqs > 0 && !packed) return fail(B200ATTN_E_BADARG, "null packed record");
    if ((total_q > 0 && !positions) || (total_slots > 0 && !slot_mapping))
        return fail(B200ATTN_E_BADARG, "null output");
    b200::PreparePrefillParams p{};
    p.packed = packed;
    p.positions = positions;
    p.slot_mapping = slot_mapping;
    p.cu_seqlens_q = cu_seqlens_q;
    p.cu_seqlens_k = cu_seqlens_k;
    p.num_seqs = num_seqs;
    p.block_size = block_size;
    cudaError_t e = b200::launch_prepare_prefill(p, total_q > total_slots ? total_q : total_slots,
                                                 static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : cuda_fail(e, "prepare_prefill launch");
}

}  // extern "C"
