// Varlen causal prefill on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), sm_100a.
// Replaces flash_attn_varlen_func on the prefill branch of Attention.forward
// (reference nanovllm/layers/attention.py:72-86), dense and paged-prefix K/V.
//
// One CTA = 256 query rows (two 128-row sub-tiles) of one (sequence, q-head); 12 warps (the last
// warpgroup gives most of its registers to the softmax groups with setmaxnreg):
//   warps 0-3   softmax group 0: one thread per query row of sub-tile 0 (TMEM lane = row)
//   warps 4-7   softmax group 1: sub-tile 1
//   warp  8     TMA producer: Q once, then 128-token K and V tiles through two 2-stage rings
//               (cp.async.bulk.tensor, SWIZZLE_128B boxes of 64 dims x 128 rows, mbarrier tx)
//   warp  9     MMA issuer (one elected lane) + TMEM owner
// TMEM (512 columns x 128 lanes of fp32): S0 | S1 | O0 | O1, 128 columns each.
//   S_i = Q_i . K_j^T      tcgen05.mma SS, M128 N128 K16 x 8, both operands K-major in smem
//   P_i = exp2(S_i*c - m)  softmax thread reads its S row with tcgen05.ld, writes bf16 P over the
//                          first 64 columns of S_i with tcgen05.st (fp32 max/sum in registers)
//   O_i += P_i . V_j       tcgen05.mma TS: A = P from TMEM, B = V tile MN-major in smem
// The issuer interleaves the two sub-tiles (PV_0[j], S_0[j+1], PV_1[j], S_1[j+1], ...) so the
// tensor pipe works on one sub-tile while the other one is in softmax. tcgen05.mma executes in
// issue order, so S_i[j+1] may overwrite P_i[j] as soon as it is issued after PV_i[j].
// O is only rescaled when a row maximum grows by more than 2^8 (lazy rescale); the row sum uses
// the same stale maximum, so O / l is exact at the end.
//
// Roofline: bf16 tensor pipe. Algorithmic flops = 4 * 128 * (visible q-k pairs) per q-head.
#include "common.h"
#include "ptx.cuh"
#include "tcgen05.cuh"

#include <cstdlib>

namespace b200 {

namespace {

constexpr int kTM = 128;                       // query rows per sub-tile (= TMEM lanes)
constexpr int kTN = 128;                       // KV tokens per tile
constexpr int kStages = 2;                     // K ring and V ring depth
constexpr int kHalfBytes = 128 * 128;          // [128 rows][64 dims] bf16, 128-byte rows, SW128
constexpr int kTileBytes = 2 * kHalfBytes;     // 32 KiB: dims 0-63 | dims 64-127
constexpr int kProducerWarp = 8;
constexpr int kMmaWarp = 9;
constexpr int kTcThreads = 384;                // 3 warpgroups: softmax 0, softmax 1, {TMA, MMA, 2 idle}
// setmaxnreg budgets. The pool is what the CTA got at launch (168 regs x 384 threads), not the
// whole register file: an inc that does not fit waits forever.
constexpr int kRegsLaunch = 168;
constexpr int kRegsSoftmax = 224;
constexpr int kRegsControl = 48;
static_assert(256 * kRegsSoftmax + 128 * kRegsControl <= kTcThreads * kRegsLaunch,
              "setmaxnreg budgets exceed the registers the CTA owns");
constexpr uint32_t kTmemCols = 512;
constexpr float kRescaleThreshold = 8.0f;      // log2 domain

constexpr int kSmemQ = 0;
constexpr int kSmemK = kSmemQ + 2 * kTileBytes;
constexpr int kSmemV = kSmemK + kStages * kTileBytes;
constexpr int kSmemBar = kSmemV + kStages * kTileBytes;
constexpr int kTcSmemBytes = kSmemBar + 256 + 1024;

struct Barriers {
    uint64_t q_full;
    uint64_t k_full[kStages], k_empty[kStages];
    uint64_t v_full[kStages], v_empty[kStages];
    uint64_t s_full[2], p_full[2], o_full[2];
    uint32_t tmem_base;
};

// Bounded spin: a protocol bug traps (launch error) instead of hanging the GPU.
__device__ __forceinline__ void wait_bar(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 28)) __trap();
    }
}

// kCascade = false: varlen causal prefill, CTA = (pair of 128-row tiles, KV head, sequence);
//   a tile row is (query token r / G, group head r % G).
// kCascade = true:  shared-prefix pass of the cascade decode (cascade.cu). CTA = (pair of 128-row
//   tiles of the B*G query rows of one KV head, prefix chunk, KV head). A tile row r is
//   (batch row r / G, group head r % G): the TMA box 64 dims x G heads x 128/G batch rows of
//   q [B, Hq, 128] lands in shared memory in exactly that order. Keys are the tiles of this
//   chunk of the P shared pages (block ids from row b0's table), all visible: no masking. The
//   epilogue writes normalised fp32 partials + log2-sum-exp for the decode merge.
// kEmu of every 4 exponential pairs are evaluated on the FMA pipe (tc::exp2_fma_pipe), the rest on
// the MUFU.
template <bool kCascade, int kEmu>
__global__ void __launch_bounds__(kTcThreads, 1)
prefill_tc_kernel(const __grid_constant__ CUtensorMap q_map,
                  const __grid_constant__ CUtensorMap k_map,
                  const __grid_constant__ CUtensorMap v_map, const PrefillParams p) {
    int seq = 0, h = 0, kvh = 0, q_begin = 0, Lq = 0, k_begin = 0, Lk = 0, offset = 0;
    int j0 = 0;       // first 128-token KV tile of this CTA (cascade: start of the chunk)
    int chunk = 0;    // cascade: prefix chunk = workspace slot ws_slot0 + chunk
    int q_c1 = 0;     // TMA head coordinate of the Q boxes
    int row0[2], n_kv[2], q_c2[2];
    if constexpr (kCascade) {
        // Let the dependent grid (the streaming decode over the per-row remainders, launched
        // with programmatic stream serialization) start now; it waits on this grid only where it
        // reads the partials written below.
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        const PrefixChunks pc = prefix_chunks_of(p.plan[0], p.block_size, p.prefix_chunks);
        chunk = blockIdx.y;
        if (chunk >= pc.valid) return;  // uniform for the whole CTA (also: no shared prefix)
        kvh = blockIdx.z;
        seq = p.plan[1];                // row whose table names the shared pages
        j0 = chunk * pc.per;
        const int n = min(pc.per, pc.tiles - j0);
        const int rows_per_sub = kTM / p.group;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            row0[i] = (static_cast<int>(blockIdx.x) * 2 + i) * rows_per_sub;  // first batch row
            n_kv[i] = row0[i] < p.batch ? n : 0;
            q_c2[i] = row0[i];
        }
        q_c1 = kvh * p.group;
    } else {
        // GQA packing: a 128-row sub-tile = 128/G query tokens x the G heads of one KV head
        // (row r = token r / G, head r % G; one TMA box 64 dims x G heads x 128/G tokens), so a
        // K/V tile is loaded once per KV head and short query blocks still fill the tile.
        seq = blockIdx.z;
        kvh = blockIdx.y;
        q_begin = p.cu_seqlens_q[seq];
        Lq = p.cu_seqlens_q[seq + 1] - q_begin;
        k_begin = p.cu_seqlens_k[seq];
        Lk = p.cu_seqlens_k[seq + 1] - k_begin;
        const int toks_per_sub = kTM / p.group;
        const int n_pairs = (Lq + 2 * toks_per_sub - 1) / (2 * toks_per_sub);
        if (static_cast<int>(blockIdx.x) >= n_pairs) return;  // uniform for the whole CTA
        const int pair = n_pairs - 1 - blockIdx.x;            // heaviest (last) tokens first
        offset = Lk - Lq;                                     // token i sees keys j <= i + offset
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            row0[i] = (pair * 2 + i) * toks_per_sub;          // first query token of the sub-tile
            const int toks_end = min(Lq, row0[i] + toks_per_sub);
            const int vis = row0[i] < Lq ? min(Lk, toks_end + offset) : 0;  // visible keys [0, vis)
            n_kv[i] = vis > 0 ? (vis + kTN - 1) / kTN : 0;
            q_c2[i] = q_begin + row0[i];
        }
        q_c1 = kvh * p.group;
    }
    const int g_shift = __ffs(p.group) - 1;  // G is a power of two (128 % G == 0)
    (void)h;
    const int n_max = max(n_kv[0], n_kv[1]);

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    Barriers* bar = reinterpret_cast<Barriers*>(smem + kSmemBar);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        mbar_init(&bar->q_full, 1);
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&bar->k_full[s], 1);
            mbar_init(&bar->k_empty[s], 1);
            mbar_init(&bar->v_full[s], 1);
            mbar_init(&bar->v_empty[s], 1);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            mbar_init(&bar->s_full[i], 1);
            mbar_init(&bar->p_full[i], kTM);
            mbar_init(&bar->o_full[i], 1);
        }
        fence_mbar_init();
    }
    if (warp == kMmaWarp) tc::tmem_alloc(&bar->tmem_base, kTmemCols);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = bar->tmem_base;

    // setmaxnreg sits inside each role branch so ptxas budgets every role separately.
    if (warp == kProducerWarp) {
        // ============================ TMA producer ============================
        tc::reg_dec<kRegsControl>();
        if (lane == 0 && n_max > 0) {
            tma_prefetch_desc(&q_map);
            tma_prefetch_desc(&k_map);
            tma_prefetch_desc(&v_map);
            const int n_q = n_kv[1] > 0 ? 2 : 1;  // sub-tile 0 is loaded whenever anything runs
            mbar_expect_tx(&bar->q_full, n_q * kTileBytes);
            for (int i = 0; i < n_q; ++i) {
                uint8_t* dst = smem + kSmemQ + i * kTileBytes;
                tma_load_3d(dst, &q_map, &bar->q_full, 0, q_c1, q_c2[i]);
                tma_load_3d(dst + kHalfBytes, &q_map, &bar->q_full, 64, q_c1, q_c2[i]);
            }
            const bool paged = kCascade || p.block_tables != nullptr;
            const int32_t* table =
                paged ? p.block_tables + static_cast<int64_t>(seq) * p.table_row_stride : nullptr;
            for (int j = 0; j < n_max; ++j) {
                const int s = j % kStages;
                const uint32_t ph = (j / kStages) & 1u;
                const int tok0 = (j0 + j) * kTN;
                int slot;
                if (paged) {
                    // a 128-token tile never straddles a page (block_size % 128 == 0)
                    slot = table[tok0 / p.block_size] * p.block_size + tok0 % p.block_size;
                } else {
                    slot = k_begin + tok0;
                }
                wait_bar(&bar->k_empty[s], ph ^ 1u);
                mbar_expect_tx(&bar->k_full[s], kTileBytes);
                uint8_t* kd = smem + kSmemK + s * kTileBytes;
                tma_load_3d(kd, &k_map, &bar->k_full[s], 0, kvh, slot);
                tma_load_3d(kd + kHalfBytes, &k_map, &bar->k_full[s], 64, kvh, slot);
                wait_bar(&bar->v_empty[s], ph ^ 1u);
                mbar_expect_tx(&bar->v_full[s], kTileBytes);
                uint8_t* vd = smem + kSmemV + s * kTileBytes;
                tma_load_3d(vd, &v_map, &bar->v_full[s], 0, kvh, slot);
                tma_load_3d(vd + kHalfBytes, &v_map, &bar->v_full[s], 64, kvh, slot);
            }
        }
    } else if (warp == kMmaWarp) {
        // ============================ MMA issuer ============================
        tc::reg_dec<kRegsControl>();
        constexpr uint32_t idesc_s = tc::idesc_bf16_f32(kTM, kTN, false);
        constexpr uint32_t idesc_o = tc::idesc_bf16_f32(kTM, kHeadDim, true);
        const uint32_t q_addr = smem_u32(smem + kSmemQ);
        const uint32_t k_addr = smem_u32(smem + kSmemK);
        const uint32_t v_addr = smem_u32(smem + kSmemV);

        auto issue_s = [&](int i, int stage) {  // S_i = Q_i . K^T   (head_dim 128 = 2 halves x 4)
            if (lane == 0) {
                const uint32_t a0 = q_addr + i * kTileBytes;
                const uint32_t b0 = k_addr + stage * kTileBytes;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint32_t off = (kk >> 2) * kHalfBytes + (kk & 3) * 32;
                    tc::mma_ss(tmem + i * kTN, tc::desc_kmajor(a0 + off), tc::desc_kmajor(b0 + off),
                               idesc_s, kk > 0);
                }
                tc::commit(&bar->s_full[i]);
            }
            __syncwarp();
        };
        auto issue_pv = [&](int i, int stage, bool accumulate) {  // O_i (+)= P_i . V
            if (lane == 0) {
                const uint32_t b0 = v_addr + stage * kTileBytes;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {  // 16 tokens per instruction
                    tc::mma_ts(tmem + 2 * kTN + i * kHeadDim, tmem + i * kTN + kk * 8,
                               tc::desc_mnmajor(b0 + kk * 16 * 128, kHalfBytes), idesc_o,
                               accumulate || kk > 0);
                }
            }
            __syncwarp();
        };

        if (n_max > 0) {
            wait_bar(&bar->q_full, 0);
            wait_bar(&bar->k_full[0], 0);
            tc::fence_after_sync();
#pragma unroll
            for (int i = 0; i < 2; ++i)
                if (n_kv[i] > 0) issue_s(i, 0);
            if (lane == 0) tc::commit(&bar->k_empty[0]);
            __syncwarp();
        }
        for (int j = 0; j < n_max; ++j) {
            const int s = j % kStages;
            wait_bar(&bar->v_full[s], (j / kStages) & 1u);
            // Rows at or beyond Lk inside the tile are masked (P = 0), but 0 * NaN would still
            // poison O: clear those V rows (unwritten cache slots may hold anything).
            const int valid = kCascade ? kTN : Lk - j * kTN;  // shared pages are full
            if (valid < kTN) {
                uint8_t* vs = smem + kSmemV + s * kTileBytes;
                const int r_lo = max(valid, 0);
                for (int idx = r_lo * 16 + lane; idx < kTN * 16; idx += 32) {
                    const int r = idx >> 4, c = idx & 15;
                    *reinterpret_cast<int4*>(vs + (c >> 3) * kHalfBytes + r * 128 + (c & 7) * 16) =
                        make_int4(0, 0, 0, 0);
                }
                fence_proxy_async();
                __syncwarp();
            }
            const int sn = (j + 1) % kStages;
            if (j + 1 < n_max) wait_bar(&bar->k_full[sn], ((j + 1) / kStages) & 1u);
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                if (j < n_kv[i]) {
                    wait_bar(&bar->p_full[i], j & 1u);
                    tc::fence_after_sync();
                    issue_pv(i, s, j > 0);
                    if (j == n_kv[i] - 1) {
                        if (lane == 0) tc::commit(&bar->o_full[i]);
                        __syncwarp();
                    }
                }
                if (j + 1 < n_kv[i]) issue_s(i, sn);
            }
            if (lane == 0) {
                tc::commit(&bar->v_empty[s]);
                if (j + 1 < n_max) tc::commit(&bar->k_empty[sn]);
            }
            __syncwarp();
        }
    } else if (warp < kProducerWarp) {
        // ============================ softmax / correction / epilogue ============================
        tc::reg_inc<kRegsSoftmax>();
        const int i = warp >> 2;                       // sub-tile
        const int r = (warp & 3) * 32 + lane;          // row inside the sub-tile = TMEM lane
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const uint32_t s_tmem = tmem + lane_base + i * kTN;
        const uint32_t o_tmem = tmem + lane_base + 2 * kTN + i * kHeadDim;
        const int row = row0[i] + (r >> g_shift);      // query token (cascade: batch row)
        const int head = kvh * p.group + (r & (p.group - 1));
        const int n = n_kv[i];
        const float c = p.scale_log2;
        float m_ref = -INFINITY;  // scaled (log2-domain) reference maximum of this row
        float l = 0.f;

        for (int j = 0; j < n; ++j) {
            wait_bar(&bar->s_full[i], j & 1u);
            tc::fence_after_sync();
            uint32_t sr[128];
            tc::tmem_ld128(s_tmem, sr);

            const int key0 = j * kTN;
            // columns <= lim are visible (cascade: every key of the shared prefix is)
            const int lim = kCascade ? kTN - 1 : min(row + offset, Lk - 1) - key0;
            if (lim < kTN - 1) {
#pragma unroll
                for (int t = 0; t < 128; ++t)
                    sr[t] = t <= lim ? sr[t] : 0xff800000u;  // -inf
            }
            float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
            for (int t = 0; t < 128; t += 8) {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    mx4[u] = tc::fmax3(mx4[u], __uint_as_float(sr[t + 2 * u]),
                                       __uint_as_float(sr[t + 2 * u + 1]));
            }
            const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
            const float m_new = fmaxf(m_ref, mx * c);
            if (j == 0) {
                m_ref = m_new;
            } else if (const bool grow = m_new - m_ref > kRescaleThreshold;
                       __any_sync(0xffffffffu, grow)) {
                // PV_i[j-1] has completed (it was issued before S_i[j], whose commit we waited on).
                // The branch is warp-uniform (TMEM access is .sync.aligned) but the decision is
                // per row: rows that did not grow are multiplied by exactly 1, so a row's result
                // never depends on its neighbours.
                const float alpha = grow ? fast_exp2(m_ref - m_new) : 1.f;
                l *= alpha;
                m_ref = grow ? m_new : m_ref;
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) {
                    uint32_t orow[32];
                    tc::tmem_ld32(o_tmem + q4 * 32, orow);
#pragma unroll
                    for (int t = 0; t < 32; ++t)
                        orow[t] = __float_as_uint(__uint_as_float(orow[t]) * alpha);
                    tc::tmem_st32(o_tmem + q4 * 32, orow);
                }
            }
            const float mu = m_ref == -INFINITY ? 0.f : m_ref;
            // p = exp2(s*c - mu): packed f32x2 FMA / ADD halve the issue slots next to MUFU.ex2
            const float2 c2 = make_float2(c, c), nmu2 = make_float2(-mu, -mu);
            float2 ls[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {  // 32 packed columns at a time: P regs die early
                uint32_t pr[32];
#pragma unroll
                for (int t = 0; t < 32; ++t) {
                    const int e = hh * 64 + 2 * t;
                    float2 x = make_float2(__uint_as_float(sr[e]), __uint_as_float(sr[e + 1]));
                    x = __ffma2_rn(x, c2, nmu2);
                    if ((t & 3) < kEmu) {
                        x = tc::exp2_fma_pipe(x);
                    } else {
                        x.x = fast_exp2(x.x);
                        x.y = fast_exp2(x.y);
                    }
                    ls[t & 1] = __fadd2_rn(ls[t & 1], x);
                    pr[t] = pack_bf16x2(x.x, x.y);
                }
                tc::tmem_st32(s_tmem + hh * 32, pr);
            }
            l += (ls[0].x + ls[0].y) + (ls[1].x + ls[1].y);
            tc::wait_st();
            tc::fence_before_sync();
            mbar_arrive(&bar->p_full[i]);
        }

        if constexpr (kCascade) {
            // ---- epilogue: normalised fp32 partial + log2-sum-exp of this prefix chunk ----
            const int b = row;
            if (n > 0) {  // warp-uniform: n depends on the sub-tile only
                wait_bar(&bar->o_full[i], 0);
                tc::fence_after_sync();
                const int64_t qrow =
                    static_cast<int64_t>(b) * p.num_q_heads + head;
                const int64_t slot = qrow * p.ws_slots + p.ws_slot0 + chunk;
                const float inv = l > 0.f ? 1.f / l : 0.f;
                float* og = p.ws_o + slot * kHeadDim;
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) {
                    uint32_t orow[32];
                    tc::tmem_ld32(o_tmem + q4 * 32, orow);
                    if (b < p.batch) {
#pragma unroll
                        for (int v4 = 0; v4 < 8; ++v4) {
                            float4 f;
                            f.x = __uint_as_float(orow[v4 * 4 + 0]) * inv;
                            f.y = __uint_as_float(orow[v4 * 4 + 1]) * inv;
                            f.z = __uint_as_float(orow[v4 * 4 + 2]) * inv;
                            f.w = __uint_as_float(orow[v4 * 4 + 3]) * inv;
                            *reinterpret_cast<float4*>(og + q4 * 32 + v4 * 4) = f;
                        }
                    }
                }
                if (b < p.batch) p.ws_lse[slot] = l > 0.f ? m_ref + log2f(l) : -INFINITY;
            }
        } else {
        // ---- epilogue: O / l -> bf16 -> global (row-predicated 16-byte stores) ----
        const bool row_ok = row < Lq;
        uint8_t* orow_g = static_cast<uint8_t*>(p.out) +
                          ((static_cast<int64_t>(q_begin) + row) * p.num_q_heads + head) * 256;
        if (n > 0) {
            wait_bar(&bar->o_full[i], 0);
            tc::fence_after_sync();
            const float inv = l > 0.f ? 1.f / l : 0.f;
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
                uint32_t orow[32];
                tc::tmem_ld32(o_tmem + q4 * 32, orow);
                if (row_ok) {
#pragma unroll
                    for (int v4 = 0; v4 < 4; ++v4) {
                        int4 pk;
                        pk.x = pack_bf16x2(__uint_as_float(orow[v4 * 8 + 0]) * inv,
                                           __uint_as_float(orow[v4 * 8 + 1]) * inv);
                        pk.y = pack_bf16x2(__uint_as_float(orow[v4 * 8 + 2]) * inv,
                                           __uint_as_float(orow[v4 * 8 + 3]) * inv);
                        pk.z = pack_bf16x2(__uint_as_float(orow[v4 * 8 + 4]) * inv,
                                           __uint_as_float(orow[v4 * 8 + 5]) * inv);
                        pk.w = pack_bf16x2(__uint_as_float(orow[v4 * 8 + 6]) * inv,
                                           __uint_as_float(orow[v4 * 8 + 7]) * inv);
                        st_v4(orow_g + q4 * 64 + v4 * 16, pk);
                    }
                }
            }
        } else if (row_ok) {
            // no visible key for this sub-tile (only possible when Lk < Lq): zeros, as flash-attn
#pragma unroll
            for (int v = 0; v < 16; ++v) st_v4(orow_g + v * 16, make_int4(0, 0, 0, 0));
        }
        }  // !kCascade
    }

    else {
        tc::reg_dec<kRegsControl>();  // idle warps of the control warpgroup
    }

    tc::fence_before_sync();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem, kTmemCols);
    }
}

}  // namespace

namespace {

// B200ATTN_EXP_EMU=0..3 overrides how many of every 4 exponential pairs leave the MUFU.
int exp_emu_from_env(int dflt) {
    static int v = [] {
        const char* e = std::getenv("B200ATTN_EXP_EMU");
        return (e && e[0] >= '0' && e[0] <= '3') ? e[0] - '0' : -1;
    }();
    return v < 0 ? dflt : v;
}

template <bool kCascade, int kEmu>
cudaError_t launch_tc(dim3 grid, const PrefillParams& p, const CUtensorMap& q_map,
                      const CUtensorMap& k_map, const CUtensorMap& v_map, cudaStream_t stream) {
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(prefill_tc_kernel<kCascade, kEmu>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kTcSmemBytes);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    prefill_tc_kernel<kCascade, kEmu><<<grid, kTcThreads, kTcSmemBytes, stream>>>(q_map, k_map,
                                                                                  v_map, p);
    count_launch();
    return cudaGetLastError();
}

template <bool kCascade>
cudaError_t launch_tc_emu(int emu, dim3 grid, const PrefillParams& p, const CUtensorMap& q_map,
                          const CUtensorMap& k_map, const CUtensorMap& v_map,
                          cudaStream_t stream) {
    switch (emu) {
        case 0: return launch_tc<kCascade, 0>(grid, p, q_map, k_map, v_map, stream);
        case 1: return launch_tc<kCascade, 1>(grid, p, q_map, k_map, v_map, stream);
        case 3: return launch_tc<kCascade, 3>(grid, p, q_map, k_map, v_map, stream);
        default: return launch_tc<kCascade, 2>(grid, p, q_map, k_map, v_map, stream);
    }
}

}  // namespace

constexpr int kDefaultExpEmu = 0;

cudaError_t launch_prefill_tc(const PrefillParams& p, const CUtensorMap& q_map,
                              const CUtensorMap& k_map, const CUtensorMap& v_map,
                              cudaStream_t stream) {
    const int rows = p.max_seqlen_q * p.group;  // query rows per KV head of the longest sequence
    dim3 grid((rows + 2 * kTM - 1) / (2 * kTM), p.num_kv_heads, p.num_seqs);
    return launch_tc_emu<false>(exp_emu_from_env(kDefaultExpEmu), grid, p, q_map, k_map, v_map,
                                stream);
}

cudaError_t launch_cascade_prefix(const PrefillParams& p, const CUtensorMap& q_map,
                                  const CUtensorMap& k_map, const CUtensorMap& v_map,
                                  cudaStream_t stream) {
    const int rows = p.batch * p.group;  // query rows per KV head
    dim3 grid((rows + 2 * kTM - 1) / (2 * kTM), p.prefix_chunks, p.num_kv_heads);
    return launch_tc_emu<true>(exp_emu_from_env(kDefaultExpEmu), grid, p, q_map, k_map, v_map,
                               stream);
}

}  // namespace b200
