// Inline-PTX wrappers for the 5th-generation tensor cores of sm_100a: tcgen05.mma with operands in
// shared memory (SS) or tensor memory (TS), TMEM allocation, tcgen05.ld/st, commit and fences,
// plus the 64-bit shared-memory matrix descriptors and the 32-bit instruction descriptor the MMA
// consumes. Hand-written from the PTX ISA; used by the prefill kernel (prefill_tc.cu).
#pragma once
#include <cstdint>

#include "ptx.cuh"

namespace b200 {
namespace tc {

// ----------------------------------------------------------------------------------------------
// Descriptors
// ----------------------------------------------------------------------------------------------
// Shared-memory matrix descriptor (tcgen05 flavour: version field = 1), SWIZZLE_128B layouts.
//   bits  0..13  start address >> 4          bits 16..29  leading-dimension byte offset >> 4
//   bits 32..45  stride byte offset >> 4     bits 46..47  descriptor version (1 on sm_100)
//   bits 49..51  base offset (0: tiles are 1024-byte aligned)      bits 61..63  layout (2 = SW128)
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes,
                                                    uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= 1ull << 46;
    d |= 2ull << 61;
    return d;
}
// K-major operand tile [rows][64 bf16] (128-byte rows, 8-row groups 1024 B apart). The leading
// offset is not used by the hardware for swizzled K-major layouts (canonical value 16 B).
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t smem_addr) {
    return smem_desc_sw128(smem_addr, 16, 1024);
}
// MN-major operand: [k rows][64 bf16 of MN] blocks; the next 64 MN elements are `mn_block_bytes`
// further, the next 8 k rows 1024 B further.
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t smem_addr, uint32_t mn_block_bytes) {
    return smem_desc_sw128(smem_addr, mn_block_bytes, 1024);
}

// Instruction descriptor for kind::f16 with bf16 (or fp16) A/B and fp32 accumulate.
//   [4,6) D format (1 = f32)  [7,10) A format (0 = f16, 1 = bf16)  [10,13) B format (same codes)
//   bit 15 A major (0 = K)    bit 16 B major (0 = K, 1 = MN)   [17,23) N >> 3   [24,29) M >> 4
__host__ __device__ constexpr uint32_t idesc_f32acc(int m, int n, bool b_mn_major, bool f16) {
    return (1u << 4) | ((f16 ? 0u : 1u) << 7) | ((f16 ? 0u : 1u) << 10) |
           ((b_mn_major ? 1u : 0u) << 16) | (static_cast<uint32_t>(n >> 3) << 17) |
           (static_cast<uint32_t>(m >> 4) << 24);
}
__host__ __device__ constexpr uint32_t idesc_bf16_f32(int m, int n, bool b_mn_major) {
    return idesc_f32acc(m, n, b_mn_major, false);
}

// ----------------------------------------------------------------------------------------------
// TMEM allocation (one warp, .sync.aligned)
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(smem_result)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                 : "memory");
}

// ----------------------------------------------------------------------------------------------
// Ordering
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void fence_before_sync() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void fence_after_sync() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void wait_st() {
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
// mbarrier arrives (count 1) once every tcgen05 async op this thread issued so far has completed.
__device__ __forceinline__ void commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                     smem_u32(bar))
                 : "memory");
}

// ----------------------------------------------------------------------------------------------
// MMA: D[tmem] (+)= A . B, M = 128, K = 16 per instruction, issued by ONE thread
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                       uint32_t idesc, bool accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate ? 1u : 0u)
        : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc,
                                       uint32_t idesc, bool accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate ? 1u : 0u)
        : "memory");
}
// Fully static forms: TMEM addresses and the accumulate flag are compile-time constants (the CTA
// owns the whole TMEM, base 0), the descriptors are constant expressions of the caller. ptxas
// then feeds UTCHMMA from uniform moves instead of an ELECT / R2UR broadcast loop per operand.
// The WHOLE WARP must call these, converged: one lane is elected inside (elect.sync), so the
// instruction is not under a divergent branch.
template <uint32_t kDTmem, bool kAccumulate>
__device__ __forceinline__ void mma_ss_static(uint64_t a_desc, uint64_t b_desc, uint32_t idesc) {
    asm volatile(
        "{\n\t.reg .pred p, e;\n\t.reg .b32 d;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "mov.b32 d, %0;\n\t"
        "elect.sync _|e, 0xffffffff;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::f16 [d], %1, %2, %3, p;\n\t}"
        ::"n"(kDTmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "n"(kAccumulate ? 1 : 0)
        : "memory");
}
template <uint32_t kDTmem, uint32_t kATmem, bool kAccumulate>
__device__ __forceinline__ void mma_ts_static(uint64_t b_desc, uint32_t idesc) {
    asm volatile(
        "{\n\t.reg .pred p, e;\n\t.reg .b32 d, a;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "mov.b32 d, %0;\n\tmov.b32 a, %1;\n\t"
        "elect.sync _|e, 0xffffffff;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::f16 [d], [a], %2, %3, p;\n\t}"
        ::"n"(kDTmem), "n"(kATmem), "l"(b_desc), "r"(idesc), "n"(kAccumulate ? 1 : 0)
        : "memory");
}

// ----------------------------------------------------------------------------------------------
// TMEM <-> registers, shape 32x32b: lane t of the warp touches TMEM lane (base lane + t), 32
// consecutive 32-bit columns. A warp may only touch lanes 32*(warp_id%4) .. +31.
// The load waits for its own completion inside the asm block so the outputs are valid when the
// statement retires (the compiler cannot hoist their uses above a separate wait).
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
          "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
          "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
          "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31};"
        ::"r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]),
          "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]),
          "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]),
          "r"(r[29]), "r"(r[30]), "r"(r[31]), "r"(taddr)
        : "memory");
}


// 128 consecutive columns of this thread's TMEM lane: four x32 loads in flight, one wait.
__device__ __forceinline__ void tmem_ld128(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "{\n\t.reg .b32 a1, a2, a3;\n\t"
        "add.u32 a1, %128, 32;\n\tadd.u32 a2, %128, 64;\n\tadd.u32 a3, %128, 96;\n\t"
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%128];\n\t"
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [a1];\n\t"
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%64, %65, %66, %67, %68, %69, %70, %71, %72, %73, %74, %75, %76, %77, %78, %79, %80, %81, %82, %83, %84, %85, %86, %87, %88, %89, %90, %91, %92, %93, %94, %95}, [a2];\n\t"
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%96, %97, %98, %99, %100, %101, %102, %103, %104, %105, %106, %107, %108, %109, %110, %111, %112, %113, %114, %115, %116, %117, %118, %119, %120, %121, %122, %123, %124, %125, %126, %127}, [a3];\n\t"
        "tcgen05.wait::ld.sync.aligned;\n\t}"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63]), "=r"(r[64]), "=r"(r[65]), "=r"(r[66]), "=r"(r[67]), "=r"(r[68]), "=r"(r[69]), "=r"(r[70]), "=r"(r[71]), "=r"(r[72]), "=r"(r[73]), "=r"(r[74]), "=r"(r[75]), "=r"(r[76]), "=r"(r[77]), "=r"(r[78]), "=r"(r[79]), "=r"(r[80]), "=r"(r[81]), "=r"(r[82]), "=r"(r[83]), "=r"(r[84]), "=r"(r[85]), "=r"(r[86]), "=r"(r[87]), "=r"(r[88]), "=r"(r[89]), "=r"(r[90]), "=r"(r[91]), "=r"(r[92]), "=r"(r[93]), "=r"(r[94]), "=r"(r[95]), "=r"(r[96]), "=r"(r[97]), "=r"(r[98]), "=r"(r[99]), "=r"(r[100]), "=r"(r[101]), "=r"(r[102]), "=r"(r[103]), "=r"(r[104]), "=r"(r[105]), "=r"(r[106]), "=r"(r[107]), "=r"(r[108]), "=r"(r[109]), "=r"(r[110]), "=r"(r[111]), "=r"(r[112]), "=r"(r[113]), "=r"(r[114]), "=r"(r[115]), "=r"(r[116]), "=r"(r[117]), "=r"(r[118]), "=r"(r[119]), "=r"(r[120]), "=r"(r[121]), "=r"(r[122]), "=r"(r[123]), "=r"(r[124]), "=r"(r[125]), "=r"(r[126]), "=r"(r[127])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_st64(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x64.b32 [%64], "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63};"
        ::"r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]), "r"(r[32]), "r"(r[33]), "r"(r[34]), "r"(r[35]), "r"(r[36]), "r"(r[37]), "r"(r[38]), "r"(r[39]), "r"(r[40]), "r"(r[41]), "r"(r[42]), "r"(r[43]), "r"(r[44]), "r"(r[45]), "r"(r[46]), "r"(r[47]), "r"(r[48]), "r"(r[49]), "r"(r[50]), "r"(r[51]), "r"(r[52]), "r"(r[53]), "r"(r[54]), "r"(r[55]), "r"(r[56]), "r"(r[57]), "r"(r[58]), "r"(r[59]), "r"(r[60]), "r"(r[61]), "r"(r[62]), "r"(r[63]), "r"(taddr)
        : "memory");
}

// Register re-budgeting between warp groups (whole warpgroup executes it, value multiple of 8).
template <int N>
__device__ __forceinline__ void reg_inc() {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N));
}
template <int N>
__device__ __forceinline__ void reg_dec() {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N));
}
// 2^x for a pair of fp32 values WITHOUT the MUFU: Cody-Waite split x = floor(x) + f, a cubic for
// 2^f on [0, 1) (max relative error ~1e-4, far below the bf16 rounding P gets anyway), and the
// integer part added straight into the exponent field. The round-down add against 1.5 * 2^23
// leaves floor(x) in the low mantissa bits of `xr`, so (bits(xr) << 23) is floor(x) << 23.
// The softmax of the tcgen05 prefill is MUFU-bound (16 ex2/clk/SM against 8192 MMA flop/clk/SM);
// moving a fraction of the exponentials to the FMA pipe takes the MUFU off the critical path.
// x must be <= 127; values below -127 are clamped (result ~2^-127, i.e. 0 after bf16 rounding).
__device__ __forceinline__ float2 exp2_fma_pipe(float2 x) {
    const float kMagic = 12582912.f;  // 2^23 + 2^22
    x.x = fmaxf(x.x, -127.f);
    x.y = fmaxf(x.y, -127.f);
    float2 xr;
    asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
        "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %4};\n\t"
        "add.rm.ftz.f32x2 rd, ra, rb;\n\t"
        "mov.b64 {%0, %1}, rd;\n\t}"
        : "=f"(xr.x), "=f"(xr.y)
        : "f"(x.x), "f"(x.y), "f"(kMagic));
    const float2 fl = __fadd2_rn(xr, make_float2(-kMagic, -kMagic));     // floor(x), exact
    const float2 f = __ffma2_rn(fl, make_float2(-1.f, -1.f), x);         // x - floor(x) in [0, 1)
    float2 p = __ffma2_rn(make_float2(0.077119089663028717f, 0.077119089663028717f), f,
                          make_float2(0.227564394474029541f, 0.227564394474029541f));
    p = __ffma2_rn(p, f, make_float2(0.695146143436431885f, 0.695146143436431885f));
    p = __ffma2_rn(p, f, make_float2(1.f, 1.f));
    float2 r;
    r.x = __int_as_float(__float_as_int(p.x) + (__float_as_int(xr.x) << 23));
    r.y = __int_as_float(__float_as_int(p.y) + (__float_as_int(xr.y) << 23));
    return r;
}

__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

}  // namespace tc
}  // namespace b200
