// tcgen05 (5th-generation tensor core) building blocks for sm_100a: TMEM allocation, UMMA shared-memory
// and instruction descriptors, MMA issue/commit, TMEM loads, and the bf16x3 operand split that gives
// fp32-accurate products on the bf16 tensor pipe.  Inline PTX; no CUTLASS.
//
// Operand storage ("X layout"): a matrix X[row][feature] (R rows, F features, bf16) is stored as
//     byte offset(row, feature) = (feature / 8) * C + row * 16 + (feature % 8) * 2,   C = R * 16 + 16
// i.e. for every 8-feature chunk, all rows' 16-byte pieces are contiguous; the 16 bytes of padding per
// chunk rotate the shared-memory banks from chunk to chunk, so a warp that owns 32 features of one row
// (the epilogues) stores and loads without bank conflicts.  This is the canonical un-swizzled UMMA
// layout in BOTH orientations, so one copy serves as
//   * a K-major operand   (M/N = row,     K = feature): LBO = C (next K chunk), SBO = 128 (next 8 rows)
//   * an MN-major operand (M/N = feature, K = row):     SBO = C (next MN chunk), LBO = 128 (next 8 K)
// (core matrix = 8 x 16 B = 128 contiguous bytes in either view).
//
// fp32 accuracy from bf16 tensor cores: a = a1 + a2 + a3 with each part a bf16 (8+8+8 mantissa bits),
// a*b ~= a1b1 + (a1b2 + a2b1) + (a1b3 + a2b2 + a3b1): six MMAs accumulated in fp32 in TMEM; the
// dropped terms are below 2^-24 |a||b|.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace rrl {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- X layout ---------------------------------------------------------------------------------
__host__ __device__ constexpr int x_chunk_bytes(int R) { return R * 16 + 16; }
__host__ __device__ constexpr int x_plane_bytes(int R, int F) { return (F / 8) * x_chunk_bytes(R); }
__device__ __forceinline__ int x_off_bytes(int row, int feature, int R) { return (feature >> 3) * x_chunk_bytes(R) + row * 16 + (feature & 7) * 2; }

// ---- bf16x3 split -------------------------------------------------------------------------------
struct Split3 {
    __nv_bfloat16 p1, p2, p3;
};
__device__ __forceinline__ Split3 split3(float a)
{
    Split3 s;
    s.p1 = __float2bfloat16_rn(a);
    const float r1 = a - __bfloat162float(s.p1);  // exact
    s.p2 = __float2bfloat16_rn(r1);
    const float r2 = r1 - __bfloat162float(s.p2);  // exact
    s.p3 = __float2bfloat16_rn(r2);
    return s;
}

// Truncating variant kept in fp32 containers: t1 + t2 + t3 == a up to 2^-24 |a|, each t_k has only its
// top 16 bits set (so it IS a bf16 once the upper half-word is taken).  Two LOPs + two FADDs per
// element; pairs are packed into bf16x2 words with one byte-permute per plane.
struct Trunc3 {
    float t1, t2, t3;
};
__device__ __forceinline__ Trunc3 trunc3(float a)
{
    Trunc3 s;
    s.t1 = __uint_as_float(__float_as_uint(a) & 0xffff0000u);
    const float r1 = a - s.t1;  // exact
    s.t2 = __uint_as_float(__float_as_uint(r1) & 0xffff0000u);
    s.t3 = r1 - s.t2;           // exact; its low half-word is dropped when packed
    return s;
}
// bf16x2 word {lo = top half of x, hi = top half of y}
__device__ __forceinline__ uint32_t pack_hi16(float x, float y) { return __byte_perm(__float_as_uint(x), __float_as_uint(y), 0x7632); }
__device__ __forceinline__ unsigned short hi16(float x) { return (unsigned short)(__float_as_uint(x) >> 16); }

// 32 lanes x 16 / 8 consecutive columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]);
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8])
{
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

// ---- descriptors -----------------------------------------------------------------------------------
// shared-memory matrix descriptor, SWIZZLE_NONE, Blackwell version field = 1
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
// instruction descriptor for kind::f16 with bf16 inputs and fp32 accumulation
__host__ __device__ constexpr uint32_t idesc_bf16(int M, int N, bool a_mn_major, bool b_mn_major)
{
    return (1u << 4)                       // c_format = F32
           | (1u << 7)                     // a_format = BF16
           | (1u << 10)                    // b_format = BF16
           | ((a_mn_major ? 1u : 0u) << 15)
           | ((b_mn_major ? 1u : 0u) << 16)
           | ((uint32_t)(N >> 3) << 17)
           | ((uint32_t)(M >> 4) << 24);
}

// ---- TMEM ---------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_dst, uint32_t ncols)  // one full warp
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols)  // the same warp
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy smem writes -> visible to the tensor core (async proxy)
__device__ __forceinline__ void fence_smem_to_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread
__device__ __forceinline__ void mma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"((uint32_t)accumulate)
        : "memory");
}
// Warp-uniform issue: every lane of a converged warp calls this with identical operands and one elected
// lane issues.  Keeping the call site convergent lets the compiler hold the descriptors in uniform
// registers; under `if (tid == 0)` it wraps every MMA in an ELECT/R2UR waterfall loop (~57 cycles each).
__device__ __forceinline__ void mma_bf16_elect(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"((uint32_t)accumulate)
        : "memory");
}
// A operand taken from tensor memory (lane = row of A, 32-bit column = two consecutive K elements, the lower K
// index in the lower half-word); B from shared memory as usual.
__device__ __forceinline__ void mma_bf16_ts_elect(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, bool accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"((uint32_t)accumulate)
        : "memory");
}
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile(
        "{\n\t"
        ".reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, q;\n\t"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
// all previously issued MMAs of this thread -> arrive on the mbarrier when they have completed
__device__ __forceinline__ void mma_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- mbarrier -----------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}\n" ::"r"(smem_u32(bar)) : "memory");
}

// one arrival that also announces `bytes` of asynchronous (bulk-copy) traffic still to land on this barrier
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// TMA bulk copy (no tensor map): `bytes` contiguous bytes global -> this CTA's shared memory, completion counted on `bar`.
// Source, destination and size are multiples of 16.
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// ---- TMEM -> registers: 32 lanes x 32 consecutive 32-bit columns (thread = lane) -------------------------
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32])
{
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16])
{
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// registers -> TMEM: 32 lanes x 8 consecutive 32-bit columns (thread = lane)
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};\n" ::"r"(taddr), "r"(v[0]), "r"(v[1]),
                 "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// 32 lanes x 64 consecutive columns in ONE instruction and one wait (four x16 loads cost four round trips)
__device__ __forceinline__ void tmem_ld64(uint32_t taddr, float (&v)[64])
{
    uint32_t r[64];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
        "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
        "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]),
          "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]),
          "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]),
          "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]),
          "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 64; ++i) v[i] = __uint_as_float(r[i]);
}

// two x16 loads in flight, one wait
__device__ __forceinline__ void tmem_ld16x2(uint32_t taddr_a, float (&a)[16], uint32_t taddr_b, float (&b)[16])
{
    uint32_t r[16], q[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr_a)
        : "memory");
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3]), "=r"(q[4]), "=r"(q[5]), "=r"(q[6]), "=r"(q[7]), "=r"(q[8]),
          "=r"(q[9]), "=r"(q[10]), "=r"(q[11]), "=r"(q[12]), "=r"(q[13]), "=r"(q[14]), "=r"(q[15])
        : "r"(taddr_b)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = __uint_as_float(r[i]), b[i] = __uint_as_float(q[i]);
}

// registers -> TMEM without the wait: follow a batch of these with tmem_st_wait()
__device__ __forceinline__ void tmem_st8_nowait(uint32_t taddr, const uint32_t (&v)[8])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};\n" ::"r"(taddr), "r"(v[0]), "r"(v[1]),
                 "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// registers -> TMEM: 32 lanes x 16 consecutive 32-bit columns (thread = lane)
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float (&v)[16])
{
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};\n" ::"r"(taddr),
        "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
        "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
        "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
        "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15]))
        : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// The six bf16x3 products of one GEMM, smallest terms first.  Plane p (0,1,2) of an operand starts at
// base + p * plane_bytes; per K step (16 elements of K) the start address advances by a_kstep / b_kstep.
struct OperandDesc {
    uint32_t base, plane_bytes, lbo, sbo, kstep;
};
__device__ __forceinline__ void issue_gemm_bf16x3(uint32_t d_tmem, const OperandDesc &A, const OperandDesc &B, uint32_t idesc,
                                                  int ksteps, bool accumulate_into_d)
{
    const int pa[6] = {0, 1, 2, 0, 1, 0}, pb[6] = {2, 1, 0, 1, 0, 0};
    bool acc = accumulate_into_d;
#pragma unroll
    for (int t = 0; t < 6; ++t) {
        const uint32_t a0 = A.base + pa[t] * A.plane_bytes, b0 = B.base + pb[t] * B.plane_bytes;
        for (int s = 0; s < ksteps; ++s) {
            mma_bf16(d_tmem, smem_desc(a0 + s * A.kstep, A.lbo, A.sbo), smem_desc(b0 + s * B.kstep, B.lbo, B.sbo), idesc, acc);
            acc = true;
        }
    }
}

// The same GEMM issued by a whole (converged) warp: descriptors are built from warp-uniform values and one
// elected lane issues each MMA.  Measured on B200 (tests' rate probe, M=128, K=16 per MMA): 52 cycles per
// MMA at N=64 and 64 at N=128 (= the tensor-core floor) against 57+ for the single-thread form above.
template <int KSTEPS>
__device__ __forceinline__ void issue_gemm_bf16x3_warp(uint32_t d_tmem, const OperandDesc &A, const OperandDesc &B, uint32_t idesc,
                                                       bool accumulate_into_d)
{
    const uint64_t da = smem_desc(A.base, A.lbo, A.sbo), db = smem_desc(B.base, B.lbo, B.sbo);
    const int pa[6] = {0, 1, 2, 0, 1, 0}, pb[6] = {2, 1, 0, 1, 0, 0};
#pragma unroll
    for (int t = 0; t < 6; ++t) {
#pragma unroll
        for (int s = 0; s < KSTEPS; ++s) {
            const uint32_t oa = (pa[t] * A.plane_bytes + s * A.kstep) >> 4, ob = (pb[t] * B.plane_bytes + s * B.kstep) >> 4;
            mma_bf16_elect(d_tmem, da + oa, db + ob, idesc, accumulate_into_d || t > 0 || s > 0);
        }
    }
}
// A from TMEM: the three planes of A start at columns a0, a1, a2; a K step (16 elements) is 8 columns
template <int KSTEPS>
__device__ __forceinline__ void issue_gemm_bf16x3_ts_warp(uint32_t d_tmem, uint32_t a0, uint32_t a1, uint32_t a2, const OperandDesc &B,
                                                          uint32_t idesc, bool accumulate_into_d)
{
    const uint64_t db = smem_desc(B.base, B.lbo, B.sbo);
    const int pa[6] = {0, 1, 2, 0, 1, 0}, pb[6] = {2, 1, 0, 1, 0, 0};
#pragma unroll
    for (int t = 0; t < 6; ++t) {
        const uint32_t a = pa[t] == 0 ? a0 : pa[t] == 1 ? a1 : a2;
#pragma unroll
        for (int s = 0; s < KSTEPS; ++s) {
            const uint32_t ob = (pb[t] * B.plane_bytes + s * B.kstep) >> 4;
            mma_bf16_ts_elect(d_tmem, a + 8 * s, db + ob, idesc, accumulate_into_d || t > 0 || s > 0);
        }
    }
}
// commit from the same elected lane that issued (elect.sync picks the same leader for the same mask)
__device__ __forceinline__ void mma_commit_warp(uint64_t *bar)
{
    asm volatile(
        "{\n\t"
        ".reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t"
        "}\n" ::"r"(smem_u32(bar))
        : "memory");
}

}  // namespace tc
}  // namespace rrl
