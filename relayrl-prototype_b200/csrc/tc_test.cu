// Self-test of the tcgen05 building blocks (tc_common.cuh): one CTA computes one bf16x3 GEMM in each
// of the three operand orientations the fused learner kernel uses.  Exported for tests only
// (tests/test_gpu_tc.py); not part of include/relayrl_b200.h.
#include "common.cuh"
#include "tc_common.cuh"

namespace rrl {
namespace {

using namespace tc;

// mode 0: D[128 x 64]  = A[128 x 128] * B[64 x 128]^T      (A K-major, B K-major)     -- forward GEMM
// mode 1: D[128 x 64]  = A[128 x 128]^T * B[64 x 128]^T    (A MN-major, B K-major)    -- data-gradient GEMM
// mode 2: D[128 x 128] = A[64 x 128]^T * B[64 x 128]       (A MN-major, B MN-major)   -- weight-gradient GEMM
// mode 3: D[128 x 128] = A[128 x 0:64] * B[64 x 128]       (A K-major,  B MN-major)   -- weight-gradient GEMM with
//                                                                                         the transposed dZ2 planes
// mode 4: D[128 x 128] = A[64 x 128]^T * B[128 x 0:64]^T   (A^T in TENSOR MEMORY, B K-major)  -- weight-gradient GEMM
//                                                                                            with H1^T kept in TMEM
// All inputs row-major fp32 [rows][128] in global memory.
__global__ void __launch_bounds__(128, 1) tc_gemm_test_kernel(int mode, const float *__restrict__ A,
                                                              const float *__restrict__ B, float *__restrict__ D)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int RA = (mode == 2 || mode == 4) ? 64 : 128, RB = mode == 4 ? 128 : 64;
    const uint32_t planeA = x_plane_bytes(RA, 128), planeB = x_plane_bytes(RB, 128);
    unsigned char *sa = smem, *sb = smem + 3 * planeA;

    for (int i = tid; i < RA * 128; i += 128) {
        const int r = i >> 7, f = i & 127;
        const Split3 s = split3(A[i]);
        const int o = x_off_bytes(r, f, RA);
        *reinterpret_cast<__nv_bfloat16 *>(sa + o) = s.p1;
        *reinterpret_cast<__nv_bfloat16 *>(sa + planeA + o) = s.p2;
        *reinterpret_cast<__nv_bfloat16 *>(sa + 2 * planeA + o) = s.p3;
    }
    for (int i = tid; i < RB * 128; i += 128) {
        const int r = i >> 7, f = i & 127;
        const Split3 s = split3(B[i]);
        const int o = x_off_bytes(r, f, RB);
        *reinterpret_cast<__nv_bfloat16 *>(sb + o) = s.p1;
        *reinterpret_cast<__nv_bfloat16 *>(sb + planeB + o) = s.p2;
        *reinterpret_cast<__nv_bfloat16 *>(sb + 2 * planeB + o) = s.p3;
    }
    if (warp == 0) tmem_alloc(&tmem_base_s, 256);
    if (tid == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    fence_smem_to_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_base_s;

    const int N = mode >= 2 ? 128 : 64;
    if (mode == 4) {
        // thread = lane v: A^T[v][k = row], two rows per 32-bit column, three planes of 32 columns at 128, 160, 192
#pragma unroll 1
        for (int c0 = 0; c0 < 32; c0 += 8) {
            uint32_t w[3][8];
            for (int c = 0; c < 8; ++c) {
                const Trunc3 lo = trunc3(A[(size_t)(2 * (c0 + c)) * 128 + tid]), hi = trunc3(A[(size_t)(2 * (c0 + c) + 1) * 128 + tid]);
                w[0][c] = pack_hi16(lo.t1, hi.t1), w[1][c] = pack_hi16(lo.t2, hi.t2), w[2][c] = pack_hi16(lo.t3, hi.t3);
            }
            for (int pl = 0; pl < 3; ++pl) tmem_st8(tmem + ((uint32_t)(32 * warp) << 16) + 128 + 32 * pl + c0, w[pl]);
        }
        fence_before_sync();
        __syncthreads();
        fence_after_sync();
        if (warp == 0) {
            OperandDesc b;
            b.base = smem_u32(sb), b.plane_bytes = planeB;
            b.lbo = x_chunk_bytes(RB), b.sbo = 128, b.kstep = 2 * x_chunk_bytes(RB);
            issue_gemm_bf16x3_ts_warp<4>(tmem, tmem + 128, tmem + 160, tmem + 192, b, idesc_bf16(128, 128, false, false), false);
            mma_commit_warp(&bar);
        }
    }
    if (mode != 4 && tid == 0) {
        OperandDesc a, b;
        a.base = smem_u32(sa), a.plane_bytes = planeA;
        b.base = smem_u32(sb), b.plane_bytes = planeB;
        int ksteps;
        uint32_t idesc;
        if (mode == 0) {  // K = features
            a.lbo = x_chunk_bytes(RA), a.sbo = 128, a.kstep = 2 * x_chunk_bytes(RA);
            b.lbo = x_chunk_bytes(RB), b.sbo = 128, b.kstep = 2 * x_chunk_bytes(RB);
            ksteps = 8, idesc = idesc_bf16(128, 64, false, false);
        } else if (mode == 1) {  // A: M = features, K = rows(128); B: K = features
            a.sbo = x_chunk_bytes(RA), a.lbo = 128, a.kstep = 2 * 128;
            b.lbo = x_chunk_bytes(RB), b.sbo = 128, b.kstep = 2 * x_chunk_bytes(RB);
            ksteps = 8, idesc = idesc_bf16(128, 64, true, false);
        } else if (mode == 2) {  // both: M/N = features, K = rows(64)
            a.sbo = x_chunk_bytes(RA), a.lbo = 128, a.kstep = 2 * 128;
            b.sbo = x_chunk_bytes(RB), b.lbo = 128, b.kstep = 2 * 128;
            ksteps = 4, idesc = idesc_bf16(128, 128, true, true);
        } else {  // A: M = rows(128), K = features 0..63; B: N = features, K = rows(64)
            a.lbo = x_chunk_bytes(RA), a.sbo = 128, a.kstep = 2 * x_chunk_bytes(RA);
            b.sbo = x_chunk_bytes(RB), b.lbo = 128, b.kstep = 2 * 128;
            ksteps = 4, idesc = idesc_bf16(128, 128, false, true);
        }
        issue_gemm_bf16x3(tmem, a, b, idesc, ksteps, false);
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    fence_after_sync();

    // epilogue: warp w owns TMEM lanes 32w..32w+31 (= rows of D)
    for (int c0 = 0; c0 < N; c0 += 32) {
        float v[32];
        tmem_ld32(tmem + ((uint32_t)(32 * warp) << 16) + c0, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) D[(size_t)tid * N + c0 + i] = v[i];
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

}  // namespace
}  // namespace rrl

extern "C" int32_t rrl_debug_tc_gemm(int32_t mode, const float *A, const float *B, float *D, void *stream)
{
    using namespace rrl;
    if (mode < 0 || mode > 4 || !A || !B || !D) return RRL_EINVAL;
    const size_t smem = 3 * (size_t)rrl::tc::x_plane_bytes((mode == 2 || mode == 4) ? 64 : 128, 128) + 3 * (size_t)rrl::tc::x_plane_bytes(mode == 4 ? 128 : 64, 128);
    if (cudaFuncSetAttribute(tc_gemm_test_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return RRL_ECUDA;
    tc_gemm_test_kernel<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(mode, A, B, D);
    return cudaGetLastError() == cudaSuccess ? RRL_OK : RRL_ECUDA;
}

// ---- micro-benchmark: how fast does one thread issue, and the tensor core retire, M=128 x N x 16 MMAs? -----
namespace rrl {
namespace {
using namespace tc;
template <bool TS>
__global__ void __launch_bounds__(128, 1) tc_rate_kernel(int N, int a_mn, int b_mn, int count, int uniform_issue, long long *out)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t planeA = x_plane_bytes(128, 128), planeB = x_plane_bytes(N, 128);
    for (uint32_t i = tid; i < (planeA + planeB) / 4; i += 128) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (warp == 0) tmem_alloc(&tmem_base_s, 256);
    if (tid == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    fence_smem_to_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_base_s;
    if (uniform_issue && warp == 0) {
        // the whole warp runs the loop; descriptors stay warp-uniform; one elected lane issues
        const uint32_t abase = smem_u32(smem), bbase = smem_u32(smem + planeA);
        const uint32_t a_lbo = a_mn ? 128 : x_chunk_bytes(128), a_sbo = a_mn ? x_chunk_bytes(128) : 128;
        const uint32_t b_lbo = b_mn ? 128 : x_chunk_bytes(N), b_sbo = b_mn ? x_chunk_bytes(N) : 128;
        const uint32_t a_k = a_mn ? 256 : 2 * x_chunk_bytes(128), b_k = b_mn ? 256 : 2 * x_chunk_bytes(N);
        const uint32_t idesc = idesc_bf16(128, N, a_mn != 0, b_mn != 0);
        const uint64_t da = smem_desc(abase, a_lbo, a_sbo), db = smem_desc(bbase, b_lbo, b_sbo);
        const int ksteps = (a_mn || b_mn) ? min(8, N / 16) : 8;
        const long long t0 = clock64();
        for (int i = 0; i < count; i += 4) {
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                const int ss = s % ksteps;
                if (TS)  // A from tensor memory (columns 128..191), N <= 128
                    mma_bf16_ts_elect(tmem, tmem + 128 + 8 * ss, db + ((ss * b_k) >> 4), idesc_bf16(128, N, false, b_mn != 0), (i + s) > 0);
                else
                    mma_bf16_elect(tmem, da + ((ss * a_k) >> 4), db + ((ss * b_k) >> 4), idesc, (i + s) > 0);
            }
        }
        if (elect_one()) mma_commit(&bar);
        __syncwarp();
        const long long t1 = clock64();
        mbar_wait(&bar, 0);
        const long long t2 = clock64();
        if (tid == 0) out[0] = t1 - t0, out[1] = t2 - t0;
    }
    if (!uniform_issue && tid == 0) {
        OperandDesc a, b;
        a.base = smem_u32(smem), b.base = smem_u32(smem + planeA);
        if (a_mn) a.sbo = x_chunk_bytes(128), a.lbo = 128, a.kstep = 256;
        else a.lbo = x_chunk_bytes(128), a.sbo = 128, a.kstep = 2 * x_chunk_bytes(128);
        if (b_mn) b.sbo = x_chunk_bytes(N), b.lbo = 128, b.kstep = 256;
        else b.lbo = x_chunk_bytes(N), b.sbo = 128, b.kstep = 2 * x_chunk_bytes(N);
        const uint32_t idesc = idesc_bf16(128, N, a_mn != 0, b_mn != 0);
        const int ksteps = (a_mn || b_mn) ? min(8, N / 16) : 8;
        const long long t0 = clock64();
        for (int i = 0; i < count; ++i) {
            const int s = i % ksteps;
            mma_bf16(tmem, smem_desc(a.base + s * a.kstep, a.lbo, a.sbo), smem_desc(b.base + s * b.kstep, b.lbo, b.sbo), idesc, i > 0);
        }
        mma_commit(&bar);
        const long long t1 = clock64();
        mbar_wait(&bar, 0);
        const long long t2 = clock64();
        out[0] = t1 - t0, out[1] = t2 - t0;
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}
}  // namespace
}  // namespace rrl

extern "C" int32_t rrl_debug_tc_rate(int32_t N, int32_t a_mn, int32_t b_mn, int32_t count, int32_t uniform_issue, long long *out_dev, void *stream)
{
    using namespace rrl;
    if (N < 16 || N > 256 || N % 16 || count < 4 || count % 4 || !out_dev || (a_mn == 2 && (N > 128 || !uniform_issue))) return RRL_EINVAL;
    const size_t smem = (size_t)rrl::tc::x_plane_bytes(128, 128) + rrl::tc::x_plane_bytes(N, 128);
    if (a_mn == 2) {
        if (cudaFuncSetAttribute(tc_rate_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return RRL_ECUDA;
        tc_rate_kernel<true><<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(N, 0, b_mn, count, uniform_issue, out_dev);
    } else {
        if (cudaFuncSetAttribute(tc_rate_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return RRL_ECUDA;
        tc_rate_kernel<false><<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(N, a_mn, b_mn, count, uniform_issue, out_dev);
    }
    return cudaGetLastError() == cudaSuccess ? RRL_OK : RRL_ECUDA;
}
