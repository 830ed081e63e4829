// K4-K6 on the 5th-generation tensor cores: the fused policy / value forward+backward pass with its
// three 128x128 GEMMs per 64-transition tile issued as tcgen05.mma (bf16x3 split => fp32-accurate,
// accumulators in TMEM).  Same contract, same outputs as the CUDA-core kernel in mlp.cu (which remains
// for hidden != 128, wide heads, large obs_dim and forward-only evaluation).
//
// Per tile (M = 64 transitions), with X[row][feature] operands in the un-swizzled "X layout"
// (tc_common.cuh) as three bf16 planes each:
//   stage 1:  H1 = relu(obs W1^T + b1)               CUDA cores (K = obs_dim is tiny) -> planes + relu bit mask
//   G1:       D1^T[out x row] = W2 * H1^T            A = W2 (K-major),  B = H1 (K-major),  N = 64
//   epilogue 1 (thread = hidden unit, 16 rows): h2 = relu(D1 + b2); head = W3 h2 (warp reduce-scatter);
//       loss, dhead per row; dz2 = dhead W3 [h2 > 0] -> planes stored transposed ([unit][row]: the owner of a
//       unit writes 8 rows per 16-byte store); dW3, db2 accumulate thread-locally
//   G3:       dW2[out x in] = dZ2^T * H1             A = dZ2^T (K-major), B = H1 (MN-major),   N = 128, K = 64 rows
//   G2:       D2^T[in x row] = W2^T * dZ2^T          A = W2 (MN-major),   B = dZ2^T (MN-major), N = 64
//   epilogue 2 (thread = hidden unit): dz1 = D2 [h1 > 0]; dW1, db1 accumulate thread-locally
// Schedule (the operand planes are single buffered -- W2 + H1 + dZ2 fill shared memory -- so the overlap
// comes from the order of work): G3 goes first; as soon as it retires the H1 planes are free and stage 1 of
// the NEXT tile runs under G2; G1 of the next tile is issued before epilogue 2 of this one and runs under it.
// The tile inputs are fetched from global memory one tile ahead.  dW2 is drained from TMEM into per-thread
// f32 sums after every tile; every TC_FLUSH (32) tiles everything spills into the CTA's float64 partials
// exactly like the CUDA-core kernel.
#include "mlp_internal.cuh"
#include "tc_common.cuh"

namespace rrl {

namespace {

using namespace tc;

constexpr int TC_H = 128;
constexpr int TC_MAX_D = 8;
constexpr int TC_MAX_OUT = 4;
constexpr int W_PLANE = tc::x_plane_bytes(TC_H, TC_H);    // bytes of one bf16 plane of W2
constexpr int W_CHUNK = tc::x_chunk_bytes(TC_H);
// H1 and dZ2 are kept transposed: X layout of [unit][row], so that the thread that owns a unit stores 8 rows per 16 bytes
constexpr int DZ_PLANE = tc::x_plane_bytes(TC_H, TILE_M), DZ_CHUNK = tc::x_chunk_bytes(TC_H);
// TMEM columns: D1[2] (64 each; after epilogue 1 has read it, D1[k&1] is reused for planes 0 and 1 of H1^T, G3's A
// operand), plane 2 of H1^T (32), D2 (64), dW2^T (128, one tile), ACC (128, the window's dW2^T sum)
constexpr uint32_t TM_D1 = 0, TM_P2 = 128, TM_D2 = 160, TM_DW2 = 224, TM_ACC = 352, TM_COLS = 512;
// f32 accumulation window of this kernel: 32 tiles = 2048 transitions, then the float64 spill (the CUDA-core kernel
// uses 16; here a spill costs two barriers and 32 reductions per thread: 5.5 % of a tile at 16, 2.7 % at 32)
constexpr int TC_FLUSH = 2 * FLUSH_TILES;
// Accuracy note.  The tensor core adds each MMA into its fp32 accumulator by CHOPPING (toward zero), not rounding:
// -4.3e-8 of |accumulator| per MMA on average (tools/tc_bias.py, profiles/r1_summary.md).  The small cross terms of the
// bf16x3 product are issued first, so what matters are the K/16 main MMAs at the end.  When the terms of a sum share a sign
// the result comes out up to ~2e-7 (relative) short -- within float32 rounding of any single number, but SYSTEMATIC.  On
// C1-sized batches (~2500 transitions) late in training, where the value targets are large and the residuals small, that
// made every v about 2e-7 too small and the value-head bias gradient 1.1e-5 short; Adam normalises gradients, and a
// quarter of the weights drifted by up to 1.8 lr per epoch against the float64 run (tools/diag_c1_grad.py, measured).
// A constant correction factor was tried and rejected: for sums whose terms cancel (fresh weights) there is no such bias
// and the factor then costs the golden epoch its 1e-5 agreement.  Hence TC_MIN_ROWS: batches too small to fill the
// pipeline anyway (where the effect is visible) take the CUDA-core kernel, whose float32 FMAs round to nearest.
constexpr int64_t TC_MIN_ROWS = 32768;  // ~3.5 tiles per SM; below it the CUDA-core kernel runs (see the note above)
constexpr int NBUF = 5;  // input buffers: tile k-1 (epilogue 2, which runs on into the next iteration), k (epilogue 1), k+1 (stage 1),
                         // k+2 (landed), k+3 (in flight from the top of the next iteration)
constexpr int TC_THREADS = 512;            // 16 worker warps: TMEM lane quarter q = warp & 3, row group rq = warp >> 2
// + one warp that only issues the MMAs (the issuing thread is held until its MMAs retire: the queue is a few
// entries deep).  17 warps are allocated as 20, i.e. 96 registers per thread: the dW2 sums therefore live in
// TMEM (columns TM_ACC..), not in registers.
constexpr int TC_CTA_THREADS = TC_THREADS + 32;
constexpr int TC_RG = TC_THREADS / 128;    // row groups per tile
constexpr int TC_RPT = TILE_M / TC_RG;     // rows per thread in the epilogues (16)

// shared-memory map (bytes)
constexpr int OFF_W2 = 0;
constexpr int OFF_H1 = OFF_W2 + 3 * W_PLANE;
constexpr int OFF_DZ = OFF_H1 + 3 * DZ_PLANE;  // (H1 is kept transposed like dZ2: [unit][row])
constexpr int OFF_OBS = OFF_DZ + 3 * DZ_PLANE;                // f32 [NBUF][64][8]
constexpr int OFF_W1 = OFF_OBS + NBUF * TILE_M * TC_MAX_D * 4;  // f32 [128][8]
constexpr int OFF_B1 = OFF_W1 + TC_H * TC_MAX_D * 4;          // f32 [128]
constexpr int OFF_B2 = OFF_B1 + TC_H * 4;
constexpr int OFF_W3 = OFF_B2 + TC_H * 4;                     // f32 [4][128]
constexpr int OFF_B3 = OFF_W3 + TC_MAX_OUT * TC_H * 4;        // f32 [4] (+pad)
constexpr int OFF_VPART = OFF_B3 + 64;                        // f32 [4 quarters][4][64]
constexpr int OFF_MASK = OFF_VPART + 4 * TC_MAX_OUT * TILE_M * 4;      // f32 [NBUF][64][4]
constexpr int OFF_ROWF = OFF_MASK + NBUF * TILE_M * TC_MAX_OUT * 4;    // f32 [NBUF][3][64]: ret | adv, logp_old, act
constexpr int OFF_RED = OFF_ROWF + NBUF * 3 * TILE_M * 4;              // double [16]
constexpr int OFF_BAR = OFF_RED + 128;                                 // 5 mbarriers + tmem base
constexpr int TC_SMEM_BYTES = OFF_BAR + 64;
static_assert(TC_SMEM_BYTES <= 227 * 1024, "tensor-core learner kernel exceeds shared memory");
static_assert(TILE_M * TC_MAX_D == TC_THREADS, "one observation element per thread in the input fetch");
static_assert(TC_RG * TC_H * (2 + TC_MAX_OUT + TC_MAX_D) * 4 + TILE_M * TC_MAX_OUT * 4 <= 3 * DZ_PLANE, "spill staging must fit in the dZ2 planes");

// 16 values per lane: lane l returns the total over the 32 lanes for c == (l & 15)
__device__ __forceinline__ float reduce_scatter16(float (&p)[16], int lane)
{
#pragma unroll
    for (int i = 0; i < 16; ++i) p[i] += __shfl_xor_sync(0xffffffffu, p[i], 16);
#pragma unroll
    for (int w = 8; w >= 1; w >>= 1) {
        const bool upper = (lane & w) != 0;
#pragma unroll
        for (int i = 0; i < w; ++i) {
            const float send = upper ? p[i] : p[i + w];
            const float keep = upper ? p[i + w] : p[i];
            p[i] = keep + __shfl_xor_sync(0xffffffffu, send, w);
        }
    }
    return p[0];
}

#ifdef RRL_TC_TIMING
__device__ unsigned long long g_tc_timing[16];
#define TCT_DECL long long tct_prev = clock64()
#define TCT(slot)                                                                      \
    do {                                                                               \
        const long long tct_now = clock64();                                           \
        if (blockIdx.x == 0 && lane == 0) atomicAdd(&g_tc_timing[slot], (unsigned long long)(tct_now - tct_prev)); \
        tct_prev = tct_now;                                                            \
    } while (0)
#else
#define TCT_DECL
#define TCT(slot)
#endif

// D4 = ceil(obs_dim / 4) as a compile-time constant: with a run-time bound the obs_dim = 4 case issued every FFMA and
// LDS of the unused upper half predicated off -- half of all FFMA issue slots
template <int MODE, int D4>
__global__ void __launch_bounds__(TC_CTA_THREADS, 1) mlp_tc_kernel(MlpArgs a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char *w2_p = smem + OFF_W2, *h1_p = smem + OFF_H1, *dz_p = smem + OFF_DZ;
    float *obs_s = reinterpret_cast<float *>(smem + OFF_OBS), *w1_s = reinterpret_cast<float *>(smem + OFF_W1);
    float *b1_s = reinterpret_cast<float *>(smem + OFF_B1), *b2_s = reinterpret_cast<float *>(smem + OFF_B2);
    float *w3_s = reinterpret_cast<float *>(smem + OFF_W3), *b3_s = reinterpret_cast<float *>(smem + OFF_B3);
    float *vpart_s = reinterpret_cast<float *>(smem + OFF_VPART), *mask_s = reinterpret_cast<float *>(smem + OFF_MASK);
    float *rowf_s = reinterpret_cast<float *>(smem + OFF_ROWF);
    double *red_s = reinterpret_cast<double *>(smem + OFF_RED);
    // bar1/2/3: G1/G2/G3 retired (tensor core -> workers); go1/go23: operands staged (workers -> MMA warp)
    uint64_t *bar1 = reinterpret_cast<uint64_t *>(smem + OFF_BAR), *bar2 = bar1 + 1, *bar3 = bar1 + 2, *go1 = bar1 + 3, *go23 = bar1 + 4;
    uint32_t *tmem_s = reinterpret_cast<uint32_t *>(bar1 + 5);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q = warp & 3, rq = warp >> 2;  // TMEM lane quarter of this warp, which group of 16 rows
    const int j = 32 * q + lane;             // the hidden unit this thread owns in both epilogues
    const int d = a.d, out = a.out;
    const NetDims nd{d, TC_H, out};

    // ---- one-time set-up: weights -> shared memory (W2 as three bf16 planes), TMEM, mbarriers ------
    {
        const float *W1 = a.params + nd.oW1(), *W2 = a.params + nd.oW2(), *W3 = a.params + nd.oW3();
        for (int i = tid; i < TC_H * TC_H; i += TC_CTA_THREADS) {
            const Trunc3 s = trunc3(W2[i]);
            const int o = x_off_bytes(i >> 7, i & 127, TC_H);  // row = out unit, feature = in unit
            *reinterpret_cast<unsigned short *>(w2_p + o) = hi16(s.t1);
            *reinterpret_cast<unsigned short *>(w2_p + W_PLANE + o) = hi16(s.t2);
            *reinterpret_cast<unsigned short *>(w2_p + 2 * W_PLANE + o) = hi16(s.t3);
        }
        for (int i = tid; i < TC_H * TC_MAX_D; i += TC_CTA_THREADS) {
            const int u = i / TC_MAX_D, c = i - u * TC_MAX_D;
            w1_s[i] = c < d ? W1[u * d + c] : 0.f;
        }
        for (int i = tid; i < TC_MAX_OUT * TC_H; i += TC_CTA_THREADS) w3_s[i] = i < out * TC_H ? W3[i] : 0.f;
        if (tid < TC_H) b1_s[tid] = a.params[nd.ob1() + tid], b2_s[tid] = a.params[nd.ob2() + tid];
        if (tid < TC_MAX_OUT) b3_s[tid] = tid < out ? a.params[nd.ob3() + tid] : 0.f;
        if (warp == 0) tmem_alloc(tmem_s, TM_COLS);
        if (tid == 0) {
            mbar_init(bar1, 1);
            mbar_init(bar2, 1);
            mbar_init(bar3, 1);
            mbar_init(go1, 1);
            mbar_init(go23, 1);
            mbar_fence_init();
        }
    }
    fence_smem_to_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_s;
    const uint32_t tmem_lane = tmem + ((uint32_t)(32 * q) << 16);

    float mean = 0.f, sd = 1.f;
    if (MODE != MODE_VALUE) mean = (float)a.stats[3], sd = (float)a.stats[4];

    // operand descriptors (tc_common.cuh: K-major -> LBO = R*16, SBO = 128; MN-major -> SBO = R*16, LBO = 128)
    OperandDesc W2_k, W2_mn, H1_k, H1_mn, DZ_k, DZ_mn;
    W2_k.base = W2_mn.base = smem_u32(w2_p), W2_k.plane_bytes = W2_mn.plane_bytes = W_PLANE;
    W2_k.lbo = W_CHUNK, W2_k.sbo = 128, W2_k.kstep = 2 * W_CHUNK;
    W2_mn.sbo = W_CHUNK, W2_mn.lbo = 128, W2_mn.kstep = 256;
    // H1^T [unit][row]: MN-major as G1's B (N = row, K = unit) -- the same view dZ2^T has as G2's B
    H1_k.base = H1_mn.base = smem_u32(h1_p), H1_k.plane_bytes = H1_mn.plane_bytes = DZ_PLANE;
    H1_k.lbo = DZ_CHUNK, H1_k.sbo = 128, H1_k.kstep = 2 * DZ_CHUNK;
    H1_mn.sbo = DZ_CHUNK, H1_mn.lbo = 128, H1_mn.kstep = 256;
    // dZ2^T [unit][row]: K-major as G3's A (M = unit, K = row), MN-major as G2's B (N = row, K = unit)
    DZ_k.base = DZ_mn.base = smem_u32(dz_p), DZ_k.plane_bytes = DZ_mn.plane_bytes = DZ_PLANE;
    DZ_k.lbo = DZ_CHUNK, DZ_k.sbo = 128, DZ_k.kstep = 2 * DZ_CHUNK;
    DZ_mn.sbo = DZ_CHUNK, DZ_mn.lbo = 128, DZ_mn.kstep = 256;
    constexpr uint32_t ID_G1 = idesc_bf16(128, 64, false, true), ID_G2 = idesc_bf16(128, 64, true, true),
                       ID_G3 = idesc_bf16(128, 128, false, false);

    // thread-local gradient accumulators for hidden unit j over this thread's quarter of the rows.
    // dW2 row j, this warp's 32 in-units: drained from TMEM after every tile.  The tensor core adds into
    // its fp32 accumulator with truncation (measured bias -4e-8 per MMA), so letting dW2 ride in TMEM
    // across tiles cost ~4e-6 of accuracy after 8 tiles; 24 MMAs per drain keep it at the CUDA-core level.
    // (the sums themselves sit in TMEM columns TM_ACC.., same lane/column ownership as the drain)
    float dW3acc[TC_MAX_OUT], dW1acc[4 * D4], db1acc = 0.f, db2acc = 0.f, db3acc[TC_MAX_OUT];
#pragma unroll
    for (int i = 0; i < TC_MAX_OUT; ++i) dW3acc[i] = 0.f, db3acc[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 4 * D4; ++i) dW1acc[i] = 0.f;
    double l0 = 0.0, l1 = 0.0, l2 = 0.0;
    bool first_flush = true;
    double *my_partials = a.partials + (size_t)blockIdx.x * a.n_params;
    const int64_t num_tiles = (a.n + TILE_M - 1) / TILE_M;

    // ---- per-tile building blocks ----------------------------------------------------------------------
    // the inputs of a tile travel global -> shared memory buffer b by cp.async (no registers), one tile ahead;
    // out-of-range elements are zero filled (src-size 0)
    auto cp4 = [&](void *dst, const float *src, bool ok) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_u32(dst)), "l"(ok ? src : a.obs), "r"(ok ? 4 : 0) : "memory");
    };
    auto fetch_async = [&](int64_t tile, int b) {
        const int64_t row0 = tile * TILE_M;
        const int rows = (int)min((int64_t)TILE_M, a.n - row0);
        {
            const int m = tid / TC_MAX_D, c = tid - m * TC_MAX_D;
            cp4(&obs_s[b * TILE_M * TC_MAX_D + tid], a.obs + (row0 + m) * d + c, m < rows && c < d);
        }
        if (MODE != MODE_VALUE && tid < TILE_M * TC_MAX_OUT) {
            const int m = tid / TC_MAX_OUT, c = tid - m * TC_MAX_OUT;
            cp4(&mask_s[b * TILE_M * TC_MAX_OUT + tid], a.mask + (row0 + m) * out + c, m < rows && c < out);
        }
        if (tid >= TC_THREADS - TILE_M) {  // the last two warps fetch the per-row scalars
            const int m = tid - (TC_THREADS - TILE_M);
            const bool ok = m < rows;
            float *rf = rowf_s + b * 3 * TILE_M;
            if (MODE == MODE_VALUE) {
                cp4(&rf[m], a.ret + row0 + m, ok);
            } else {
                cp4(&rf[m], a.adv + row0 + m, ok);
                cp4(&rf[TILE_M + m], a.logp_old + row0 + m, ok);
                cp4(&rf[2 * TILE_M + m], a.act + row0 + m, ok);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto inputs_landed = [&]() { asm volatile("cp.async.wait_all;" ::: "memory"); };
    // H1 of 16 rows for the hidden unit this thread owns, as bf16x3 words (two rows per word) + the relu bit mask.  Used
    // twice per tile with the same thread mapping and operation order, so both copies of H1 are identical: by stage 1
    // (-> shared-memory planes, G1's B operand) and by epilogue 1 (-> tensor memory, G3's A operand).
    auto h1_words = [&](int b, uint32_t (&hk)[3][TC_RPT / 2], uint32_t &hbits) {
        const float *ob = obs_s + b * TILE_M * TC_MAX_D;
        float4 w1r[D4];
#pragma unroll
        for (int i4 = 0; i4 < D4; ++i4) w1r[i4] = *reinterpret_cast<const float4 *>(&w1_s[j * TC_MAX_D + 4 * i4]);
        const float b1 = b1_s[j];
        hbits = 0u;
#pragma unroll
        for (int c = 0; c < TC_RPT; c += 2) {
            Trunc3 s2[2];
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int m = TC_RPT * rq + c + t;
                float acc = b1;
#pragma unroll
                for (int i4 = 0; i4 < D4; ++i4) {  // the observation row is one broadcast load for the whole warp
                    const float4 x = *reinterpret_cast<const float4 *>(&ob[m * TC_MAX_D + 4 * i4]);
                    acc = fmaf(x.x, w1r[i4].x, acc), acc = fmaf(x.y, w1r[i4].y, acc);
                    acc = fmaf(x.z, w1r[i4].z, acc), acc = fmaf(x.w, w1r[i4].w, acc);
                }
                hbits |= (acc > 0.f ? 1u : 0u) << (c + t);
                s2[t] = trunc3(fmaxf(acc, 0.f));
            }
            hk[0][c >> 1] = pack_hi16(s2[0].t1, s2[1].t1), hk[1][c >> 1] = pack_hi16(s2[0].t2, s2[1].t2),
            hk[2][c >> 1] = pack_hi16(s2[0].t3, s2[1].t3);
        }
    };
    // stage 1: H1^T planes ([unit][row], like dZ2: the owner of a unit stores 8 rows per 16 bytes)
    auto stage1 = [&](int b) {
        uint32_t hk[3][TC_RPT / 2], hbits;
        h1_words(b, hk, hbits);
#pragma unroll
        for (int pl = 0; pl < 3; ++pl) {
            unsigned char *dst = h1_p + pl * DZ_PLANE + (2 * rq) * DZ_CHUNK + j * 16;  // rows 16 rq .. +7, then +8 .. +15
            *reinterpret_cast<uint4 *>(dst) = make_uint4(hk[pl][0], hk[pl][1], hk[pl][2], hk[pl][3]);
            *reinterpret_cast<uint4 *>(dst + DZ_CHUNK) = make_uint4(hk[pl][4], hk[pl][5], hk[pl][6], hk[pl][7]);
        }
    };
    // the 16 worker warps meet at a named barrier; the MMA warp never joins it
    auto worker_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(TC_THREADS) : "memory"); };

    if (warp == TC_THREADS / 32) {
        // ---- the MMA warp: follows the workers' hand-offs tile by tile ------------------------------------
        // order on the tensor pipe: G1(0); then per tile k: G1(k+1) (runs under epilogue 1 of tile k), G3(k), G2(k)
        TCT_DECL;
        auto g1 = [&](int kk) {
            mbar_wait(go1, (uint32_t)(kk & 1));
            TCT(3);
            fence_after_sync();
#ifdef TC_DBG_SKIP_G1  // timing experiment only (make tc_variants; wrong results): one MMA instead of 48
            mma_bf16_elect(tmem + TM_D1 + 64 * (kk & 1), smem_desc(W2_k.base, W2_k.lbo, W2_k.sbo), smem_desc(H1_mn.base, H1_mn.lbo, H1_mn.sbo), ID_G1, false);
#else
            issue_gemm_bf16x3_warp<TC_H / 16>(tmem + TM_D1 + 64 * (kk & 1), W2_k, H1_mn, ID_G1, false);
#endif
            mma_commit_warp(bar1);
            TCT(4);
        };
        int k = 0;
        if ((int64_t)blockIdx.x < num_tiles) g1(0);
        for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++k) {
            if (tile + gridDim.x < num_tiles) g1(k + 1);
            mbar_wait(go23, (uint32_t)(k & 1));
            TCT(0);
            fence_after_sync();
            // G3: dW2^T[in x out] = H1^T * dZ2, A = H1^T from tensor memory (planes 0, 1 in the tile's D1 buffer)
            const uint32_t hT = tmem + TM_D1 + 64 * (k & 1);
            issue_gemm_bf16x3_ts_warp<TILE_M / 16>(tmem + TM_DW2, hT, hT + 32, tmem + TM_P2, DZ_k, ID_G3, false);
            mma_commit_warp(bar3);
            TCT(1);
#ifdef TC_DBG_SKIP_G2
            mma_bf16_elect(tmem + TM_D2, smem_desc(W2_mn.base, W2_mn.lbo, W2_mn.sbo), smem_desc(DZ_mn.base, DZ_mn.lbo, DZ_mn.sbo), ID_G2, false);
#else
            issue_gemm_bf16x3_warp<TC_H / 16>(tmem + TM_D2, W2_mn, DZ_mn, ID_G2, false);
#endif
            mma_commit_warp(bar2);
            TCT(2);
        }
    } else {

    uint32_t hb_carry = 0u;
    // ---- epilogue 2 of tile kk (inputs in buffer b3): dz1 = D2 [h1 > 0]; dW1, db1 ------------------------------
    // (split in two: the accumulator read-out, which must precede the hand-off that lets G2 of the NEXT tile overwrite D2,
    //  and the arithmetic, which runs after that hand-off, under the next tile's G3 and G2 -- with epilogue 2 as a whole
    //  ahead of epilogue 1 the workers waited 10 % of every tile for G2 to retire, measured with the timing variants)
    auto epilogue2_wait = [&](int kk) {
        mbar_wait(bar2, (uint32_t)(kk & 1));
        fence_after_sync();
    };
    auto epilogue2_load = [&](float (&dh)[TC_RPT]) { tmem_ld16(tmem_lane + TM_D2 + TC_RPT * rq, dh); };
    // hb: bit c = h1 > 0 for row 16 rq + c of that tile (set by this thread in its epilogue 1)
    auto epilogue2_math = [&](int b3, const float (&dh)[TC_RPT], uint32_t hb) {
        const float *ob = obs_s + b3 * TILE_M * TC_MAX_D;
        float dbl = 0.f;
#pragma unroll
        for (int c = 0; c < TC_RPT; ++c) {
            const int m = TC_RPT * rq + c;
            const float dz1 = ((hb >> c) & 1u) ? dh[c] : 0.f;
            dbl += dz1;
#pragma unroll
            for (int i4 = 0; i4 < D4; ++i4) {
                const float4 t = *reinterpret_cast<const float4 *>(&ob[m * TC_MAX_D + 4 * i4]);
                dW1acc[4 * i4] = fmaf(dz1, t.x, dW1acc[4 * i4]);
                dW1acc[4 * i4 + 1] = fmaf(dz1, t.y, dW1acc[4 * i4 + 1]);
                dW1acc[4 * i4 + 2] = fmaf(dz1, t.z, dW1acc[4 * i4 + 2]);
                dW1acc[4 * i4 + 3] = fmaf(dz1, t.w, dW1acc[4 * i4 + 3]);
            }
        }
        db1acc += dbl;
    };
    // ---- drain tile kk's dW2^T (TMEM lane = in unit j, this warp's 32 out units) into the window sum in TMEM ----
    auto drain = [&](int kk) {
        mbar_wait(bar3, (uint32_t)(kk & 1));  // (also: G3(kk) no longer reads the dZ2 planes)
        fence_after_sync();
        const bool first_in_window = kk % TC_FLUSH == 0;
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
            float v[16], s[16];
            if (!first_in_window) {  // (warp-uniform) both loads in flight, one wait
                tmem_ld16x2(tmem_lane + TM_DW2 + 32 * rq + 16 * hf, v, tmem_lane + TM_ACC + 32 * rq + 16 * hf, s);
#pragma unroll
                for (int c = 0; c < 16; ++c) v[c] += s[c];
            } else {
                tmem_ld16(tmem_lane + TM_DW2 + 32 * rq + 16 * hf, v);
            }
            tmem_st16(tmem_lane + TM_ACC + 32 * rq + 16 * hf, v);
        }
    };
    // ---- spill the f32 window into the CTA's float64 partials (staging in the dZ2 planes, idle at that point) ----
    auto spill = [&]() {
        // each float64 partial is owned by one thread: the first window stores, later windows add with fire-and-
        // forget reductions (no load latency on the path; one thread per address => fixed order).  dW2^T sits in
        // TMEM with lane = in unit: for a fixed out unit the 32 lanes of a warp hit 32 consecutive doubles.
        auto put = [&](int64_t p, float v) {
            if (first_flush)
                __stcg(my_partials + p, (double)v);
            else
                atomicAdd(my_partials + p, (double)v);
        };
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
            float v[16];
            tmem_ld16(tmem_lane + TM_ACC + 32 * rq + 16 * hf, v);
#pragma unroll
            for (int c = 0; c < 16; ++c) put(nd.oW2() + (int64_t)(32 * rq + 16 * hf + c) * TC_H + j, v[c]);
        }
        // per-unit accumulators: the row groups meet in shared memory
        float *stage = reinterpret_cast<float *>(dz_p);  // [TC_RG][128][S], S = 2 + MAX_OUT + MAX_D
        constexpr int S = 2 + TC_MAX_OUT + TC_MAX_D;
        float *mine = stage + (rq * TC_H + j) * S;
        mine[0] = db1acc, mine[1] = db2acc;
#pragma unroll
        for (int i = 0; i < TC_MAX_OUT; ++i) mine[2 + i] = dW3acc[i];
#pragma unroll
        for (int i = 0; i < TC_MAX_D; ++i) mine[2 + TC_MAX_OUT + i] = i < 4 * D4 ? dW1acc[i] : 0.f;
        float *stage3 = stage + TC_RG * TC_H * S;  // [64][MAX_OUT] per-row dhead sums
        if (q == 0 && lane < TC_RPT)
#pragma unroll
            for (int i = 0; i < TC_MAX_OUT; ++i) stage3[(TC_RPT * rq + lane) * TC_MAX_OUT + i] = db3acc[i];
        worker_sync();
        if (tid < TC_H) {
            auto tot = [&](int kk) {
                float t = 0.f;
#pragma unroll
                for (int g_ = 0; g_ < TC_RG; ++g_) t += stage[(g_ * TC_H + tid) * S + kk];
                return t;
            };
            put(nd.ob1() + tid, tot(0));
            put(nd.ob2() + tid, tot(1));
            for (int o_ = 0; o_ < out; ++o_) put(nd.oW3() + (int64_t)o_ * TC_H + tid, tot(2 + o_));
            for (int i = 0; i < d; ++i) put(nd.oW1() + (int64_t)tid * d + i, tot(2 + TC_MAX_OUT + i));
        }
        if (tid < out) {
            float s = 0.f;
            for (int m = 0; m < TILE_M; ++m) s += stage3[m * TC_MAX_OUT + tid];
            put(nd.ob3() + tid, s);
        }
        db1acc = 0.f, db2acc = 0.f;
#pragma unroll
        for (int i = 0; i < TC_MAX_OUT; ++i) dW3acc[i] = 0.f, db3acc[i] = 0.f;
#pragma unroll
        for (int i = 0; i < 4 * D4; ++i) dW1acc[i] = 0.f;
        first_flush = false;
        worker_sync();  // the staging area is read: epilogue 1 may write the dZ2 planes again
    };

    // ---- pipeline prologue: inputs and H1 of the first tile, G1 in flight ---------------------------------
    if ((int64_t)blockIdx.x < num_tiles) {
        fetch_async(blockIdx.x, 0);
        inputs_landed();
        worker_sync();
        stage1(0);
        fence_smem_to_async();
        fence_before_sync();
        worker_sync();
        if (tid == 0) mbar_arrive(go1);
        if ((int64_t)blockIdx.x + gridDim.x < num_tiles) fetch_async((int64_t)blockIdx.x + gridDim.x, 1);
    }

    int k = 0;
    TCT_DECL;
    for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++k) {
        const int b3 = k % NBUF, nb3 = (k + 1) % NBUF, pb3 = (k + NBUF - 1) % NBUF;  // this, next, previous tile's inputs
        const int64_t next = tile + gridDim.x, next2 = next + gridDim.x;
        const bool has_next = next < num_tiles;
        const float *rf = rowf_s + b3 * 3 * TILE_M;
        const int rows_k = (int)min((int64_t)TILE_M, a.n - tile * TILE_M);

        mbar_wait(bar1, (uint32_t)(k & 1));  // G1(k) retired: D1[k & 1] is ready and the H1 planes are free
        TCT(5);
        fence_after_sync();

        // ---- [A] stage 1 of the next tile; its G1 then runs under everything below --------------------------
        // the inputs of tile k+2 start their way in (its buffer held tile k-2, dead since the previous iteration);
        // those of tile k+1 landed before the barrier that ended the previous iteration
        if (next2 < num_tiles) fetch_async(next2, (k + 2) % NBUF);
        if (has_next) {
            if (k == 0) {  // tile 1 was fetched in the prologue
                inputs_landed();
                worker_sync();
            }
            stage1(nb3);
            fence_smem_to_async();
            fence_before_sync();
            worker_sync();
            if (tid == 0) mbar_arrive(go1);
        }
        TCT(8);

        // ---- [D] the previous tile's dW2 (its G3 retired long ago); every TC_FLUSH tiles the spill (the per-unit sums of
        //      epilogue 2 lag one tile behind: they join the next window) --------------------------------------------------
        const uint32_t hb_prev = hb_carry;  // tile k-1's relu mask, for its epilogue 2 at the end of this iteration
        if (k > 0) {
            drain(k - 1);
            TCT(7);
            if (k % TC_FLUSH == 0) {
                epilogue2_wait(k - 1);  // the spill stages in the dZ2 planes: G2 of tile k-1 must have read them
                spill();
            }
            TCT(11);
        }

        // ---- [B] epilogue 1 (thread = hidden unit j, rows 16*rq .. 16*rq+15) ---------------------------------------
        {
            float h2[TC_RPT];
            tmem_ld16(tmem_lane + TM_D1 + 64 * (k & 1) + TC_RPT * rq, h2);
            const float b2 = b2_s[j];
#pragma unroll
            for (int c = 0; c < TC_RPT; ++c) h2[c] = fmaxf(h2[c] + b2, 0.f);
            for (int o_ = 0; o_ < out; ++o_) {
                const float w3 = w3_s[o_ * TC_H + j];
                float p[TC_RPT];
#pragma unroll
                for (int c = 0; c < TC_RPT; ++c) p[c] = h2[c] * w3;
                const float tot = reduce_scatter16(p, lane);  // row 16*rq + (lane & 15), summed over this warp's 32 units
                if (lane < TC_RPT) vpart_s[(q * TC_MAX_OUT + o_) * TILE_M + TC_RPT * rq + lane] = tot;
            }
            TCT(12);
            fence_before_sync();
            worker_sync();  // (also: every warp has read D1[k & 1])
            fence_after_sync();
            TCT(13);
            // H1^T of this tile -> tensor memory (G3's A operand): recomputed exactly as stage 1 did; planes 0 and 1 go where
            // D1[k & 1] was, plane 2 has its own 32 columns
            {
                uint32_t hk[3][TC_RPT / 2], hbits;
                h1_words(b3, hk, hbits);
                const uint32_t hT = tmem_lane + TM_D1 + 64 * (k & 1) + (TC_RPT / 2) * rq;
                tmem_st8_nowait(hT, hk[0]);
                tmem_st8_nowait(hT + 32, hk[1]);
                tmem_st8_nowait(tmem_lane + TM_P2 + (TC_RPT / 2) * rq, hk[2]);
                tmem_st_wait();
                hb_carry = hbits;  // the relu mask epilogue 2 needs one tile later: same thread, stays in a register
            }
            TCT(14);
            // per-row loss and dhead: lane l of every warp works out row 16*rq + (l & 15); the four unit
            // quarters do it redundantly (it is a handful of instructions), only q == 0 keeps the sums
            float g_lane;
            int at_lane = 0;
            {
                const int m = TC_RPT * rq + (lane & (TC_RPT - 1));
                const bool ok = m < rows_k;
                const bool keeper = q == 0 && lane < TC_RPT;
                float head[TC_MAX_OUT];
#pragma unroll
                for (int o_ = 0; o_ < TC_MAX_OUT; ++o_)
                    head[o_] = ((vpart_s[(0 * TC_MAX_OUT + o_) * TILE_M + m] + vpart_s[(1 * TC_MAX_OUT + o_) * TILE_M + m]) +
                                (vpart_s[(2 * TC_MAX_OUT + o_) * TILE_M + m] + vpart_s[(3 * TC_MAX_OUT + o_) * TILE_M + m])) +
                               b3_s[o_];
                if (MODE == MODE_VALUE) {
                    const float diff = head[0] - rf[m];
                    g_lane = ok ? a.inv_n * (2.f * diff) : 0.f;  // REINFORCE.py:160 backward
                    if (q == 0) {  // warp-uniform branch: the float64 pipe is slow, keep it to four warps
                        if (ok && keeper) l0 += (double)diff * (double)diff;
                        if (keeper) db3acc[0] += g_lane;
                    }
                } else {
                    const float *mk = mask_s + b3 * TILE_M * TC_MAX_OUT;
                    float mx = -INFINITY;
#pragma unroll
                    for (int o_ = 0; o_ < TC_MAX_OUT; ++o_)
                        if (o_ < out) {
                            head[o_] = head[o_] + ((ok ? mk[m * TC_MAX_OUT + o_] : 1.f) - 1.f) * 1e8f;  // kernel.py:23-26
                            mx = fmaxf(mx, head[o_]);
                        }
                    float se = 0.f;
#pragma unroll
                    for (int o_ = 0; o_ < TC_MAX_OUT; ++o_)
                        if (o_ < out) se += expf(head[o_] - mx);
                    at_lane = ok ? min(max((int)rf[2 * TILE_M + m], 0), out - 1) : 0;
                    float ent = 0.f, lg_a = 0.f;
#pragma unroll
                    for (int o_ = 0; o_ < TC_MAX_OUT; ++o_)
                        if (o_ < out) {
                            ent += fmaxf(head[o_], -3.402823466e38f) * (expf(head[o_] - mx) / se);
                            if (o_ == at_lane) lg_a = head[o_];
                        }
                    const float advn = ok ? __fdiv_rn(rf[m] - mean, sd) : 0.f;  // replay_buffer.py:105
                    if (q == 0 && ok && keeper) {
                        l0 += (double)(lg_a * advn);  // logp is the raw masked logit
                        l1 += (double)(rf[TILE_M + m] - lg_a);
                        l2 += (double)ent;
                    }
                    g_lane = ok ? (-a.inv_n) * advn : 0.f;
                    if (keeper)
#pragma unroll
                        for (int o_ = 0; o_ < TC_MAX_OUT; ++o_) db3acc[o_] += (o_ == at_lane) ? g_lane : 0.f;
                }
            }
            TCT(15);
            // dz2 = dhead W3 [h2 > 0] -> transposed planes (two 16-byte pieces of 8 rows per plane); dW3, db2
            float dbl = 0.f;
            const float w3j = w3_s[j];
            uint32_t pk[3][TC_RPT / 2];
#pragma unroll
            for (int c = 0; c < TC_RPT; c += 2) {
                Trunc3 s2[2];
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    const float g = __shfl_sync(0xffffffffu, g_lane, c + t);
                    float dz;
                    if (MODE == MODE_VALUE) {
                        dz = h2[c + t] > 0.f ? g * w3j : 0.f;
                        dW3acc[0] = fmaf(g, h2[c + t], dW3acc[0]);
                    } else {
                        const int at = __shfl_sync(0xffffffffu, at_lane, c + t);
                        dz = h2[c + t] > 0.f ? g * w3_s[at * TC_H + j] : 0.f;
                        const float gh = g * h2[c + t];
#pragma unroll
                        for (int o_ = 0; o_ < TC_MAX_OUT; ++o_) dW3acc[o_] += (o_ == at) ? gh : 0.f;
                    }
                    dbl += dz;
                    s2[t] = trunc3(dz);
                }
                pk[0][c >> 1] = pack_hi16(s2[0].t1, s2[1].t1), pk[1][c >> 1] = pack_hi16(s2[0].t2, s2[1].t2), pk[2][c >> 1] = pack_hi16(s2[0].t3, s2[1].t3);
            }
            if (k > 0) epilogue2_wait(k - 1);  // G2 of tile k-1 has read the dZ2 planes (retired long ago by now) and D2 is complete
#pragma unroll
            for (int pl = 0; pl < 3; ++pl) {
                unsigned char *dst = dz_p + pl * DZ_PLANE + (2 * rq) * DZ_CHUNK + j * 16;  // rows 16 rq .. +7, then +8 .. +15
                *reinterpret_cast<uint4 *>(dst) = make_uint4(pk[pl][0], pk[pl][1], pk[pl][2], pk[pl][3]);
                *reinterpret_cast<uint4 *>(dst + DZ_CHUNK) = make_uint4(pk[pl][4], pk[pl][5], pk[pl][6], pk[pl][7]);
            }
            db2acc += dbl;
        }
        float dh_prev[TC_RPT];  // D2 of tile k-1, read out before G2 of this tile may overwrite it
        if (k > 0) epilogue2_load(dh_prev);
        inputs_landed();  // tile k+2's inputs (fetched a whole iteration ago) become visible with the barrier below
        fence_smem_to_async();
        fence_before_sync();
        worker_sync();
        if (tid == 0) mbar_arrive(go23);  // the MMA warp issues G3, then G2
        TCT(6);
        // ---- epilogue 2 of the previous tile, under this tile's G3 and G2 ---------------------------------------------
        if (k > 0) epilogue2_math(pb3, dh_prev, hb_prev);
        TCT(10);
    }
    // ---- pipeline tail: the last tile's drain, epilogue 2 and the final spill ------------------------------------
    if (k > 0) {
        drain(k - 1);
        epilogue2_wait(k - 1);
        float dh_last[TC_RPT];
        epilogue2_load(dh_last);
        epilogue2_math((k - 1) % NBUF, dh_last, hb_carry);
        spill();
    }
    }  // worker warps

    // ---- loss sums (rows live in lanes 0..15 of the q == 0 warps) --------------------------------------------
#pragma unroll
    for (int dd = 16; dd > 0; dd >>= 1) {
        l0 += __shfl_xor_sync(0xffffffffu, l0, dd);
        l1 += __shfl_xor_sync(0xffffffffu, l1, dd);
        l2 += __shfl_xor_sync(0xffffffffu, l2, dd);
    }
    if (lane == 0 && q == 0 && warp < TC_THREADS / 32) red_s[rq * 4 + 0] = l0, red_s[rq * 4 + 1] = l1, red_s[rq * 4 + 2] = l2;
    fence_before_sync();
    __syncthreads();
    if (tid == 0 && a.loss_partials) {
        double *lp = a.loss_partials + (size_t)blockIdx.x * 4;
        lp[0] = (red_s[0] + red_s[4]) + (red_s[8] + red_s[12]);
        lp[1] = (red_s[1] + red_s[5]) + (red_s[9] + red_s[13]);
        lp[2] = (red_s[2] + red_s[6]) + (red_s[10] + red_s[14]);
        lp[3] = 0.0;
    }
    if (warp == 0) tmem_dealloc(tmem, TM_COLS);
}

}  // namespace

#ifdef RRL_TC_TIMING
extern "C" int32_t rrl_debug_tc_timing(unsigned long long *out16, int32_t reset)
{
    if (out16 && cudaMemcpyFromSymbol(out16, g_tc_timing, sizeof(g_tc_timing)) != cudaSuccess) return RRL_ECUDA;
    if (reset) {
        unsigned long long z[16] = {};
        if (cudaMemcpyToSymbol(g_tc_timing, z, sizeof(z)) != cudaSuccess) return RRL_ECUDA;
    }
    return RRL_OK;
}
#endif

bool tc_pass_supported(const NetDims &nd, int mode, bool with_grad, int64_t rows)
{
    return rows >= TC_MIN_ROWS && with_grad && nd.H == TC_H && nd.d <= TC_MAX_D && nd.out <= TC_MAX_OUT && (mode == MODE_VALUE || mode == MODE_POLICY_REF);
}

size_t tc_smem_bytes() { return TC_SMEM_BYTES; }

template <int MODE, int D4>
static int32_t tc_launch_one(const MlpArgs &a, int grid, cudaStream_t stream, ErrBuf &err)
{
    RRL_CUDA_OK(cudaFuncSetAttribute(mlp_tc_kernel<MODE, D4>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES), err);
    mlp_tc_kernel<MODE, D4><<<grid, TC_CTA_THREADS, TC_SMEM_BYTES, stream>>>(a);
    RRL_CUDA_OK(cudaGetLastError(), err);
    return RRL_OK;
}

int32_t tc_pass_launch(const MlpArgs &a, const NetDims &nd, int mode, int grid, cudaStream_t stream, ErrBuf &err)
{
    const bool wide = nd.d > 4;
    if (mode == MODE_VALUE) return wide ? tc_launch_one<MODE_VALUE, 2>(a, grid, stream, err) : tc_launch_one<MODE_VALUE, 1>(a, grid, stream, err);
    return wide ? tc_launch_one<MODE_POLICY_REF, 2>(a, grid, stream, err) : tc_launch_one<MODE_POLICY_REF, 1>(a, grid, stream, err);
}

}  // namespace rrl
