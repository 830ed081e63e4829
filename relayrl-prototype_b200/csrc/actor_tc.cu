// K8 on the 5th-generation tensor cores: the batched actor step for wide layers (BASELINE config C4: 65 536
// environments, obs_dim 128, 2 x 256 hidden units, 18 actions).  Same contract and outputs as actor.cu, which stays
// for every other shape; replaces the same reference code (kernel.py:99-107,128-135 as called from
// RelayRLAgentZmq::request_for_action, src/network/client/agent_zmq.rs:458-571).
//
// One persistent CTA per SM walks tiles of 128 environments.  Environments are the M dimension of every MMA
// (TMEM lane = environment), hidden units the N dimension in chunks of 128, so that the epilogue thread that owns an
// environment reads its whole row of pre-activations from its own TMEM lane, and the next layer's A operand is written
// straight back into tensor memory (bf16 pairs per 32-bit column) without touching shared memory:
//   obs tile  -> three bf16 planes in shared memory (K-major)                      CUDA cores, once per tile
//   layer 1   : ACC[128 env x 128 units] = obs * W1^T   (A: shared memory, B: streamed)   tcgen05.mma, bf16x3
//   epilogue 1: h1 = relu(ACC + b1) -> three bf16 planes in TMEM columns 128..511
//   layer 2   : ACC = h1 * W2^T                         (A: tensor memory, B: streamed)
//   epilogue 2: h2 = relu(ACC + b2); head += h2 * W3^T  (thread = environment, W3 broadcast from shared memory)
//   tail      : masked softmax, Philox draw, outputs (actor_internal.cuh)
// for the policy net and then the value net.  The weights do not fit in shared memory (W2 alone is 384 KB as bf16x3),
// so they are packed ONCE per rrl_actor_set_params into the exact order and layout the kernel consumes and stream from
// L2 through a ring of 12 KB stages filled by TMA bulk copies (cp.async.bulk, no tensor map, this GPU's memory only).
//
// fp32 accuracy: a*b = a1b1 + (a1b2 + a2b1) + (a1b3 + a2b2 + a3b1) with bf16 parts (tc_common.cuh).  The tensor core
// adds into its fp32 accumulator by chopping, -4e-8 of |accumulator| per MMA (csrc/mlp_tc.cu "Accuracy note"), so the
// five small products of every K step are accumulated first (cross pass) and the K/16 main products last (main pass,
// which re-streams plane 1 of the weights: 4/3 of the bytes, 16 instead of 96 MMAs at full accumulator size).
#include "actor_internal.cuh"
#include <type_traits>

#include "tc_common.cuh"

namespace rrl {

namespace {

using namespace tc;

constexpr int AT_M = 128;           // environments per tile
constexpr int AT_NCHUNK = 128;      // hidden units per accumulator chunk
constexpr int AT_MAX_DK = 128;      // obs_dim rounded up to 16
constexpr int AT_MAX_A = 32;
constexpr int AT_SUB = 4096;        // bytes of one (k-step, plane) piece of a weight chunk: 128 units x 16 k, bf16
constexpr int AT_BLOCK = 3 * AT_SUB;  // one ring stage
constexpr int AT_STAGES = 6;
constexpr int AT_WORKERS = 256;     // 8 epilogue warps: TMEM lane quarter q = warp & 3, column half hf = warp >> 2
constexpr int AT_THREADS = AT_WORKERS + 64;  // + the MMA warp + the producer warp
constexpr uint32_t AT_TM_ACC = 0, AT_TM_A = 128, AT_TM_COLS = 512;  // accumulator | three planes of 128 columns

struct PackShape {
    int dK, ks1, ks2, nc, blocks1, blocks2;  // K steps of the two layers, unit chunks, ring blocks per (layer, chunk)
    __host__ __device__ int per_net() const { return nc * (blocks1 + blocks2); }
};
__host__ __device__ inline PackShape pack_shape(int d, int H)
{
    PackShape s;
    s.dK = (d + 15) & ~15, s.ks1 = s.dK / 16, s.ks2 = H / 16, s.nc = H / AT_NCHUNK;
    s.blocks1 = s.ks1 + (s.ks1 + 2) / 3, s.blocks2 = s.ks2 + (s.ks2 + 2) / 3;
    return s;
}

// Packed stream of one net: for layer in (1, 2): for chunk c: [cross blocks: K step s -> planes 1, 2, 3 of that step]
// [main blocks: group g -> plane 1 of K steps 3g, 3g+1, 3g+2].  A piece holds unit n (of the chunk), k (of the step) at
// byte (k / 8) * 2048 + n * 16 + (k % 8) * 2: the un-swizzled K-major core-matrix layout (LBO 2048, SBO 128).
__global__ void __launch_bounds__(256) actor_pack_kernel(const float *__restrict__ params, int d, int H, int out,
                                                         unsigned short *__restrict__ pack)
{
    const PackShape ps = pack_shape(d, H);
    const NetDims nd{d, H, out};
    const int64_t total = (int64_t)ps.per_net() * (AT_BLOCK / 2);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int blk = (int)(i / (AT_BLOCK / 2));
        const int e = (int)(i - (int64_t)blk * (AT_BLOCK / 2));
        const int sub = e / (AT_SUB / 2), w = e - sub * (AT_SUB / 2);  // element within the piece
        const int kc = w / 1024, n = (w - kc * 1024) / 8, k8 = w & 7;
        int layer = 1, per = ps.blocks1, ks = ps.ks1;
        if (blk >= ps.nc * ps.blocks1) blk -= ps.nc * ps.blocks1, layer = 2, per = ps.blocks2, ks = ps.ks2;
        const int c = blk / per, b = blk - c * per;
        int step, plane;
        if (b < ks)
            step = b, plane = sub;
        else
            step = 3 * (b - ks) + sub, plane = 0;
        const int unit = c * AT_NCHUNK + n, k = step * 16 + kc * 8 + k8;
        float v = 0.f;
        if (step < ks) {
            if (layer == 1)
                v = k < d ? params[nd.oW1() + (int64_t)unit * d + k] : 0.f;
            else
                v = params[nd.oW2() + (int64_t)unit * H + k];
        }
        const Trunc3 t = trunc3(v);
        pack[i] = hi16(plane == 0 ? t.t1 : plane == 1 ? t.t2 : t.t3);
    }
}

// shared-memory map (bytes)
constexpr int OBS_CHUNK = tc::x_chunk_bytes(AT_M);                    // 8 k of all 128 rows (+16 bytes of bank rotation)
constexpr int OBS_PLANE = (AT_MAX_DK / 8) * OBS_CHUNK;
constexpr int OFF_OBS = 0;
constexpr int OFF_RING = OFF_OBS + 3 * OBS_PLANE;
constexpr int OFF_W3PI = OFF_RING + AT_STAGES * AT_BLOCK;             // f32 [32][H], rows >= A zero
constexpr int OFF_W3V = OFF_W3PI + AT_MAX_A * 256 * 4;                // f32 [4][H], row 0 = the value head
constexpr int OFF_BIAS = OFF_W3V + 4 * 256 * 4;                       // f32 b1 pi | b2 pi | b1 v | b2 v, 256 each
constexpr int OFF_B3 = OFF_BIAS + 4 * 256 * 4;                        // f32 [32] policy + [1] value (+ pad)
constexpr int OFF_COMB = OFF_B3 + 256;                                // f32 [128][33]: the two column halves meet here
constexpr int OFF_VCOMB = OFF_COMB + AT_M * (AT_MAX_A + 1) * 4;       // f32 [128]
constexpr int OFF_BAR = OFF_VCOMB + AT_M * 4;                         // mbarriers: full[6] empty[6] acc_full acc_empty a_ready planes_ready
constexpr int AT_SMEM_BYTES = OFF_BAR + (2 * AT_STAGES + 4) * 8 + 8;
static_assert(AT_SMEM_BYTES <= 227 * 1024, "tensor-core actor kernel exceeds shared memory");
static_assert(OFF_RING % 128 == 0 && OFF_BAR % 8 == 0, "alignment");

// developer timing variants (make actor_variants; wrong results, never shipped)
#ifdef AT_DBG_NO_FULL_WAIT
#define AT_WAIT_FULL(bar, ph) ((void)0)
#else
#define AT_WAIT_FULL(bar, ph) mbar_wait(bar, ph)
#endif
#ifdef AT_DBG_NO_MMA
#define AT_MMA(x) ((void)0)
#else
#define AT_MMA(x) x
#endif

template <int H>
__global__ void __launch_bounds__(AT_THREADS, 1) actor_tc_kernel(ActorArgs a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char *obs_p = smem + OFF_OBS, *ring = smem + OFF_RING;
    float *w3pi_s = reinterpret_cast<float *>(smem + OFF_W3PI), *w3v_s = reinterpret_cast<float *>(smem + OFF_W3V);
    float *bias_s = reinterpret_cast<float *>(smem + OFF_BIAS), *b3_s = reinterpret_cast<float *>(smem + OFF_B3);
    float *comb_s = reinterpret_cast<float *>(smem + OFF_COMB), *vcomb_s = reinterpret_cast<float *>(smem + OFF_VCOMB);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + OFF_BAR), *empty = full + AT_STAGES;
    uint64_t *acc_full = empty + AT_STAGES, *acc_empty = acc_full + 1, *a_ready = acc_full + 2, *planes_ready = acc_full + 3;
    uint32_t *tmem_s = reinterpret_cast<uint32_t *>(planes_ready + 1);

    constexpr int NC = H / AT_NCHUNK;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int d = a.d, A = a.A, A4 = (A + 3) & ~3;
    const int nets = a.with_v ? 2 : 1;
    const PackShape ps = pack_shape(d, H);
    const NetDims pd{d, H, A}, vd{d, H, 1};
    const int64_t tiles = (a.batch + AT_M - 1) / AT_M;

    // ---- one-time set-up: head weights and biases -> shared memory, mbarriers, TMEM ------------------------------
    for (int i = tid; i < AT_MAX_A * H; i += AT_THREADS) w3pi_s[i] = i < A * H ? a.pi.params[pd.oW3() + i] : 0.f;
    for (int i = tid; i < 4 * H; i += AT_THREADS) w3v_s[i] = (a.with_v && i < H) ? a.vf.params[vd.oW3() + i] : 0.f;
    for (int i = tid; i < H; i += AT_THREADS) {
        bias_s[i] = a.pi.params[pd.ob1() + i], bias_s[256 + i] = a.pi.params[pd.ob2() + i];
        bias_s[512 + i] = a.with_v ? a.vf.params[vd.ob1() + i] : 0.f, bias_s[768 + i] = a.with_v ? a.vf.params[vd.ob2() + i] : 0.f;
    }
    if (tid < AT_MAX_A) b3_s[tid] = tid < A ? a.pi.params[pd.ob3() + tid] : 0.f;
    if (tid == AT_MAX_A) b3_s[tid] = a.with_v ? a.vf.params[vd.ob3()] : 0.f;
    if (warp == 0) tmem_alloc(tmem_s, AT_TM_COLS);
    if (tid == 0) {
        for (int s = 0; s < AT_STAGES; ++s) mbar_init(&full[s], 1), mbar_init(&empty[s], 1);
        mbar_init(acc_full, 1);
        mbar_init(acc_empty, AT_WORKERS / 32);
        mbar_init(a_ready, AT_WORKERS / 32);
        mbar_init(planes_ready, AT_WORKERS / 32);
        mbar_fence_init();
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_s;

    if (warp == AT_WORKERS / 32 + 1) {
        // ---- producer: the packed weights of each net, block after block, into the ring -------------------------------
        if (lane == 0) {
            uint32_t st = 0, ph = 0;
            const int per_net = ps.per_net();
            for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x)
                for (int net = 0; net < nets; ++net) {
                    const unsigned char *src = net ? a.vf.pack : a.pi.pack;
                    for (int b = 0; b < per_net; ++b) {
                        mbar_wait(&empty[st], ph ^ 1u);  // (passes at once the first time round)
                        mbar_arrive_expect_tx(&full[st], AT_BLOCK);
                        bulk_g2s(ring + st * AT_BLOCK, src + (size_t)b * AT_BLOCK, AT_BLOCK, &full[st]);
                        if (++st == AT_STAGES) st = 0, ph ^= 1u;
                    }
                }
        }
    } else if (warp == AT_WORKERS / 32) {
        // ---- the MMA warp (converged; one elected lane issues) --------------------------------------------------------
        constexpr uint32_t IDESC = idesc_bf16(AT_M, AT_NCHUNK, false, false);
        const uint32_t ring_u = smem_u32(ring), obs_u = smem_u32(obs_p);
        // Descriptors are built once per ring block (one move into uniform registers) and stepped by compile-time constants:
        // with a descriptor computed per MMA the issue loop ran at 161 cycles per MMA instead of the tensor core's 64.
        constexpr uint64_t SUBQ = AT_SUB >> 4, PLQ = OBS_PLANE >> 4;
        uint32_t st = 0, ph = 0, chunk_n = 0, tcount = 0, net_n = 0;
        auto next_stage = [&]() {
            if (++st == AT_STAGES) st = 0, ph ^= 1u;
        };
        // wait_at: the K step of layer 2 before which the planes written by the LAST epilogue of layer 1 must be complete (the
        // earlier K steps read planes of earlier chunks, so layer 2 starts while that epilogue is still splitting its values)
        auto run_chunk = [&](auto l1_tag, int ks, int wait_at, uint32_t wait_parity) {
            constexpr bool L1 = decltype(l1_tag)::value;
            const uint32_t acc_t = tmem + AT_TM_ACC;
            for (int s = 0; s < ks; ++s) {  // cross pass: the five small products of K step s
                if (s == wait_at) {
                    mbar_wait(planes_ready, wait_parity);
                    fence_after_sync();
                }
                AT_WAIT_FULL(&full[st], ph);
                const uint64_t bd = smem_desc(ring_u + st * AT_BLOCK, 2048, 128);
                if (L1) {
                    const uint64_t ad = smem_desc(obs_u + s * 2 * OBS_CHUNK, OBS_CHUNK, 128);
                    AT_MMA(mma_bf16_elect(acc_t, ad, bd + 2 * SUBQ, IDESC, s > 0));
                    AT_MMA(mma_bf16_elect(acc_t, ad + PLQ, bd + SUBQ, IDESC, true));
                    AT_MMA(mma_bf16_elect(acc_t, ad + 2 * PLQ, bd, IDESC, true));
                    AT_MMA(mma_bf16_elect(acc_t, ad, bd + SUBQ, IDESC, true));
                    AT_MMA(mma_bf16_elect(acc_t, ad + PLQ, bd, IDESC, true));
                } else {
                    const uint32_t at = tmem + AT_TM_A + 8 * s;
                    AT_MMA(mma_bf16_ts_elect(acc_t, at, bd + 2 * SUBQ, IDESC, s > 0));
                    AT_MMA(mma_bf16_ts_elect(acc_t, at + 128, bd + SUBQ, IDESC, true));
                    AT_MMA(mma_bf16_ts_elect(acc_t, at + 256, bd, IDESC, true));
                    AT_MMA(mma_bf16_ts_elect(acc_t, at, bd + SUBQ, IDESC, true));
                    AT_MMA(mma_bf16_ts_elect(acc_t, at + 128, bd, IDESC, true));
                }
                mma_commit_warp(&empty[st]);
                next_stage();
            }
            for (int s0 = 0; s0 < ks; s0 += 3) {  // main pass: plane 1 x plane 1, three K steps per block
                AT_WAIT_FULL(&full[st], ph);
                const uint64_t bd = smem_desc(ring_u + st * AT_BLOCK, 2048, 128);
#pragma unroll
                for (int j = 0; j < 3; ++j)
                    if (s0 + j < ks) {
                        if (L1)
                            AT_MMA(mma_bf16_elect(acc_t, smem_desc(obs_u + (s0 + j) * 2 * OBS_CHUNK, OBS_CHUNK, 128), bd + j * SUBQ, IDESC, true));
                        else
                            AT_MMA(mma_bf16_ts_elect(acc_t, tmem + AT_TM_A + 8 * (s0 + j), bd + j * SUBQ, IDESC, true));
                    }
                mma_commit_warp(&empty[st]);
                next_stage();
            }
        };
        for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++tcount) {
            mbar_wait(a_ready, tcount & 1u);  // this tile's observation planes are in shared memory
            fence_after_sync();
            for (int net = 0; net < nets; ++net, ++net_n)
                for (int layer = 1; layer <= 2; ++layer)
                    for (int c = 0; c < NC; ++c, ++chunk_n) {
                        mbar_wait(acc_empty, (chunk_n & 1u) ^ 1u);  // the epilogue has read the previous chunk
                        fence_after_sync();
                        if (layer == 1)
                            run_chunk(std::true_type{}, ps.ks1, -1, 0u);
                        else
                            run_chunk(std::false_type{}, H / 16, c == 0 ? (NC - 1) * (AT_NCHUNK / 16) : -1, net_n & 1u);
                        mma_commit_warp(acc_full);
                    }
        }
    } else {
        // ---- the eight worker warps: thread = environment (TMEM lane), column half hf ------------------------------------
        const int q = warp & 3, hf = warp >> 2;
        const int row = 32 * q + lane;
        const uint32_t tmem_lane = tmem + ((uint32_t)(32 * q) << 16);
        const bool vec = (d & 7) == 0 && (reinterpret_cast<uintptr_t>(a.obs) & 15) == 0;
        auto workers_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(AT_WORKERS) : "memory"); };
        auto arrive = [&](uint64_t *bar) {  // one arrival per warp, after every lane's TMEM / shared-memory traffic
            fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar);
        };
        // observations of one tile -> three bf16 planes (16-byte pieces: one environment, 8 consecutive features)
        auto convert_obs = [&](int64_t tile) {
            const int kcs = ps.dK / 8;
            const int64_t row0 = tile * AT_M;
            for (int p = tid; p < AT_M * kcs; p += AT_WORKERS) {
                const int r = p / kcs, kc = p - r * kcs, k0 = 8 * kc;
                const bool ok = row0 + r < a.batch;
                float x[8];
                if (vec) {
                    float4 u = make_float4(0.f, 0.f, 0.f, 0.f), w = u;
                    if (ok && k0 < d) {
                        const float4 *src = reinterpret_cast<const float4 *>(a.obs + (row0 + r) * d + k0);
                        u = __ldcs(src), w = __ldcs(src + 1);
                    }
                    x[0] = u.x, x[1] = u.y, x[2] = u.z, x[3] = u.w, x[4] = w.x, x[5] = w.y, x[6] = w.z, x[7] = w.w;
                } else {
#pragma unroll
                    for (int i = 0; i < 8; ++i) x[i] = (ok && k0 + i < d) ? __ldcs(a.obs + (row0 + r) * d + k0 + i) : 0.f;
                }
                uint32_t w1[4], w2[4], w3[4];
#pragma unroll
                for (int i = 0; i < 8; i += 2) {
                    const Trunc3 s0 = trunc3(x[i]), s1 = trunc3(x[i + 1]);
                    w1[i >> 1] = pack_hi16(s0.t1, s1.t1), w2[i >> 1] = pack_hi16(s0.t2, s1.t2), w3[i >> 1] = pack_hi16(s0.t3, s1.t3);
                }
                unsigned char *dst = obs_p + kc * OBS_CHUNK + r * 16;
                *reinterpret_cast<uint4 *>(dst) = make_uint4(w1[0], w1[1], w1[2], w1[3]);
                *reinterpret_cast<uint4 *>(dst + OBS_PLANE) = make_uint4(w2[0], w2[1], w2[2], w2[3]);
                *reinterpret_cast<uint4 *>(dst + 2 * OBS_PLANE) = make_uint4(w3[0], w3[1], w3[2], w3[3]);
            }
            fence_smem_to_async();  // generic-proxy writes -> visible to the tensor core
            arrive(a_ready);
        };
        // this thread's 64 accumulator columns of the current chunk, bias added, relu applied
        auto load_chunk = [&](float (&v)[64], const float *bias, int c) {
            tmem_ld64(tmem_lane + AT_TM_ACC + 64 * hf, v);
            const float *bb = bias + AT_NCHUNK * c + 64 * hf;
#pragma unroll
            for (int i = 0; i < 64; i += 4) {
                const float4 b = *reinterpret_cast<const float4 *>(bb + i);
                v[i] = fmaxf(v[i] + b.x, 0.f), v[i + 1] = fmaxf(v[i + 1] + b.y, 0.f);
                v[i + 2] = fmaxf(v[i + 2] + b.z, 0.f), v[i + 3] = fmaxf(v[i + 3] + b.w, 0.f);
            }
        };

        if ((int64_t)blockIdx.x < tiles) convert_obs(blockIdx.x);
        uint32_t chunk_n = 0;
        for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
            const int64_t grow = tile * AT_M + row;
            const bool valid = grow < a.batch;
            for (int net = 0; net < nets; ++net) {
                const float *b1 = bias_s + 512 * net, *b2 = b1 + 256;
                // ---- layer 1: h1 -> the A operand of layer 2 in tensor memory --------------------------------------------
                for (int c = 0; c < NC; ++c, ++chunk_n) {
                    mbar_wait(acc_full, chunk_n & 1u);
                    fence_after_sync();
                    float v[64];
                    load_chunk(v, b1, c);
                    const bool last = c + 1 == NC;
                    arrive(acc_empty);  // the values are in registers: the next chunk (layer 1, or the first K steps of layer 2) may start
#ifndef AT_DBG_NO_EPI
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        uint32_t p1[8], p2[8], p3[8];
#pragma unroll
                        for (int i = 0; i < 16; i += 2) {
                            const Trunc3 s0 = trunc3(v[16 * g + i]), s1 = trunc3(v[16 * g + i + 1]);
                            p1[i >> 1] = pack_hi16(s0.t1, s1.t1), p2[i >> 1] = pack_hi16(s0.t2, s1.t2), p3[i >> 1] = pack_hi16(s0.t3, s1.t3);
                        }
                        const uint32_t col = tmem_lane + AT_TM_A + 64 * c + 32 * hf + 8 * g;  // unit / 2
                        tmem_st8_nowait(col, p1), tmem_st8_nowait(col + 128, p2), tmem_st8_nowait(col + 256, p3);
                    }
                    tmem_st_wait();
#else
                    if (v[3] == 123.f) vcomb_s[row] = v[5];
#endif
                    if (last) {
                        arrive(planes_ready);  // layer 2 may now read the planes of this chunk's units too
                        // every MMA that reads the observation planes has retired: the next tile's can be built under layer 2
                        if (net == nets - 1 && tile + gridDim.x < tiles) convert_obs(tile + gridDim.x);
                    }
                }
                // ---- layer 2 and the head -----------------------------------------------------------------------------------
                float head[AT_MAX_A];
#pragma unroll
                for (int i = 0; i < AT_MAX_A; ++i) head[i] = 0.f;
                const float *w3 = net ? w3v_s : w3pi_s;
                const int outs4 = net ? 4 : A4;
                for (int c = 0; c < NC; ++c, ++chunk_n) {
                    mbar_wait(acc_full, chunk_n & 1u);
                    fence_after_sync();
                    float v[64];
                    load_chunk(v, b2, c);
                    arrive(acc_empty);  // the values are in registers: the tensor core moves on while the head is summed
                    const float *wc = w3 + AT_NCHUNK * c + 64 * hf;
#ifdef AT_DBG_NO_EPI
                    head[0] += v[7];
                    if (false)
#endif
#pragma unroll
                    for (int a0 = 0; a0 < AT_MAX_A; a0 += 4)
                        if (a0 < outs4) {
#pragma unroll
                            for (int i = 0; i < 64; i += 4) {
#pragma unroll
                                for (int t = 0; t < 4; ++t) {
                                    const float4 w = *reinterpret_cast<const float4 *>(wc + (a0 + t) * H + i);
                                    float s = head[a0 + t];
                                    s = fmaf(v[i], w.x, s), s = fmaf(v[i + 1], w.y, s), s = fmaf(v[i + 2], w.z, s), s = fmaf(v[i + 3], w.w, s);
                                    head[a0 + t] = s;
                                }
                            }
                        }
                }
                // ---- the two column halves meet; one thread per environment finishes ---------------------------------------
                if (net == 0) {
                    float *mine = comb_s + row * (AT_MAX_A + 1);
                    if (hf == 1) {
#pragma unroll
                        for (int i = 0; i < AT_MAX_A; ++i)
                            if (i < A) mine[i] = head[i];
                    }
                    workers_sync();
                    if (hf == 0) {
#pragma unroll
                        for (int i = 0; i < AT_MAX_A; ++i)
                            if (i < A) mine[i] = (head[i] + mine[i]) + b3_s[i];
                        if (valid) actor_policy_tail(a, mine, 1, grow);
                    }
                } else {
                    if (hf == 1) vcomb_s[row] = head[0];
                    workers_sync();
                    if (hf == 0 && valid && a.v) a.v[grow] = (head[0] + vcomb_s[row]) + b3_s[AT_MAX_A];
                }
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, AT_TM_COLS);
}

}  // namespace

bool actor_tc_supported(int d, int H, int A, int64_t batch)
{
    return (H == 128 || H == 256) && d >= 1 && d <= AT_MAX_DK && A >= 1 && A <= AT_MAX_A && batch >= AT_M;
}

size_t actor_tc_pack_bytes(int d, int H)
{
    if (!actor_tc_supported(d, H, 1, AT_M)) return 0;
    return (size_t)pack_shape(d, H).per_net() * AT_BLOCK;
}

int32_t actor_tc_pack_launch(const float *params, int d, int H, int out, unsigned char *pack, cudaStream_t stream, ErrBuf &err)
{
    actor_pack_kernel<<<148, 256, 0, stream>>>(params, d, H, out, reinterpret_cast<unsigned short *>(pack));
    RRL_CUDA_OK(cudaGetLastError(), err);
    return RRL_OK;
}

int32_t actor_tc_forward_launch(const ActorArgs &a, int H, cudaStream_t stream, ErrBuf &err)
{
    int dev = 0, sms = 148;
    RRL_CUDA_OK(cudaGetDevice(&dev), err);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t tiles = (a.batch + AT_M - 1) / AT_M;
    const int grid = (int)(tiles < sms ? tiles : sms);
    if (H == 256) {
        RRL_CUDA_OK(cudaFuncSetAttribute(actor_tc_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM_BYTES), err);
        actor_tc_kernel<256><<<grid, AT_THREADS, AT_SMEM_BYTES, stream>>>(a);
    } else {
        RRL_CUDA_OK(cudaFuncSetAttribute(actor_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM_BYTES), err);
        actor_tc_kernel<128><<<grid, AT_THREADS, AT_SMEM_BYTES, stream>>>(a);
    }
    RRL_CUDA_OK(cudaGetLastError(), err);
    return RRL_OK;
}

size_t actor_pack_bytes(int d, int H) { return actor_tc_pack_bytes(d, H); }
int32_t actor_pack_launch(const float *params, int d, int H, int out, unsigned char *pack, cudaStream_t stream, ErrBuf &err)
{
    return actor_tc_pack_launch(params, d, H, out, pack, stream, err);
}

}  // namespace rrl
