// K4-K7: fused policy / value MLP forward + backward, and Adam.
//
// Replaces (paths relative to relayrl_framework/src/native/python/algorithms/REINFORCE/):
//   DiscretePolicyNetwork.forward + compute_loss_pi + autograd   (kernel.py:23-46, REINFORCE.py:141-156)
//   BaselineValueNetwork.forward + compute_loss_vf + autograd     (kernel.py:78-84,  REINFORCE.py:158-160)
//   torch.optim.Adam.step, single-tensor CPU path                  (REINFORCE.py:47,50,108,115)
//
// One persistent CTA per SM walks 64-transition tiles.  For a tile the whole chain
//   obs -> relu(W1 obs + b1) -> relu(W2 h1 + b2) -> head -> loss -> dhead -> dz2 -> dz1
// lives in shared memory / registers: nothing of size [N, H] ever touches HBM; per transition the
// kernel reads obs (4d B) plus 8-12 B of targets and writes nothing.  Weights are shared-memory
// resident for the life of the CTA.  Weight gradients accumulate in registers (dW2: a
// (H/16)x(H/16) block per thread) and thread-owned shared slots (dW1, dW3, biases) in f32 for
// FLUSH_TILES tiles, then spill into per-CTA float64 partials in global memory; a second kernel
// sums the partials over CTAs in a fixed order (deterministic, no atomics) into the flat f32
// gradient that NCCL all-reduces and Adam consumes.
//
// This file is the fp32 CUDA-core implementation of the three HxH GEMMs per tile.
#include "mlp_internal.cuh"

#include <stdlib.h>

namespace rrl {

namespace {

struct SmemLayout {
    int dpad, dp, ldh, S;
    int w2, w1, w3, b1, b2, b3, obs, h1, h2, head, mask, rowf, rowi, dw1, dw3, db3, red;
    int total_floats;
};

__host__ __device__ inline SmemLayout smem_layout(int d, int H, int out, bool grad)
{
    SmemLayout L;
    L.dpad = (d + 3) & ~3;
    L.dp = L.dpad + 4;
    L.ldh = H + 4;
    L.S = MLP_THREADS / H;
    int o = 0;
    L.w2 = o, o += H * L.ldh;
    L.w1 = o, o += H * L.dp;
    L.w3 = o, o += out * H;
    L.b1 = o, o += H;
    L.b2 = o, o += H;
    L.b3 = o, o += MAX_OUT;
    L.obs = o, o += TILE_M * L.dp;
    L.h1 = o, o += TILE_M * L.ldh;
    L.h2 = o, o += TILE_M * L.ldh;
    L.head = o, o += TILE_M * out;
    L.mask = o, o += TILE_M * out;
    L.rowf = o, o += 4 * TILE_M;  // per-row scalars: target/advn, logp_old, g, valid
    L.rowi = o, o += TILE_M;      // act
    L.dw1 = o, o += grad ? L.S * L.dpad * H : 0;
    L.dw3 = o, o += grad ? L.S * out * H : 0;
    L.db3 = o, o += MAX_OUT;
    o = (o + 1) & ~1;
    L.red = o, o += 2 * (MLP_THREADS / 32) * 4 + 2 * L.S * H;  // doubles for loss reduce + bias staging
    L.total_floats = o;
    return L;
}

// ---- register-tiled GEMMs on shared-memory operands ------------------------------------------
// thread (ty = tid/16, tx = tid%16) owns rows 4ty..4ty+3 of the 64-row tile.

// B is K-contiguous ([n][k], ld = ldb): column c of the thread is n = (c/4)*64 + (c%4)*16 + tx
template <int NG>
__device__ __forceinline__ void gemm_kcontig(const float *__restrict__ A_s, int lda,
                                             const float *__restrict__ B_s, int ldb, int K, int ty,
                                             int tx, float (&acc)[4][4 * NG])
{
    const float *a0 = A_s + (4 * ty) * lda;
#pragma unroll 2
    for (int k = 0; k < K; k += 4) {
        float4 a[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) a[r] = *reinterpret_cast<const float4 *>(a0 + r * lda + k);
#pragma unroll
        for (int c = 0; c < 4 * NG; ++c) {
            const int nn = (c >> 2) * 64 + (c & 3) * 16 + tx;
            const float4 b = *reinterpret_cast<const float4 *>(B_s + nn * ldb + k);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                float t = acc[r][c];
                t = fmaf(a[r].x, b.x, t);
                t = fmaf(a[r].y, b.y, t);
                t = fmaf(a[r].z, b.z, t);
                t = fmaf(a[r].w, b.w, t);
                acc[r][c] = t;
            }
        }
    }
}

// B is N-contiguous ([k][n], ld = ldb): column c of the thread is n = (c/4)*64 + tx*4 + c%4
template <int NG>
__device__ __forceinline__ void gemm_ncontig(const float *__restrict__ A_s, int lda,
                                             const float *__restrict__ B_s, int ldb, int K, int ty,
                                             int tx, float (&acc)[4][4 * NG])
{
    const float *a0 = A_s + (4 * ty) * lda;
#pragma unroll 1
    for (int k = 0; k < K; k += 4) {
        float4 a[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) a[r] = *reinterpret_cast<const float4 *>(a0 + r * lda + k);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
            for (int g = 0; g < NG; ++g) {
                const float4 b = *reinterpret_cast<const float4 *>(B_s + (k + kk) * ldb + g * 64 + tx * 4);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const float av = kk == 0 ? a[r].x : kk == 1 ? a[r].y : kk == 2 ? a[r].z : a[r].w;
                    acc[r][4 * g + 0] = fmaf(av, b.x, acc[r][4 * g + 0]);
                    acc[r][4 * g + 1] = fmaf(av, b.y, acc[r][4 * g + 1]);
                    acc[r][4 * g + 2] = fmaf(av, b.z, acc[r][4 * g + 2]);
                    acc[r][4 * g + 3] = fmaf(av, b.w, acc[r][4 * g + 3]);
                }
            }
        }
    }
}

// dW[(g2*64 + ty*4 + j2)][(g*64 + tx*4 + j)] += sum_m X[m][row] * Y[m][col]   (contraction over the tile)
template <int NG>
__device__ __forceinline__ void gemm_outer(const float *__restrict__ X_s, int ldx,
                                           const float *__restrict__ Y_s, int ldy, int ty, int tx,
                                           float (&acc)[4 * NG][4 * NG])
{
#pragma unroll 2
    for (int m = 0; m < TILE_M; ++m) {
        float4 x[NG], y[NG];
#pragma unroll
        for (int g = 0; g < NG; ++g) {
            x[g] = *reinterpret_cast<const float4 *>(X_s + m * ldx + g * 64 + ty * 4);
            y[g] = *reinterpret_cast<const float4 *>(Y_s + m * ldy + g * 64 + tx * 4);
        }
#pragma unroll
        for (int g2 = 0; g2 < NG; ++g2) {
            const float xv[4] = {x[g2].x, x[g2].y, x[g2].z, x[g2].w};
#pragma unroll
            for (int j2 = 0; j2 < 4; ++j2)
#pragma unroll
                for (int g = 0; g < NG; ++g) {
                    acc[4 * g2 + j2][4 * g + 0] = fmaf(xv[j2], y[g].x, acc[4 * g2 + j2][4 * g + 0]);
                    acc[4 * g2 + j2][4 * g + 1] = fmaf(xv[j2], y[g].y, acc[4 * g2 + j2][4 * g + 1]);
                    acc[4 * g2 + j2][4 * g + 2] = fmaf(xv[j2], y[g].z, acc[4 * g2 + j2][4 * g + 2]);
                    acc[4 * g2 + j2][4 * g + 3] = fmaf(xv[j2], y[g].w, acc[4 * g2 + j2][4 * g + 3]);
                }
        }
    }
}

// One pass over this CTA's tiles: the body of mlp_tile_kernel, and of every iteration of mlp_value_iters_kernel.
template <int H, int MODE, bool GRAD>
__device__ __forceinline__ void mlp_tile_body(const MlpArgs &a, float *smem)
{
    constexpr int NG = H / 64;
    constexpr int S = MLP_THREADS / H;   // row splits of the column passes
    constexpr int MS = TILE_M / S;       // rows per split
    const int d = a.d, out = a.out;
    const SmemLayout L = smem_layout(d, H, out, GRAD);
    float *W2_s = smem + L.w2, *W1_s = smem + L.w1, *W3_s = smem + L.w3;
    float *b1_s = smem + L.b1, *b2_s = smem + L.b2, *b3_s = smem + L.b3;
    float *obs_s = smem + L.obs, *h1_s = smem + L.h1, *h2_s = smem + L.h2;
    float *head_s = smem + L.head, *mask_s = smem + L.mask;
    float *tgt_s = smem + L.rowf, *lpo_s = tgt_s + TILE_M, *g_s = tgt_s + 2 * TILE_M, *valid_s = tgt_s + 3 * TILE_M;
    int *act_s = reinterpret_cast<int *>(smem + L.rowi);
    float *dW1_s = smem + L.dw1, *dW3_s = smem + L.dw3, *db3_s = smem + L.db3;
    double *red_s = reinterpret_cast<double *>(smem + L.red);
    float *bias_stage = reinterpret_cast<float *>(red_s + (MLP_THREADS / 32) * 4);

    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int dp = L.dp, dpad = L.dpad, ldh = L.ldh;
    const NetDims nd{d, H, out};

    // ---- weights -> shared memory (once per CTA) ------------------------------------------------
    {
        const float *W1 = a.params + nd.oW1(), *W2 = a.params + nd.oW2(), *W3 = a.params + nd.oW3();
        for (int i = tid; i < H * dp; i += MLP_THREADS) {
            const int rj = i / dp, ci = i - rj * dp;
            W1_s[i] = ci < d ? __ldcg(&W1[rj * d + ci]) : 0.f;
        }
        for (int i = tid; i < H * (H / 4); i += MLP_THREADS) {
            const int rj = i / (H / 4), c4 = i - rj * (H / 4);
            *reinterpret_cast<float4 *>(&W2_s[rj * ldh + 4 * c4]) = __ldcg(reinterpret_cast<const float4 *>(&W2[rj * H + 4 * c4]));
        }
        for (int i = tid; i < out * H; i += MLP_THREADS) W3_s[i] = __ldcg(&W3[i]);
        for (int i = tid; i < H; i += MLP_THREADS) b1_s[i] = __ldcg(&a.params[nd.ob1() + i]), b2_s[i] = __ldcg(&a.params[nd.ob2() + i]);
        if (tid < MAX_OUT) b3_s[tid] = tid < out ? __ldcg(&a.params[nd.ob3() + tid]) : 0.f, db3_s[tid] = 0.f;
        if (GRAD) {
            for (int i = tid; i < S * dpad * H; i += MLP_THREADS) dW1_s[i] = 0.f;
            for (int i = tid; i < S * out * H; i += MLP_THREADS) dW3_s[i] = 0.f;
        }
    }
    float mean = 0.f, sd = 1.f;
    if (MODE != MODE_VALUE) mean = (float)a.stats[3], sd = (float)a.stats[4];

    float dW2acc[GRAD ? 4 * NG : 1][GRAD ? 4 * NG : 1];
#pragma unroll
    for (int i = 0; i < (GRAD ? 4 * NG : 1); ++i)
#pragma unroll
        for (int j = 0; j < (GRAD ? 4 * NG : 1); ++j) dW2acc[i][j] = 0.f;
    float db1acc = 0.f, db2acc = 0.f;
    double l0 = 0.0, l1 = 0.0, l2 = 0.0;  // loss sums of rows owned by this thread (tid < 64)
    bool first_flush = true;
    int since_flush = 0;
    double *my_partials = a.partials ? a.partials + (size_t)blockIdx.x * a.n_params : nullptr;

    const int64_t num_tiles = (a.n + TILE_M - 1) / TILE_M;
    __syncthreads();

    for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int64_t row0 = tile * TILE_M;
        const int rows = (int)min((int64_t)TILE_M, a.n - row0);

        // ---- P0: tile inputs ---------------------------------------------------------------------
        for (int i = tid; i < TILE_M * dp; i += MLP_THREADS) {
            const int m = i / dp, c = i - m * dp;
            obs_s[i] = (m < rows && c < d) ? __ldcs(a.obs + (row0 + m) * d + c) : 0.f;
        }
        if (MODE != MODE_VALUE)
            for (int i = tid; i < TILE_M * out; i += MLP_THREADS)
                mask_s[i] = (i < rows * out) ? __ldcs(a.mask + row0 * out + i) : 1.f;
        if (tid < TILE_M) {
            const bool ok = tid < rows;
            valid_s[tid] = ok ? 1.f : 0.f;
            if (MODE == MODE_VALUE) {
                tgt_s[tid] = ok ? a.ret[row0 + tid] : 0.f;
                act_s[tid] = 0;
            } else {
                // advantage normalisation folded into the read (replay_buffer.py:105; f32, IEEE division)
                tgt_s[tid] = ok ? __fdiv_rn(a.adv[row0 + tid] - mean, sd) : 0.f;
                lpo_s[tid] = ok ? a.logp_old[row0 + tid] : 0.f;
                int at = ok ? (int)a.act[row0 + tid] : 0;  // act.long() (kernel.py:34)
                act_s[tid] = min(max(at, 0), out - 1);
            }
        }
        __syncthreads();

        // ---- P1: h1 = relu(obs W1^T + b1) ------------------------------------------------------------
        {
            float acc[4][4 * NG];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4 * NG; ++c) acc[r][c] = 0.f;
            gemm_kcontig<NG>(obs_s, dp, W1_s, dp, dpad, ty, tx, acc);
#pragma unroll
            for (int c = 0; c < 4 * NG; ++c) {
                const int nn = (c >> 2) * 64 + (c & 3) * 16 + tx;
                const float b = b1_s[nn];
#pragma unroll
                for (int r = 0; r < 4; ++r) h1_s[(4 * ty + r) * ldh + nn] = fmaxf(acc[r][c] + b, 0.f);
            }
        }
        __syncthreads();

        // ---- P2: h2 = relu(h1 W2^T + b2) -------------------------------------------------------------
        {
            float acc[4][4 * NG];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4 * NG; ++c) acc[r][c] = 0.f;
            gemm_kcontig<NG>(h1_s, ldh, W2_s, ldh, H, ty, tx, acc);
#pragma unroll
            for (int c = 0; c < 4 * NG; ++c) {
                const int nn = (c >> 2) * 64 + (c & 3) * 16 + tx;
                const float b = b2_s[nn];
#pragma unroll
                for (int r = 0; r < 4; ++r) h2_s[(4 * ty + r) * ldh + nn] = fmaxf(acc[r][c] + b, 0.f);
            }
        }
        __syncthreads();

        // ---- P3: head = h2 W3^T + b3  (thread: row m, outputs a = ag, ag+4, ...) ------------------------
        {
            const int m = tid & (TILE_M - 1), ag = tid >> 6;
            float hacc[HEAD_ACC];
#pragma unroll
            for (int t = 0; t < HEAD_ACC; ++t) hacc[t] = 0.f;
            const int na = (out - ag + 3) / 4;  // number of outputs this thread owns
            if (na > 0) {
                for (int j = 0; j < H; j += 4) {
                    const float4 h = *reinterpret_cast<const float4 *>(&h2_s[m * ldh + j]);
#pragma unroll
                    for (int t = 0; t < HEAD_ACC; ++t)
                        if (t < na) {
                            const float4 w = *reinterpret_cast<const float4 *>(&W3_s[(ag + 4 * t) * H + j]);
                            float s = hacc[t];
                            s = fmaf(h.x, w.x, s), s = fmaf(h.y, w.y, s), s = fmaf(h.z, w.z, s), s = fmaf(h.w, w.w, s);
                            hacc[t] = s;
                        }
                }
#pragma unroll
                for (int t = 0; t < HEAD_ACC; ++t)
                    if (t < na) head_s[m * out + ag + 4 * t] = hacc[t] + b3_s[ag + 4 * t];
            }
        }
        __syncthreads();

        // ---- P4: per-row loss and dhead ---------------------------------------------------------------
        if (tid < TILE_M) {
            const int m = tid;
            const bool ok = valid_s[m] != 0.f;
            if (MODE == MODE_VALUE) {
                const float v = head_s[m];
                if (a.out_rows && ok) a.out_rows[row0 + m] = v;
                const float diff = v - tgt_s[m];
                if (ok) l0 += (double)diff * (double)diff;
                // mean backward (1/N), then pow backward 2*diff (REINFORCE.py:160)
                g_s[m] = ok ? a.inv_n * (2.f * diff) : 0.f;
            } else {
                // logits = x + (mask-1)*1e8 ; probs = softmax(logits)   (kernel.py:23-28)
                float mx = -INFINITY;
                for (int q = 0; q < out; ++q) {
                    const float lg = head_s[m * out + q] + (mask_s[m * out + q] - 1.f) * 1e8f;
                    head_s[m * out + q] = lg;
                    mx = fmaxf(mx, lg);
                }
                float se = 0.f;
                for (int q = 0; q < out; ++q) se += expf(head_s[m * out + q] - mx);
                const int at = act_s[m];
                const float lg_a = head_s[m * out + at];
                float ent = 0.f;
                for (int q = 0; q < out; ++q) {
                    const float lg = head_s[m * out + q];
                    const float p = expf(lg - mx) / se;
                    ent += fmaxf(lg, -3.402823466e38f) * p;  // clamp(min=finfo.min) (REINFORCE.py:150-151)
                    if (a.out_rows && ok) a.out_rows[(row0 + m) * out + q] = lg;
                    if (MODE == MODE_POLICY_LSM) mask_s[m * out + q] = p;  // probs for the dense dhead
                }
                // reference: logp is the raw masked logit (kernel.py:33-37)
                const float logp = MODE == MODE_POLICY_REF ? lg_a : (lg_a - mx) - logf(se);
                const float advn = tgt_s[m];
                if (ok) {
                    l0 += (double)(logp * advn);
                    l1 += (double)(lpo_s[m] - logp);
                    l2 += (double)ent;
                }
                // loss = -(logp*adv).mean(): d/dlogp = -(1/N) * adv
                g_s[m] = ok ? (-a.inv_n) * advn : 0.f;
            }
        }
        __syncthreads();

        if constexpr (GRAD) {
            // ---- P5: dW3/db3, dz2 = (dhead W3) * [h2 > 0], db2  (thread: column j, row split s) -------
            {
                const int j = tid % H, s = tid / H;
                float *dW3_mine = dW3_s + s * out * H;
                float dbl = 0.f;
                for (int m = s * MS; m < (s + 1) * MS; ++m) {
                    const float h = h2_s[m * ldh + j];
                    const float g = g_s[m];
                    float dh;
                    if (MODE != MODE_POLICY_LSM) {
                        const int at = act_s[m];
                        dW3_mine[at * H + j] = fmaf(g, h, dW3_mine[at * H + j]);
                        dh = g * W3_s[at * H + j];
                    } else {
                        // log-softmax head: dlogit[q] = g * ([q == act] - p[q])
                        const int at = act_s[m];
                        dh = 0.f;
                        for (int q = 0; q < out; ++q) {
                            const float dl = g * ((q == at ? 1.f : 0.f) - mask_s[m * out + q]);
                            dW3_mine[q * H + j] = fmaf(dl, h, dW3_mine[q * H + j]);
                            dh = fmaf(dl, W3_s[q * H + j], dh);
                        }
                    }
                    const float dz = h > 0.f ? dh : 0.f;
                    h2_s[m * ldh + j] = dz;
                    dbl += dz;
                }
                db2acc += dbl;
                if (tid < out) {
                    float sb = 0.f;
                    for (int m = 0; m < TILE_M; ++m) {
                        const float g = g_s[m];
                        if (MODE != MODE_POLICY_LSM)
                            sb += (act_s[m] == tid) ? g : 0.f;
                        else
                            sb += g * ((act_s[m] == tid ? 1.f : 0.f) - mask_s[m * out + tid]);
                    }
                    db3_s[tid] += sb;
                }
            }
            __syncthreads();

            // ---- P6: dW2 += dz2^T h1 ---------------------------------------------------------------------
            gemm_outer<NG>(h2_s, ldh, h1_s, ldh, ty, tx, dW2acc);
            __syncthreads();

            // ---- P7: dz1 = (dz2 W2) * [h1 > 0]  (overwrites h1) ---------------------------------------------
            {
                float acc[4][4 * NG];
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int c = 0; c < 4 * NG; ++c) acc[r][c] = 0.f;
                gemm_ncontig<NG>(h2_s, ldh, W2_s, ldh, H, ty, tx, acc);
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int g = 0; g < NG; ++g) {
                        float4 *p = reinterpret_cast<float4 *>(&h1_s[(4 * ty + r) * ldh + g * 64 + tx * 4]);
                        const float4 hv = *p;
                        float4 o;
                        o.x = hv.x > 0.f ? acc[r][4 * g + 0] : 0.f;
                        o.y = hv.y > 0.f ? acc[r][4 * g + 1] : 0.f;
                        o.z = hv.z > 0.f ? acc[r][4 * g + 2] : 0.f;
                        o.w = hv.w > 0.f ? acc[r][4 * g + 3] : 0.f;
                        *p = o;
                    }
            }
            __syncthreads();

            // ---- P8: dW1 += dz1^T obs, db1 -----------------------------------------------------------------
            {
                const int j = tid % H, s = tid / H;
                float *dW1_mine = dW1_s + s * dpad * H;
                float dbl = 0.f;
                for (int m = s * MS; m < (s + 1) * MS; ++m) {
                    const float dz = h1_s[m * ldh + j];
                    dbl += dz;
                    for (int i = 0; i < d; ++i) dW1_mine[i * H + j] = fmaf(dz, obs_s[m * dp + i], dW1_mine[i * H + j]);
                }
                db1acc += dbl;
            }
            __syncthreads();

            // ---- spill the f32 window into the CTA's float64 partials -----------------------------------------
            ++since_flush;
            const bool last = tile + gridDim.x >= num_tiles;
            if (since_flush == FLUSH_TILES || last) {
                since_flush = 0;
                // biases: S partial sums per column -> staging
                bias_stage[(tid / H) * H + (tid % H)] = db1acc;
                bias_stage[S * H + (tid / H) * H + (tid % H)] = db2acc;
                db1acc = 0.f, db2acc = 0.f;
                __syncthreads();
                auto put = [&](int64_t p, float v) {
                    if (first_flush)
                        my_partials[p] = (double)v;
                    else
                        my_partials[p] += (double)v;
                };
                // W1 [H][d], b1
                for (int i = tid; i < H * d; i += MLP_THREADS) {
                    const int jj = i / d, ii = i - jj * d;
                    float v = 0.f;
                    for (int s = 0; s < S; ++s) v += dW1_s[s * dpad * H + ii * H + jj], dW1_s[s * dpad * H + ii * H + jj] = 0.f;
                    put(nd.oW1() + i, v);
                }
                if (tid < H) {
                    float v1 = 0.f, v2 = 0.f;
                    for (int s = 0; s < S; ++s) v1 += bias_stage[s * H + tid], v2 += bias_stage[S * H + s * H + tid];
                    put(nd.ob1() + tid, v1);
                    put(nd.ob2() + tid, v2);
                }
                // W2 [H][H] from registers
#pragma unroll
                for (int g2 = 0; g2 < NG; ++g2)
#pragma unroll
                    for (int j2 = 0; j2 < 4; ++j2)
#pragma unroll
                        for (int g = 0; g < NG; ++g)
#pragma unroll
                            for (int jx = 0; jx < 4; ++jx) {
                                const int o_ = g2 * 64 + ty * 4 + j2, i_ = g * 64 + tx * 4 + jx;
                                put(nd.oW2() + (int64_t)o_ * H + i_, dW2acc[4 * g2 + j2][4 * g + jx]);
                                dW2acc[4 * g2 + j2][4 * g + jx] = 0.f;
                            }
                // W3 [out][H], b3
                for (int i = tid; i < out * H; i += MLP_THREADS) {
                    float v = 0.f;
                    for (int s = 0; s < S; ++s) v += dW3_s[s * out * H + i], dW3_s[s * out * H + i] = 0.f;
                    put(nd.oW3() + i, v);
                }
                if (tid < out) put(nd.ob3() + tid, db3_s[tid]), db3_s[tid] = 0.f;
                first_flush = false;
                __syncthreads();
            }
        }
    }

    // ---- loss sums: rows live in threads 0..63 (warps 0,1) ------------------------------------------
    {
#pragma unroll
        for (int dd = 16; dd > 0; dd >>= 1) {
            l0 += __shfl_xor_sync(0xffffffffu, l0, dd);
            l1 += __shfl_xor_sync(0xffffffffu, l1, dd);
            l2 += __shfl_xor_sync(0xffffffffu, l2, dd);
        }
        if ((tid & 31) == 0 && tid < TILE_M) red_s[(tid >> 5) * 4 + 0] = l0, red_s[(tid >> 5) * 4 + 1] = l1, red_s[(tid >> 5) * 4 + 2] = l2;
        __syncthreads();
        if (tid == 0 && a.loss_partials) {
            double *lp = a.loss_partials + (size_t)blockIdx.x * 4;
            lp[0] = red_s[0] + red_s[4], lp[1] = red_s[1] + red_s[5], lp[2] = red_s[2] + red_s[6], lp[3] = 0.0;
        }
    }
}

template <int H, int MODE, bool GRAD>
__global__ void __launch_bounds__(MLP_THREADS, 1) mlp_tile_kernel(MlpArgs a)
{
    extern __shared__ __align__(16) float smem[];
    mlp_tile_body<H, MODE, GRAD>(a, smem);
}

// ---- small batches: all value iterations of an epoch in ONE launch ------------------------------------------------------
// REINFORCE.train_model runs train_vf_iters (80) Adam steps on the value net per epoch (REINFORCE.py:110-115).  On a
// CartPole-sized batch (config C1: ~2 000 transitions = 32 tiles) each step is three latency-bound launches -- pass,
// reduce, Adam: ~64 us, of which the arithmetic is ~12.  Here the CTAs stay resident (cooperative launch, one tile each)
// and iterate: pass over the own tile (mlp_tile_body, weights re-read through L2) -> grid barrier -> every thread
// reduces its share of the parameters over the CTAs' float64 partials in the same fixed order as grad_reduce_kernel
// and applies adam_kernel's operations to them -> grid barrier.  Same arithmetic in the same order as the three-launch
// path: the weights come out bit-identical (tests/test_gpu_learner.py::test_fused_value_iterations_bit_identical).
constexpr int FUSED_MAX_ITERS = 256;
struct FusedArgs {
    int iters, used;             // Adam steps; CTAs that own a tile (= the grid)
    float *params, *grad, *m, *v;
    unsigned *barrier;           // zeroed before the launch
    double *loss_first, *loss_last;  // sum of squared errors before the first / before the last step
    float w1, b2, w2, eps;       // 1-beta1, beta2, 1-beta2, eps
    float bc2_sqrt[FUSED_MAX_ITERS], neg_step[FUSED_MAX_ITERS];  // per step, formed on the host exactly as adam_launch does
};

__device__ __forceinline__ void grid_barrier(unsigned *ctr, unsigned target)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(ctr, 1u);
        unsigned seen;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(ctr) : "memory");
        } while (seen < target);
        __threadfence();
    }
    __syncthreads();
}

template <int H>
__global__ void __launch_bounds__(MLP_THREADS, 1) mlp_value_iters_kernel(MlpArgs a, const __grid_constant__ FusedArgs f)
{
    extern __shared__ __align__(16) float smem[];
    const int64_t P = a.n_params;
    const int64_t gtid = (int64_t)blockIdx.x * MLP_THREADS + threadIdx.x, gthreads = (int64_t)gridDim.x * MLP_THREADS;
    unsigned phase = 0;
    for (int it = 0; it < f.iters; ++it) {
        mlp_tile_body<H, MODE_VALUE, true>(a, smem);
        grid_barrier(f.barrier, ++phase * gridDim.x);
        // grad_reduce_kernel + adam_kernel on this thread's parameters (partials and weights through L2: other SMs wrote them)
        for (int64_t p = gtid; p < P; p += gthreads) {
            double s = 0.0;
            for (int c = 0; c < f.used; ++c) s += __ldcg(&a.partials[(size_t)c * P + p]);
            const float gi = (float)s;
            float mi = f.m[p], vi = f.v[p];
            mi = __fadd_rn(mi, __fmul_rn(f.w1, __fsub_rn(gi, mi)));
            vi = __fadd_rn(__fmul_rn(vi, f.b2), __fmul_rn(__fmul_rn(f.w2, gi), gi));
            const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), f.bc2_sqrt[it]), f.eps);
            __stcg(&f.params[p], __fadd_rn(__ldcg(&f.params[p]), __fdiv_rn(__fmul_rn(f.neg_step[it], mi), denom)));
            f.m[p] = mi, f.v[p] = vi, f.grad[p] = gi;
        }
        if (gtid == 0 && (it == 0 || it == f.iters - 1)) {
            double s = 0.0;
            for (int c = 0; c < f.used; ++c) s += __ldcg(&a.loss_partials[(size_t)c * 4]);
            if (it == 0) *f.loss_first = s;
            if (it == f.iters - 1) *f.loss_last = s;
        }
        grid_barrier(f.barrier, ++phase * gridDim.x);
    }
}

// grad[p] = sum over CTAs of the float64 partials, fixed order; losses likewise.
__global__ void __launch_bounds__(256) grad_reduce_kernel(const double *__restrict__ partials, int grid_used,
                                                         int64_t n_params, float *__restrict__ grad,
                                                         double *__restrict__ grad64,
                                                         const double *__restrict__ loss_partials,
                                                         double *__restrict__ loss_out)
{
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (partials && p < n_params) {
        double s = 0.0;
        for (int c = 0; c < grid_used; ++c) s += partials[(size_t)c * n_params + p];
        grad[p] = (float)s;
        if (grad64) grad64[p] = s;
    }
    if (loss_partials && blockIdx.x == 0 && threadIdx.x < 4) {
        double s = 0.0;
        for (int c = 0; c < grid_used; ++c) s += loss_partials[(size_t)c * 4 + threadIdx.x];
        loss_out[threadIdx.x] = s;
    }
}

__global__ void __launch_bounds__(256) grad_round_kernel(const double *__restrict__ g64, float *__restrict__ g, int64_t n)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) g[i] = (float)g64[i];
}

// torch.optim.Adam, single-tensor CPU path, f32 with unfused operations:
//   m.lerp_(g, 1-b1); v.mul_(b2).addcmul_(g, g, 1-b2);
//   denom = v.sqrt() / sqrt(1-b2^t) + eps;  p.addcdiv_(m, denom, value = -lr/(1-b1^t))
__global__ void __launch_bounds__(256) adam_kernel(float *__restrict__ p, const float *__restrict__ g,
                                                  float *__restrict__ m, float *__restrict__ v, int64_t n,
                                                  float w1 /*1-b1*/, float b2, float w2 /*1-b2*/,
                                                  float bc2_sqrt, float eps, float neg_step_size)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float gi = g[i];
    float mi = m[i], vi = v[i];
    mi = __fadd_rn(mi, __fmul_rn(w1, __fsub_rn(gi, mi)));           // lerp, weight < 0.5
    vi = __fadd_rn(__fmul_rn(vi, b2), __fmul_rn(__fmul_rn(w2, gi), gi));
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), bc2_sqrt), eps);
    p[i] = __fadd_rn(p[i], __fdiv_rn(__fmul_rn(neg_step_size, mi), denom));
    m[i] = mi, v[i] = vi;
}

template <int H, int MODE>
int32_t launch_mode(NetState &net, const MlpArgs &a, bool grad, int grid, size_t smem, cudaStream_t stream, ErrBuf &err)
{
    if (grad) {
        RRL_CUDA_OK(cudaFuncSetAttribute(mlp_tile_kernel<H, MODE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), err);
        mlp_tile_kernel<H, MODE, true><<<grid, MLP_THREADS, smem, stream>>>(a);
    } else {
        RRL_CUDA_OK(cudaFuncSetAttribute(mlp_tile_kernel<H, MODE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), err);
        mlp_tile_kernel<H, MODE, false><<<grid, MLP_THREADS, smem, stream>>>(a);
    }
    RRL_CUDA_OK(cudaGetLastError(), err);
    return RRL_OK;
}

template <int MODE>
int32_t launch_h(NetState &net, const MlpArgs &a, bool grad, int grid, size_t smem, cudaStream_t stream, ErrBuf &err)
{
    switch (net.dims.H) {
        case 64: return launch_mode<64, MODE>(net, a, grad, grid, smem, stream, err);
        case 128: return launch_mode<128, MODE>(net, a, grad, grid, smem, stream, err);
        default: err.set("hidden=%d unsupported by the fused learner kernel (64 or 128)", net.dims.H); return RRL_EINVAL;
    }
}

int32_t run_pass(NetState &net, const TrainBuffers &tb, int mode, bool with_grad, float *out_rows,
                 cudaStream_t stream, ErrBuf &err, int *launches)
{
    const NetDims &nd = net.dims;
    RRL_TRY(mlp_check_dims(nd.d, nd.H, nd.out, err));
    MlpArgs a;
    a.params = net.params, a.d = nd.d, a.out = nd.out;
    a.obs = tb.obs, a.act = tb.act, a.mask = tb.mask, a.adv = tb.adv, a.ret = tb.ret, a.logp_old = tb.logp_old;
    a.stats = tb.stats, a.n = tb.n;
    // torch forms 1/N in f32 inside mean's backward
    a.inv_n = 1.0f / (float)tb.n_global;
    a.partials = with_grad ? net.partials64 : nullptr;
    a.loss_partials = net.loss_partials, a.n_params = nd.n_params(), a.out_rows = out_rows;
    const int64_t tiles = (tb.n + TILE_M - 1) / TILE_M;
    int grid = (int)std::min<int64_t>(net.grid, std::max<int64_t>(tiles, 1));
    const size_t smem = mlp_smem_bytes(nd.d, nd.H, nd.out);
    int32_t st;
#ifdef RRL_DEVTOOLS  // developer library only: force the CUDA-core kernel for A/B comparisons
    static const bool tc_off = getenv("RRL_DISABLE_TC") != nullptr;
#else
    constexpr bool tc_off = false;
#endif
    if (!tc_off && tc_pass_supported(nd, mode, with_grad, tb.tc_any_size ? (int64_t)1 << 40 : tb.n))
        st = tc_pass_launch(a, nd, mode, grid, stream, err);
    else if (mode == MODE_VALUE)
        st = launch_h<MODE_VALUE>(net, a, with_grad, grid, smem, stream, err);
    else if (mode == MODE_POLICY_REF)
        st = launch_h<MODE_POLICY_REF>(net, a, with_grad, grid, smem, stream, err);
    else
        st = launch_h<MODE_POLICY_LSM>(net, a, with_grad, grid, smem, stream, err);
    RRL_TRY(st);
    // CTAs that received no tile leave their partials untouched: only the first min(grid, tiles) count
    const int used = (int)std::min<int64_t>(grid, tiles);
    const int64_t P = nd.n_params();
    grad_reduce_kernel<<<(int)((P + 255) / 256), 256, 0, stream>>>(with_grad ? net.partials64 : nullptr, used, P,
                                                                    net.grad, with_grad ? net.grad64 : nullptr,
                                                                    net.loss_partials, net.loss_out);
    RRL_CUDA_OK(cudaGetLastError(), err);
    if (launches) *launches += 2;
    return RRL_OK;
}

}  // namespace

size_t mlp_smem_bytes(int d, int H, int out) { return sizeof(float) * (size_t)smem_layout(d, H, out, true).total_floats; }

int32_t mlp_check_dims(int d, int H, int out, ErrBuf &err)
{
    if (H != 64 && H != 128) {
        err.set("hidden=%d unsupported by the fused learner kernel (64 or 128)", H);
        return RRL_EINVAL;
    }
    if (d < 1 || out < 1 || out > MAX_OUT) {
        err.set("obs_dim=%d / act_dim=%d outside the fused learner kernel's range (act_dim <= %d)", d, out, MAX_OUT);
        return RRL_EINVAL;
    }
    if (mlp_smem_bytes(d, H, out) > 227 * 1024) {
        err.set("obs_dim=%d act_dim=%d hidden=%d needs %zu B of shared memory per CTA (> 227 KiB)", d, out, H,
                mlp_smem_bytes(d, H, out));
        return RRL_EINVAL;
    }
    return RRL_OK;
}

int mlp_grid_for_device(int device, int H)
{
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    (void)H;
    return sms;  // one persistent CTA per SM (shared-memory resident weights)
}

int32_t policy_pass_launch(NetState &net, const TrainBuffers &tb, int logp_mode, bool with_grad, float *logits_out,
                           cudaStream_t stream, ErrBuf &err, int *launches)
{
    return run_pass(net, tb, logp_mode == RRL_LOGP_LOG_SOFTMAX ? MODE_POLICY_LSM : MODE_POLICY_REF, with_grad,
                    logits_out, stream, err, launches);
}

int32_t value_pass_launch(NetState &net, const TrainBuffers &tb, bool with_grad, float *v_out, cudaStream_t stream,
                          ErrBuf &err, int *launches)
{
    return run_pass(net, tb, MODE_VALUE, with_grad, v_out, stream, err, launches);
}

bool value_iters_fused_supported(const NetState &net, const TrainBuffers &tb, int iters, int nranks)
{
    const NetDims &nd = net.dims;
    const int64_t tiles = (tb.n + TILE_M - 1) / TILE_M;
    return nranks <= 1 && iters >= 1 && iters <= FUSED_MAX_ITERS && (nd.H == 64 || nd.H == 128) && tiles >= 1 && tiles <= net.grid &&
           net.fused_barrier != nullptr && !tc_pass_supported(nd, MODE_VALUE, true, tb.tc_any_size ? (int64_t)1 << 40 : tb.n);
}

// train_vf_iters value steps in one cooperative launch (see mlp_value_iters_kernel); loss_first / loss_last: device doubles
int32_t value_iters_fused_launch(NetState &net, const TrainBuffers &tb, int iters, double lr, double b1, double b2, double eps,
                                 double *loss_first, double *loss_last, cudaStream_t stream, ErrBuf &err, int *launches)
{
    const NetDims &nd = net.dims;
    RRL_TRY(mlp_check_dims(nd.d, nd.H, nd.out, err));
    MlpArgs a;
    a.params = net.params, a.d = nd.d, a.out = nd.out;
    a.obs = tb.obs, a.act = tb.act, a.mask = tb.mask, a.adv = tb.adv, a.ret = tb.ret, a.logp_old = tb.logp_old;
    a.stats = tb.stats, a.n = tb.n;
    a.inv_n = 1.0f / (float)tb.n_global;
    a.partials = net.partials64, a.loss_partials = net.loss_partials, a.n_params = nd.n_params(), a.out_rows = nullptr;
    FusedArgs f;
    f.iters = iters, f.used = (int)((tb.n + TILE_M - 1) / TILE_M);
    f.params = net.params, f.grad = net.grad, f.m = net.m, f.v = net.v, f.barrier = net.fused_barrier;
    f.loss_first = loss_first, f.loss_last = loss_last;
    f.w1 = (float)(1.0 - b1), f.b2 = (float)b2, f.w2 = (float)(1.0 - b2), f.eps = (float)eps;
    for (int i = 0; i < iters; ++i) {  // exactly adam_launch's scalars for steps net.step + 1 .. net.step + iters
        const double t = (double)(net.step + 1 + i);
        const double bc1 = 1.0 - pow(b1, t), bc2 = 1.0 - pow(b2, t);
        f.bc2_sqrt[i] = (float)sqrt(bc2), f.neg_step[i] = (float)(-(lr / bc1));
    }
    const size_t smem = mlp_smem_bytes(nd.d, nd.H, nd.out);
    RRL_CUDA_OK(cudaMemsetAsync(net.fused_barrier, 0, sizeof(unsigned), stream), err);
    void *args[2] = {&a, &f};
    const void *fn = nd.H == 64 ? (const void *)mlp_value_iters_kernel<64> : (const void *)mlp_value_iters_kernel<128>;
    RRL_CUDA_OK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), err);
    RRL_CUDA_OK(cudaLaunchCooperativeKernel(fn, dim3((unsigned)f.used), dim3(MLP_THREADS), args, smem, stream), err);
    net.step += iters;
    if (launches) *launches += 1;
    return RRL_OK;
}

int32_t grad_round_launch(NetState &net, cudaStream_t stream, ErrBuf &err, int *launches)
{
    const int64_t P = net.dims.n_params();
    grad_round_kernel<<<(int)((P + 255) / 256), 256, 0, stream>>>(net.grad64, net.grad, P);
    RRL_CUDA_OK(cudaGetLastError(), err);
    if (launches) *launches += 1;
    return RRL_OK;
}

int32_t adam_launch(NetState &net, double lr, double b1, double b2, double eps, cudaStream_t stream, ErrBuf &err,
                    int *launches)
{
    net.step += 1;
    const double t = (double)net.step;
    const double bc1 = 1.0 - pow(b1, t), bc2 = 1.0 - pow(b2, t);
    const double step_size = lr / bc1;
    const int64_t P = net.dims.n_params();
    adam_kernel<<<(int)((P + 255) / 256), 256, 0, stream>>>(net.params, net.grad, net.m, net.v, P, (float)(1.0 - b1),
                                                           (float)b2, (float)(1.0 - b2), (float)sqrt(bc2), (float)eps,
                                                           (float)(-step_size));
    RRL_CUDA_OK(cudaGetLastError(), err);
    if (launches) *launches += 1;
    return RRL_OK;
}

}  // namespace rrl
