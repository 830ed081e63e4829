// K8: batched actor step -- PolicyWith[out]Baseline.step for many environments at once.
//
// Replaces relayrl_framework/src/native/python/algorithms/REINFORCE/kernel.py:99-107,128-135 as
// called (batch 1, libtorch CPU, under a mutex) from RelayRLAgentZmq::request_for_action
// (src/network/client/agent_zmq.rs:458-571):
//   probs, logits = policy(obs, mask); act ~ Categorical(probs); logp_a = logits[act]; v = baseline(obs)
//
// One CTA handles 32 environments end to end: obs tile -> h1 -> h2 -> logits -> softmax ->
// sample -> (value net) -> outputs.  Activations stay in shared memory; weights stream from L2
// through a double-buffered cp.async ring in 32-wide K chunks, so any hidden size that is a
// multiple of 32 up to 256 fits.  fp32 CUDA-core math (parity: rel 1e-5 vs the torch CPU oracle).
#include "actor_internal.cuh"

namespace rrl {

namespace {

constexpr int ACT_THREADS = 256;
constexpr int ACT_TM = 32;   // environments per CTA
constexpr int ACT_KC = 32;   // K chunk
constexpr int ACT_LDB = ACT_KC + 4;
constexpr int ACT_MAX_OUT = 32;

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src, int src_bytes)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem_src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// out_s[m][n] = relu?(bias[n] + sum_k A_s[m][k] * W[n][k]),  W row-major [H][ldw] in global memory
// thread (ty = tid/32, tx = tid%32): rows 4ty..4ty+3, columns c*32 + tx.
template <int H, bool RELU>
__device__ __forceinline__ void layer_stream(const float *__restrict__ A_s, int lda, const float *__restrict__ W,
                                             int ldw, int Kp, const float *__restrict__ bias,
                                             float *__restrict__ out_s, int ldo, float *__restrict__ B_s,
                                             int tid)
{
    constexpr int NC = H / 32;
    const int ty = tid >> 5, tx = tid & 31;
    float acc[4][NC];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < NC; ++c) acc[r][c] = 0.f;
    const int nchunks = (Kp + ACT_KC - 1) / ACT_KC;
    auto issue = [&](int ch) {
        float *dst = B_s + (ch & 1) * H * ACT_LDB;
        const int k0 = ch * ACT_KC;
        for (int i = tid; i < H * (ACT_KC / 4); i += ACT_THREADS) {
            const int n = i >> 3, q = i & 7;  // ACT_KC / 4 == 8 float4 per row
            const int k = k0 + 4 * q;
            const int ok = k < Kp ? 16 : 0;   // zero-fill beyond K
            cp_async16(dst + n * ACT_LDB + 4 * q, W + (size_t)n * ldw + (ok ? k : 0), ok);
        }
        cp_async_commit();
    };
    issue(0);
    for (int ch = 0; ch < nchunks; ++ch) {
        if (ch + 1 < nchunks) {
            issue(ch + 1);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const float *Bc = B_s + (ch & 1) * H * ACT_LDB;
        const int k0 = ch * ACT_KC;
        const int kn = min(ACT_KC, Kp - k0);
        for (int kk = 0; kk < kn; kk += 4) {
            float4 a[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) a[r] = *reinterpret_cast<const float4 *>(A_s + (4 * ty + r) * lda + k0 + kk);
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const float4 b = *reinterpret_cast<const float4 *>(Bc + (c * 32 + tx) * ACT_LDB + kk);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    float t = acc[r][c];
                    t = fmaf(a[r].x, b.x, t), t = fmaf(a[r].y, b.y, t), t = fmaf(a[r].z, b.z, t), t = fmaf(a[r].w, b.w, t);
                    acc[r][c] = t;
                }
            }
        }
        __syncthreads();  // the buffer just read is refilled two iterations later
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        const int n = c * 32 + tx;
        const float b = bias[n];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const float z = acc[r][c] + b;
            out_s[(4 * ty + r) * ldo + n] = RELU ? fmaxf(z, 0.f) : z;
        }
    }
    __syncthreads();
}

// head_s[m][a] = b3[a] + sum_j h_s[m][j] * W3[a][j]   (W3 read through L1; thread: row m, a = ag + 8t)
template <int H>
__device__ __forceinline__ void head_forward(const float *__restrict__ h_s, int ldh, const float *__restrict__ W3,
                                             const float *__restrict__ b3, int out, float *__restrict__ head_s,
                                             int tid)
{
    const int m = tid & 31, ag = tid >> 5;
    float hacc[4] = {0.f, 0.f, 0.f, 0.f};
    const int na = (out - ag + 7) / 8;
    if (na > 0) {
        for (int j = 0; j < H; j += 4) {
            const float4 h = *reinterpret_cast<const float4 *>(h_s + m * ldh + j);
#pragma unroll
            for (int t = 0; t < 4; ++t)
                if (t < na) {
                    const float4 w = __ldg(reinterpret_cast<const float4 *>(W3 + (size_t)(ag + 8 * t) * H + j));
                    float s = hacc[t];
                    s = fmaf(h.x, w.x, s), s = fmaf(h.y, w.y, s), s = fmaf(h.z, w.z, s), s = fmaf(h.w, w.w, s);
                    hacc[t] = s;
                }
        }
#pragma unroll
        for (int t = 0; t < 4; ++t)
            if (t < na) head_s[m * out + ag + 8 * t] = hacc[t] + b3[ag + 8 * t];
    }
    __syncthreads();
}

template <int H>
__global__ void __launch_bounds__(ACT_THREADS, 1) actor_kernel(ActorArgs a)
{
    extern __shared__ __align__(16) float smem[];
    const int d = a.d, dK = a.dK, A = a.A;
    const int ldo = dK + 4, ldh = H + 4;
    float *obs_s = smem;
    float *h1_s = obs_s + ACT_TM * ldo;
    float *h2_s = h1_s + ACT_TM * ldh;
    float *B_s = h2_s + ACT_TM * ldh;
    float *head_s = B_s + 2 * H * ACT_LDB;
    float *vhead_s = head_s + ACT_TM * ACT_MAX_OUT;
    const int tid = threadIdx.x;
    const NetDims pd{d, H, A}, vd{d, H, 1};
    const int64_t row0 = (int64_t)blockIdx.x * ACT_TM;
    const int rows = (int)min((int64_t)ACT_TM, a.batch - row0);

    for (int i = tid; i < ACT_TM * ldo; i += ACT_THREADS) {
        const int m = i / ldo, c = i - m * ldo;
        obs_s[i] = (m < rows && c < d) ? __ldcs(a.obs + (row0 + m) * d + c) : 0.f;
    }
    __syncthreads();

    // ---- policy network (kernel.py:15-31) ---------------------------------------------------------
    layer_stream<H, true>(obs_s, ldo, a.pi.w1p, dK, dK, a.pi.params + pd.ob1(), h1_s, ldh, B_s, tid);
    layer_stream<H, true>(h1_s, ldh, a.pi.params + pd.oW2(), H, H, a.pi.params + pd.ob2(), h2_s, ldh, B_s, tid);
    head_forward<H>(h2_s, ldh, a.pi.params + pd.oW3(), a.pi.params + pd.ob3(), A, head_s, tid);

    // ---- value network (kernel.py:78-84) ----------------------------------------------------------
    if (a.with_v) {
        layer_stream<H, true>(obs_s, ldo, a.vf.w1p, dK, dK, a.vf.params + vd.ob1(), h1_s, ldh, B_s, tid);
        layer_stream<H, true>(h1_s, ldh, a.vf.params + vd.oW2(), H, H, a.vf.params + vd.ob2(), h2_s, ldh, B_s, tid);
        head_forward<H>(h2_s, ldh, a.vf.params + vd.oW3(), a.vf.params + vd.ob3(), 1, vhead_s, tid);
    }

    // ---- softmax, sample, gather (kernel.py:23-37) -----------------------------------------------
    if (tid < rows) {
        const int64_t row = row0 + tid;
        actor_policy_tail(a, head_s + tid * A, 1, row);
        if (a.with_v && a.v) a.v[row] = vhead_s[tid];
    }
}

__global__ void pad_w1_kernel(const float *__restrict__ W1, int H, int d, int dK, float *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= H * dK) return;
    const int n = i / dK, k = i - n * dK;
    out[i] = k < d ? W1[n * d + k] : 0.f;
}

size_t actor_smem_bytes(int d, int H)
{
    const int dK = (d + 3) & ~3;
    return sizeof(float) * ((size_t)ACT_TM * (dK + 4) + 2 * (size_t)ACT_TM * (H + 4) + 2 * (size_t)H * ACT_LDB +
                            2 * (size_t)ACT_TM * ACT_MAX_OUT);
}

}  // namespace

int32_t actor_check_dims(int d, int H, int A, ErrBuf &err)
{
    if (H != 64 && H != 128 && H != 256) {
        err.set("actor: hidden=%d unsupported (64, 128 or 256)", H);
        return RRL_EINVAL;
    }
    if (d < 1 || A < 1 || A > ACT_MAX_OUT) {
        err.set("actor: obs_dim=%d act_dim=%d out of range (act_dim <= %d)", d, A, ACT_MAX_OUT);
        return RRL_EINVAL;
    }
    if (actor_smem_bytes(d, H) > 227 * 1024) {
        err.set("actor: obs_dim=%d hidden=%d needs %zu B of shared memory (> 227 KiB)", d, H, actor_smem_bytes(d, H));
        return RRL_EINVAL;
    }
    return RRL_OK;
}

int32_t actor_pad_w1_launch(const float *params, int H, int d, float *w1p, cudaStream_t stream, ErrBuf &err)
{
    const int dK = (d + 3) & ~3;
    pad_w1_kernel<<<(H * dK + 255) / 256, 256, 0, stream>>>(params, H, d, dK, w1p);
    RRL_CUDA_OK(cudaGetLastError(), err);
    return RRL_OK;
}

int32_t actor_forward_launch(const float *pi_params, const float *pi_w1p, const float *vf_params,
                             const float *vf_w1p, const unsigned char *pi_pack, const unsigned char *vf_pack, int d, int H, int A, const float *obs, const float *mask,
                             int64_t batch, uint64_t seed, uint64_t offset, const int64_t *force_act, int64_t *act,
                             float *logp_a, float *v, float *probs, float *logits, cudaStream_t stream, ErrBuf &err)
{
    if (batch <= 0) return RRL_OK;
    RRL_TRY(actor_check_dims(d, H, A, err));
    ActorArgs a;
    a.pi.params = pi_params, a.pi.w1p = pi_w1p, a.vf.params = vf_params, a.vf.w1p = vf_w1p;
    a.pi.pack = pi_pack, a.vf.pack = vf_pack;
    a.d = d, a.dK = (d + 3) & ~3, a.A = A, a.with_v = vf_params != nullptr;
    a.obs = obs, a.mask = mask, a.batch = batch, a.seed = seed, a.offset = offset, a.force_act = force_act;
    a.act = act, a.logp_a = logp_a, a.v = v, a.probs = probs, a.logits = logits;
    // wide layers and enough environments to fill 128-row tiles: the tensor-core kernel (actor_tc.cu)
    if (pi_pack && (!vf_params || vf_pack) && actor_tc_supported(d, H, A, batch)) return actor_tc_forward_launch(a, H, stream, err);
    const size_t smem = actor_smem_bytes(d, H);
    const int grid = (int)((batch + ACT_TM - 1) / ACT_TM);
#define RRL_ACTOR_CASE(HH)                                                                                            \
    case HH:                                                                                                          \
        RRL_CUDA_OK(cudaFuncSetAttribute(actor_kernel<HH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), err); \
        actor_kernel<HH><<<grid, ACT_THREADS, smem, stream>>>(a);                                                     \
        break;
    switch (H) {
        RRL_ACTOR_CASE(64)
        RRL_ACTOR_CASE(128)
        RRL_ACTOR_CASE(256)
        default: err.set("actor: hidden=%d unsupported", H); return RRL_EINVAL;
    }
#undef RRL_ACTOR_CASE
    RRL_CUDA_OK(cudaGetLastError(), err);
    return RRL_OK;
}

}  // namespace rrl
