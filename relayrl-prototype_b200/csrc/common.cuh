// Shared declarations of librelayrl_b200 (sm_100a).  Host + device helpers only; the C ABI
// lives in include/relayrl_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <algorithm>
#include <string>

#include "../../include/relayrl_b200.h"

namespace rrl {

// Thrown-free error plumbing: every internal function returns a status and writes a message.
struct ErrBuf {
    std::string msg;
    void set(const char *fmt, ...);
};

#define RRL_CUDA_OK(expr, err)                                                                   \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess) {                                                                 \
            (err).set("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return RRL_ECUDA;                                                                    \
        }                                                                                        \
    } while (0)

#define RRL_TRY(expr)                 \
    do {                              \
        int32_t _s = (expr);          \
        if (_s != RRL_OK) return _s;  \
    } while (0)

// adv statistics record written by the scan (device, 5 doubles)
struct AdvStats {
    double sum, sumsq, n, mean, std;  // mean/std hold float32 values widened to double
};

// ---- scan.cu ------------------------------------------------------------------------------
size_t scan_workspace_bytes(int64_t n);
int32_t scan_launch(const float *rew, const float *val, const int64_t *offsets,
                    const float *last_vals, int64_t n_total, int64_t n_ep, double gamma, double lam,
                    bool with_baseline, float *adv, float *ret, double *stats_out, void *workspace,
                    size_t workspace_bytes, bool finalize_stats, int64_t n_scan_known /* -1: read offsets[n_ep] back */,
                    cudaStream_t stream, ErrBuf &err, int *launches);
// mean/std from (sum, sumsq, n) already in stats[0..2] (used after a cross-rank merge)
int32_t stats_finalize_launch(double *stats, cudaStream_t stream, ErrBuf &err);
int32_t stats_merge_launch(const double *gathered, int nranks, double *stats, cudaStream_t stream, ErrBuf &err);
int32_t normalise_launch(const float *adv, int64_t n, const double *stats, float *out,
                         cudaStream_t stream, ErrBuf &err);
int32_t cvt_f64_f32_launch(const double *src, float *dst, int64_t n, cudaStream_t stream, ErrBuf &err);

// ---- mlp.cu -------------------------------------------------------------------------------
struct NetDims {
    int d, H, out;  // obs_dim, hidden, output width (A for the policy, 1 for the value net)
    __host__ __device__ int64_t n_params() const { return (int64_t)H * d + H + (int64_t)H * H + H + (int64_t)out * H + out; }
    __host__ __device__ int64_t oW1() const { return 0; }
    __host__ __device__ int64_t ob1() const { return (int64_t)H * d; }
    __host__ __device__ int64_t oW2() const { return ob1() + H; }
    __host__ __device__ int64_t ob2() const { return oW2() + (int64_t)H * H; }
    __host__ __device__ int64_t oW3() const { return ob2() + H; }
    __host__ __device__ int64_t ob3() const { return oW3() + (int64_t)out * H; }
};

struct TrainBuffers {
    const float *obs, *act, *mask, *adv, *ret, *logp_old;  // device SoA
    const double *stats;                                    // AdvStats (mean/std) for on-the-fly normalisation
    int64_t n;                                              // local transitions
    int64_t n_global;                                       // N over all ranks (the mean's denominator)
    bool tc_any_size;                                       // RRL_FLAG_TC_ANY_SIZE
};

struct NetState {
    NetDims dims;
    float *params;    // flat, torch order
    float *grad;      // flat gradient (after the CTA-partial reduction)
    unsigned *fused_barrier;  // grid barrier counter of mlp_value_iters_kernel (small batches, single rank)
    double *grad64;   // data parallel only: this rank's float64 gradient sum, all-reduced in float64 and rounded
                      // to f32 once afterwards, so that N ranks and one rank differ by f64 re-association only
    float *m, *v;     // Adam moments
    double *partials64;     // [grid][n_params] per-CTA float64 gradient partials
    double *loss_partials;  // [grid][4]
    double *loss_out;       // [4] reduced: loss, kl, entropy, spare
    int64_t step;
    int grid;
};

int mlp_grid_for_device(int device, int H);
size_t mlp_smem_bytes(int d, int H, int out);
int32_t mlp_check_dims(int d, int H, int out, ErrBuf &err);
// policy: forward (+ backward if with_grad) over all transitions, gradient partials -> grad, losses -> loss_out
// logits_out ([n][out]) / v_out ([n]) are optional per-row forward taps (device), may be NULL
int32_t policy_pass_launch(NetState &net, const TrainBuffers &tb, int logp_mode, bool with_grad,
                           float *logits_out, cudaStream_t stream, ErrBuf &err, int *launches);
int32_t value_pass_launch(NetState &net, const TrainBuffers &tb, bool with_grad, float *v_out,
                          cudaStream_t stream, ErrBuf &err, int *launches);
// grad (already reduced, and all-reduced across ranks by the caller) -> Adam update of params (+ transposed copies)
// small batches on one rank: all value iterations of an epoch in one cooperative launch (mlp.cu)
bool value_iters_fused_supported(const NetState &net, const TrainBuffers &tb, int iters, int nranks);
int32_t value_iters_fused_launch(NetState &net, const TrainBuffers &tb, int iters, double lr, double b1, double b2, double eps,
                                 double *loss_first, double *loss_last, cudaStream_t stream, ErrBuf &err, int *launches);
// grad64 (all-reduced) -> grad, one rounding
int32_t grad_round_launch(NetState &net, cudaStream_t stream, ErrBuf &err, int *launches);
int32_t adam_launch(NetState &net, double lr, double b1, double b2, double eps, cudaStream_t stream,
                    ErrBuf &err, int *launches);

// ---- actor.cu -----------------------------------------------------------------------------
int32_t actor_check_dims(int d, int H, int A, ErrBuf &err);
// W1 [H][d] -> [H][roundup(d,4)] zero padded (16-byte aligned rows for cp.async)
int32_t actor_pad_w1_launch(const float *params, int H, int d, float *w1p, cudaStream_t stream, ErrBuf &err);
int32_t actor_forward_launch(const float *pi_params, const float *pi_w1p, const float *vf_params,
                             const float *vf_w1p, const unsigned char *pi_pack, const unsigned char *vf_pack, int d, int H, int A,
                             const float *obs, const float *mask, int64_t batch, uint64_t seed, uint64_t offset,
                             const int64_t *force_act, int64_t *act, float *logp_a, float *v, float *probs, float *logits,
                             cudaStream_t stream, ErrBuf &err);
// bf16x3 weight planes for the tensor-core actor kernel (actor_tc.cu); 0 bytes when the shape is not supported there
size_t actor_pack_bytes(int d, int H);
int32_t actor_pack_launch(const float *params, int d, int H, int out, unsigned char *pack, cudaStream_t stream, ErrBuf &err);

}  // namespace rrl
