// Shared between the CUDA-core (mlp.cu) and tensor-core (mlp_tc.cu) implementations of the fused
// learner pass.
#pragma once
#include "common.cuh"

namespace rrl {

constexpr int MLP_THREADS = 256;
constexpr int TILE_M = 64;       // transitions per tile
constexpr int FLUSH_TILES = 16;  // f32 accumulation window (1024 transitions) before the f64 spill
constexpr int MAX_OUT = 32;      // act_dim limit of the fused head
constexpr int HEAD_ACC = MAX_OUT / 4;

enum { MODE_VALUE = 0, MODE_POLICY_REF = 1, MODE_POLICY_LSM = 2 };

struct MlpArgs {
    const float *params;  // flat torch order
    int d, out;
    const float *obs, *act, *mask, *adv, *ret, *logp_old;
    const double *stats;  // AdvStats
    int64_t n;            // local transitions
    float inv_n;          // 1 / N_global as torch's mean-backward forms it (f32)
    double *partials;     // [grid][P] float64 gradient partials
    double *loss_partials;  // [grid][4]
    int64_t n_params;
    // optional forward taps (eval): per-row outputs
    float *out_rows;  // MODE_VALUE: v[n]; policy: logits [n][out]
};


// mlp_tc.cu: tcgen05 path (hidden = 128, obs_dim <= 16, head width <= 4, reference logp mode)
bool tc_pass_supported(const NetDims &nd, int mode, bool with_grad, int64_t rows);
size_t tc_smem_bytes();
int32_t tc_pass_launch(const MlpArgs &a, const NetDims &nd, int mode, int grid, cudaStream_t stream, ErrBuf &err);

}  // namespace rrl
