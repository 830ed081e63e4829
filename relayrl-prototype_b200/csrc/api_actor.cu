// C ABI, part 3: batched actor inference (K8).
#include <string.h>

#include "learner.cuh"

using namespace rrl;

#define A_CHECK(a)                      \
    if (!(a)) return RRL_EINVAL;        \
    if (cudaSetDevice((a)->device) != cudaSuccess) { (a)->err.set("cudaSetDevice(%d) failed", (a)->device); return RRL_ECUDA; }

extern "C" {

const char *rrl_actor_last_error(const rrl_actor_t *a) { return a ? a->err.msg.c_str() : g_create_err.msg.c_str(); }

int32_t rrl_actor_destroy(rrl_actor_t *a)
{
    if (!a) return RRL_OK;
    cudaSetDevice(a->device);
    if (a->stream) cudaStreamSynchronize(a->stream);
    cudaFree(a->pi_params), cudaFree(a->pi_w1p), cudaFree(a->vf_params), cudaFree(a->vf_w1p);
    cudaFree(a->pi_pack), cudaFree(a->vf_pack);
    cudaFree(a->d_obs), cudaFree(a->d_mask), cudaFree(a->d_logp), cudaFree(a->d_v), cudaFree(a->d_probs);
    cudaFree(a->d_logits), cudaFree(a->d_act);
    if (a->order_ev) cudaEventDestroy(a->order_ev);
    if (a->stream) cudaStreamDestroy(a->stream);
    delete a;
    return RRL_OK;
}

int32_t rrl_actor_stream(const rrl_actor_t *a, void **stream)
{
    if (!a || !stream) return RRL_EINVAL;
    *stream = static_cast<void *>(a->stream);
    return RRL_OK;
}

int32_t rrl_actor_create(int32_t obs_dim, int32_t act_dim, int32_t hidden, int32_t with_baseline, int32_t device,
                         int64_t max_batch, rrl_actor_t **out)
{
    if (!out || max_batch < 1) return RRL_EINVAL;
    *out = nullptr;
    RRL_TRY(actor_check_dims(obs_dim, hidden, act_dim, g_create_err));
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        g_create_err.set("rrl_actor_create: CUDA device %d unavailable (%d visible); this library has no CPU path", device, ndev);
        return RRL_ECUDA;
    }
    RRL_CUDA_OK(cudaSetDevice(device), g_create_err);
    rrl_actor *a = new rrl_actor();
    a->d = obs_dim, a->A = act_dim, a->H = hidden, a->with_baseline = with_baseline, a->device = device;
    a->max_batch = max_batch;
    const NetDims pd{obs_dim, hidden, act_dim}, vd{obs_dim, hidden, 1};
    const size_t dK = (obs_dim + 3) & ~3, B = (size_t)max_batch;
    bool ok = cudaStreamCreateWithFlags(&a->stream, cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&a->order_ev, cudaEventDisableTiming) == cudaSuccess;
    auto al = [&](void **p, size_t bytes) {
        if (ok) ok = cudaMalloc(p, bytes) == cudaSuccess && cudaMemset(*p, 0, bytes) == cudaSuccess;
    };
    al((void **)&a->pi_params, pd.n_params() * 4), al((void **)&a->pi_w1p, hidden * dK * 4);
    if (with_baseline) al((void **)&a->vf_params, vd.n_params() * 4), al((void **)&a->vf_w1p, hidden * dK * 4);
    if (const size_t pk = actor_pack_bytes(obs_dim, hidden)) {
        al((void **)&a->pi_pack, pk);
        if (with_baseline) al((void **)&a->vf_pack, pk);
    }
    al((void **)&a->d_obs, B * obs_dim * 4), al((void **)&a->d_mask, B * act_dim * 4), al((void **)&a->d_logp, B * 4);
    al((void **)&a->d_v, B * 4), al((void **)&a->d_probs, B * act_dim * 4), al((void **)&a->d_logits, B * act_dim * 4);
    al((void **)&a->d_act, B * 8);
    if (!ok) {
        g_create_err.set("rrl_actor_create: device allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
        rrl_actor_destroy(a);
        return RRL_ECUDA;
    }
    *out = a;
    return RRL_OK;
}

int32_t rrl_actor_set_params(rrl_actor_t *a, int32_t which, const float *flat, int64_t n)
{
    A_CHECK(a);
    if (!flat || which < 0 || which > 1 || (which == 1 && !a->with_baseline)) {
        a->err.set("rrl_actor_set_params: which=%d invalid", which);
        return RRL_EINVAL;
    }
    const NetDims nd{a->d, a->H, which == 0 ? a->A : 1};
    if (n != nd.n_params()) {
        a->err.set("parameter count %lld != %lld", (long long)n, (long long)nd.n_params());
        return RRL_EINVAL;
    }
    float *dst = which == 0 ? a->pi_params : a->vf_params, *w1p = which == 0 ? a->pi_w1p : a->vf_w1p;
    RRL_CUDA_OK(cudaMemcpyAsync(dst, flat, n * 4, cudaMemcpyHostToDevice, a->stream), a->err);
    RRL_TRY(actor_pad_w1_launch(dst, a->H, a->d, w1p, a->stream, a->err));
    if (unsigned char *pack = which == 0 ? a->pi_pack : a->vf_pack)
        RRL_TRY(actor_pack_launch(dst, a->d, a->H, nd.out, pack, a->stream, a->err));
    RRL_CUDA_OK(cudaStreamSynchronize(a->stream), a->err);
    return RRL_OK;
}

int32_t rrl_actor_step_device(rrl_actor_t *a, const float *obs, const float *mask, int64_t batch, uint64_t seed,
                              uint64_t offset, const int64_t *force_act, int64_t *act, float *logp_a, float *v,
                              float *probs, float *logits)
{
    A_CHECK(a);
    if (batch < 0 || (batch > 0 && (!obs || !mask || !act || !logp_a))) {
        a->err.set("rrl_actor_step_device: null pointer");
        return RRL_EINVAL;
    }
    if (a->order_ev) RRL_CUDA_OK(order_after_default_stream(a->stream, a->order_ev), a->err);  // inputs made on stream 0
    RRL_TRY(actor_forward_launch(a->pi_params, a->pi_w1p, a->with_baseline ? a->vf_params : nullptr, a->vf_w1p, a->pi_pack, a->vf_pack, a->d,
                                 a->H, a->A, obs, mask, batch, seed, offset, force_act, act, logp_a, v, probs, logits,
                                 a->stream, a->err));
    RRL_CUDA_OK(cudaStreamSynchronize(a->stream), a->err);  // outputs are complete when the call returns
    return RRL_OK;
}

int32_t rrl_actor_step(rrl_actor_t *a, const float *obs, const float *mask, int64_t batch, uint64_t seed,
                       uint64_t offset, int64_t *act, float *logp_a, float *v, float *probs, float *logits)
{
    A_CHECK(a);
    if (batch < 0 || batch > a->max_batch || (batch > 0 && (!obs || !mask || !act || !logp_a))) {
        a->err.set("rrl_actor_step: batch=%lld (max %lld) or null pointer", (long long)batch, (long long)a->max_batch);
        return RRL_EINVAL;
    }
    if (batch == 0) return RRL_OK;
    cudaStream_t st = a->stream;
    const size_t B = (size_t)batch;
    RRL_CUDA_OK(cudaMemcpyAsync(a->d_obs, obs, B * a->d * 4, cudaMemcpyHostToDevice, st), a->err);
    RRL_CUDA_OK(cudaMemcpyAsync(a->d_mask, mask, B * a->A * 4, cudaMemcpyHostToDevice, st), a->err);
    RRL_TRY(actor_forward_launch(a->pi_params, a->pi_w1p, a->with_baseline ? a->vf_params : nullptr, a->vf_w1p, a->pi_pack, a->vf_pack, a->d,
                                 a->H, a->A, a->d_obs, a->d_mask, batch, seed, offset, nullptr, a->d_act, a->d_logp,
                                 a->d_v, probs ? a->d_probs : nullptr, logits ? a->d_logits : nullptr, st, a->err));
    RRL_CUDA_OK(cudaMemcpyAsync(act, a->d_act, B * 8, cudaMemcpyDeviceToHost, st), a->err);
    RRL_CUDA_OK(cudaMemcpyAsync(logp_a, a->d_logp, B * 4, cudaMemcpyDeviceToHost, st), a->err);
    if (v && a->with_baseline) RRL_CUDA_OK(cudaMemcpyAsync(v, a->d_v, B * 4, cudaMemcpyDeviceToHost, st), a->err);
    if (probs) RRL_CUDA_OK(cudaMemcpyAsync(probs, a->d_probs, B * a->A * 4, cudaMemcpyDeviceToHost, st), a->err);
    if (logits) RRL_CUDA_OK(cudaMemcpyAsync(logits, a->d_logits, B * a->A * 4, cudaMemcpyDeviceToHost, st), a->err);
    RRL_CUDA_OK(cudaStreamSynchronize(st), a->err);
    return RRL_OK;
}

}  // extern "C"
