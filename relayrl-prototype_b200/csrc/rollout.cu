// Agent-side batched step service: request_for_action / flag_last_action for MANY
// environments per call, with the transitions kept packed on the device.
//
// In the reference an agent steps ONE environment per call (RelayRLAgent*::request_for_action,
// relayrl_framework/src/network/client/agent_zmq.rs:458-571, agent_grpc.rs:372-455; Python side
// src/bindings/python/network/client/o3_agent.rs:117-182): a batch-1 TorchScript step under a mutex, then five
// safetensors encodes to build the RelayRLAction that is pushed onto the trajectory (agent_zmq.rs:536-555);
// flag_last_action appends the terminal marker and ships the trajectory (:605-610, src/types/trajectory.rs:183-193).
// Here one call steps n_envs environments through K8 (csrc/actor.cu) and stores, per environment, exactly what
// that RelayRLAction would hold -- obs, sampled act, mask, the reward passed with THIS call, data{logp_a, v} -- in
// device staging laid out environment-major, so that a finished episode is one contiguous run of rows; flag_last
// appends those runs to the learner's replay buffer with device-to-device copies (the rrl_ingest_device path:
// no safetensors, no pickle, no host round trip) and records the terminal marker's reward as the path's last_val.
#include <vector>

#include "wire.h"

struct rrl_rollout {
    rrl_actor *actor = nullptr;
    rrl_learner *learner = nullptr;
    int64_t n_envs = 0, max_len = 0;
    float *obs = nullptr, *mask = nullptr, *act = nullptr, *rew = nullptr, *val = nullptr, *logp = nullptr;  // staging
    float *d_rew_in = nullptr;   // [n_envs] this call's rewards
    int *d_len = nullptr;        // [n_envs] rows already held per environment (uploaded per call)
    std::vector<int> len;        // host mirror
    cudaEvent_t copied = nullptr;
    rrl::ErrBuf err;
};

namespace rrl {
namespace {

// one thread per environment: append this step's row at (env, len[env])
__global__ void __launch_bounds__(256) rollout_store_kernel(const float *__restrict__ obs_in, const float *__restrict__ mask_in,
                                                           const float *__restrict__ rew_in, const int64_t *__restrict__ act_in,
                                                           const float *__restrict__ logp_in, const float *__restrict__ v_in,
                                                           const int *__restrict__ len, int64_t n_envs, int64_t max_len, int d,
                                                           int A, float *obs, float *mask, float *act, float *rew, float *val,
                                                           float *logp)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_envs) return;
    const int64_t row = e * max_len + len[e];
    for (int i = 0; i < d; ++i) obs[row * d + i] = obs_in[e * d + i];
    for (int i = 0; i < A; ++i) mask[row * A + i] = mask_in[e * A + i];
    act[row] = (float)act_in[e];  // the buffer keeps the action index as f32 (replay_buffer.py:21)
    rew[row] = rew_in[e];
    logp[row] = logp_in[e];
    val[row] = v_in ? v_in[e] : 0.f;
}

}  // namespace
}  // namespace rrl

using namespace rrl;

extern "C" int32_t rrl_rollout_destroy(rrl_rollout_t *r)
{
    if (!r) return RRL_OK;
    if (r->actor) {
        cudaSetDevice(r->actor->device);
        cudaStreamSynchronize(r->actor->stream);
    }
    cudaFree(r->obs), cudaFree(r->mask), cudaFree(r->act), cudaFree(r->rew), cudaFree(r->val), cudaFree(r->logp);
    cudaFree(r->d_rew_in), cudaFree(r->d_len);
    if (r->copied) cudaEventDestroy(r->copied);
    delete r;
    return RRL_OK;
}

extern "C" int32_t rrl_rollout_create(rrl_actor_t *actor, rrl_learner_t *learner, int64_t n_envs, int64_t max_len,
                                      rrl_rollout_t **out)
{
    if (!out) return RRL_EINVAL;
    *out = nullptr;
    ErrBuf &ce = g_create_err;
    if (!actor || !learner || n_envs < 1 || max_len < 1 || n_envs > actor->max_batch) {
        ce.set("rrl_rollout_create: needs an actor, a learner, 1 <= n_envs <= the actor's max_batch, max_len >= 1");
        return RRL_EINVAL;
    }
    if (actor->device != learner->cfg.device || actor->d != learner->cfg.obs_dim || actor->A != learner->cfg.act_dim ||
        (actor->with_baseline != 0) != (learner->cfg.with_baseline != 0)) {
        ce.set("rrl_rollout_create: actor and learner differ in device, obs_dim, act_dim or baseline");
        return RRL_EINVAL;
    }
    return abi_guard(ce, "rrl_rollout_create", [&]() -> int32_t {
        RRL_CUDA_OK(cudaSetDevice(actor->device), ce);
        rrl_rollout *r = new rrl_rollout();
        r->actor = actor, r->learner = learner, r->n_envs = n_envs, r->max_len = max_len;
        r->len.assign((size_t)n_envs, 0);
        const size_t rows = (size_t)n_envs * (size_t)max_len;
        bool ok = true;
        auto al = [&](void **p, size_t bytes) {
            if (ok) ok = cudaMalloc(p, bytes) == cudaSuccess;
        };
        al((void **)&r->obs, rows * actor->d * 4), al((void **)&r->mask, rows * actor->A * 4);
        al((void **)&r->act, rows * 4), al((void **)&r->rew, rows * 4), al((void **)&r->val, rows * 4), al((void **)&r->logp, rows * 4);
        al((void **)&r->d_rew_in, (size_t)n_envs * 4), al((void **)&r->d_len, (size_t)n_envs * 4);
        ok = ok && cudaEventCreateWithFlags(&r->copied, cudaEventDisableTiming) == cudaSuccess;
        if (!ok) {
            ce.set("rrl_rollout_create: device allocation of %zu staging rows failed: %s", rows, cudaGetErrorString(cudaGetLastError()));
            rrl_rollout_destroy(r);
            return RRL_ECUDA;
        }
        *out = r;
        return RRL_OK;
    });
}

extern "C" const char *rrl_rollout_last_error(const rrl_rollout_t *r) { return r ? r->err.msg.c_str() : g_create_err.msg.c_str(); }

// request_for_action(obs, mask, reward) for every environment: host buffers in, (act, logp_a, v) out.
extern "C" int32_t rrl_rollout_step(rrl_rollout_t *r, const float *obs, const float *mask, const float *reward, uint64_t seed,
                                    uint64_t offset, int64_t *act, float *logp_a, float *v)
{
    if (!r) return RRL_EINVAL;
    rrl_actor *a = r->actor;
    if (!obs || !mask || !reward || !act || !logp_a) {
        r->err.set("rrl_rollout_step: null pointer");
        return RRL_EINVAL;
    }
    if (cudaSetDevice(a->device) != cudaSuccess) {
        r->err.set("cudaSetDevice(%d) failed", a->device);
        return RRL_ECUDA;
    }
    const int64_t N = r->n_envs;
    for (int64_t e = 0; e < N; ++e)
        if (r->len[(size_t)e] >= r->max_len) {
            r->err.set("rrl_rollout_step: environment %lld already holds max_len = %lld transitions; flag its last action first",
                       (long long)e, (long long)r->max_len);
            return RRL_ECAPACITY;
        }
    cudaStream_t st = a->stream;
    const size_t B = (size_t)N;
    RRL_CUDA_OK(cudaMemcpyAsync(a->d_obs, obs, B * a->d * 4, cudaMemcpyHostToDevice, st), r->err);
    RRL_CUDA_OK(cudaMemcpyAsync(a->d_mask, mask, B * a->A * 4, cudaMemcpyHostToDevice, st), r->err);
    RRL_CUDA_OK(cudaMemcpyAsync(r->d_rew_in, reward, B * 4, cudaMemcpyHostToDevice, st), r->err);
    RRL_CUDA_OK(cudaMemcpyAsync(r->d_len, r->len.data(), B * 4, cudaMemcpyHostToDevice, st), r->err);
    int32_t s = actor_forward_launch(a->pi_params, a->pi_w1p, a->with_baseline ? a->vf_params : nullptr, a->vf_w1p, a->pi_pack, a->vf_pack, a->d, a->H,
                                     a->A, a->d_obs, a->d_mask, N, seed, offset, nullptr, a->d_act, a->d_logp, a->d_v, nullptr,
                                     nullptr, st, r->err);
    if (s != RRL_OK) return s;
    rollout_store_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(a->d_obs, a->d_mask, r->d_rew_in, a->d_act, a->d_logp,
                                                                      a->with_baseline ? a->d_v : nullptr, r->d_len, N,
                                                                      r->max_len, a->d, a->A, r->obs, r->mask, r->act, r->rew,
                                                                      r->val, r->logp);
    RRL_CUDA_OK(cudaGetLastError(), r->err);
    RRL_CUDA_OK(cudaMemcpyAsync(act, a->d_act, B * 8, cudaMemcpyDeviceToHost, st), r->err);
    RRL_CUDA_OK(cudaMemcpyAsync(logp_a, a->d_logp, B * 4, cudaMemcpyDeviceToHost, st), r->err);
    if (v && a->with_baseline) RRL_CUDA_OK(cudaMemcpyAsync(v, a->d_v, B * 4, cudaMemcpyDeviceToHost, st), r->err);
    RRL_CUDA_OK(cudaStreamSynchronize(st), r->err);  // the host reads `len` again on the next call; outputs are complete
    for (int64_t e = 0; e < N; ++e) r->len[(size_t)e] += 1;
    return RRL_OK;
}

// flag_last_action(reward) for the listed environments: their episodes go to the learner's replay buffer as they lie.
extern "C" int32_t rrl_rollout_flag_last(rrl_rollout_t *r, const int64_t *env_ids, const float *rewards, int64_t count)
{
    if (!r) return RRL_EINVAL;
    if (count < 0 || (count > 0 && (!env_ids || !rewards))) {
        r->err.set("rrl_rollout_flag_last: null pointer");
        return RRL_EINVAL;
    }
    rrl_actor *a = r->actor;
    const int d = a->d, A = a->A;
    for (int64_t i = 0; i < count; ++i) {
        const int64_t e = env_ids[i];
        if (e < 0 || e >= r->n_envs) {
            r->err.set("rrl_rollout_flag_last: environment %lld outside [0, %lld)", (long long)e, (long long)r->n_envs);
            return RRL_EINVAL;
        }
        const int64_t L = r->len[(size_t)e], row = e * r->max_len;
        const int64_t offs[2] = {0, L};
        // (the staging rows were written on the actor's stream, which rrl_rollout_step drained before returning)
        const int32_t s = rrl_ingest_device(r->learner, r->obs + row * d, r->act + row, r->mask + row * A, r->rew + row,
                                            r->val + row, r->logp + row, L, offs, &rewards[i], 1);
        if (s != RRL_OK) {
            r->err.set("rrl_rollout_flag_last: environment %lld: %s", (long long)e, rrl_last_error(r->learner));
            return s;
        }
        r->len[(size_t)e] = 0;
    }
    if (count > 0) {
        // the next step may overwrite these rows: the actor's stream waits for the learner's copies
        RRL_CUDA_OK(cudaEventRecord(r->copied, r->learner->stream), r->err);
        RRL_CUDA_OK(cudaStreamWaitEvent(a->stream, r->copied, 0), r->err);
    }
    return RRL_OK;
}
