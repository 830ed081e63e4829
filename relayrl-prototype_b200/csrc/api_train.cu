// C ABI, part 2: train_epoch (K4-K7 orchestration + NCCL), gradient taps, stateless operators.
#include <string.h>

#include "learner.cuh"

using namespace rrl;

#define H_CHECK(h)                       \
    if (!(h)) return RRL_EINVAL;         \
    if (cudaSetDevice((h)->cfg.device) != cudaSuccess) { (h)->err.set("cudaSetDevice(%d) failed", (h)->cfg.device); return RRL_ECUDA; }

static int32_t allreduce_grad(rrl_learner *h, NetState &net)
{
    if (h->cfg.nranks <= 1) return RRL_OK;
    // The sum over ranks is formed in float64 and rounded to f32 once, exactly like the single-GPU sum over CTAs
    // (grad_reduce_kernel): N ranks and one rank then differ by float64 re-association only.  139 KB: latency-bound.
    int rc = h->nccl->AllReduce(net.grad64, net.grad64, (size_t)net.dims.n_params(), NCCL_FLOAT64, NCCL_SUM, h->comm, h->stream);
    if (rc != 0) {
        h->err.set("ncclAllReduce(grad) failed: %s", h->nccl->GetErrorString(rc));
        return RRL_ENCCL;
    }
    return grad_round_launch(net, h->stream, h->err, &h->launches);
}

static TrainBuffers train_buffers(rrl_learner *h)
{
    TrainBuffers tb;
    tb.obs = h->obs, tb.act = h->act, tb.mask = h->mask, tb.adv = h->adv, tb.ret = h->ret, tb.logp_old = h->logp;
    tb.stats = h->stats, tb.n = h->ptr, tb.n_global = h->n_global;
    tb.tc_any_size = (h->cfg.flags & RRL_FLAG_TC_ANY_SIZE) != 0;
    return tb;
}

extern "C" {

int32_t rrl_train_epoch(rrl_learner_t *h, rrl_epoch_stats_t *out)
{
    H_CHECK(h);
    if (!h->finished) RRL_TRY(learner_finish(h));
    // ReplayBuffer.get(): `assert self.ptr < self.max_size` (replay_buffer.py:87)
    if (h->ptr >= h->cfg.capacity) {
        h->err.set("ReplayBuffer.get: ptr=%lld must be < capacity %lld (replay_buffer.py:87)", (long long)h->ptr,
                   (long long)h->cfg.capacity);
        return RRL_ECAPACITY;
    }
    if (h->n_global <= 0) {
        h->err.set("rrl_train_epoch: empty buffer");
        return RRL_ESTATE;
    }
    cudaStream_t st = h->stream;
    const int launches0 = h->launches;
    const TrainBuffers tb = train_buffers(h);
    const rrl_config_t &c = h->cfg;
    RRL_CUDA_OK(cudaMemsetAsync(h->loss_sync, 0, 8 * sizeof(double), st), h->err);

    // ---- policy: one Adam step (REINFORCE.py:105-108). pi_l_old equals loss_pi (no update in between)
    RRL_CUDA_OK(cudaEventRecord(h->ev[2], st), h->err);
    RRL_TRY(policy_pass_launch(h->pi, tb, c.logp_mode, true, nullptr, st, h->err, &h->launches));
    RRL_CUDA_OK(cudaMemcpyAsync(h->loss_sync, h->pi.loss_out, 3 * sizeof(double), cudaMemcpyDeviceToDevice, st), h->err);
    RRL_TRY(allreduce_grad(h, h->pi));
    RRL_TRY(adam_launch(h->pi, c.pi_lr, c.adam_beta1, c.adam_beta2, c.adam_eps, st, h->err, &h->launches));
    RRL_CUDA_OK(cudaEventRecord(h->ev[3], st), h->err);

    // ---- value: train_vf_iters Adam steps (REINFORCE.py:110-115)
    if (c.with_baseline) {
        if (c.train_vf_iters == 0) {
            RRL_TRY(value_pass_launch(h->vf, tb, false, nullptr, st, h->err, &h->launches));
            RRL_CUDA_OK(cudaMemcpyAsync(h->loss_sync + 3, h->vf.loss_out, sizeof(double), cudaMemcpyDeviceToDevice, st), h->err);
            RRL_CUDA_OK(cudaMemcpyAsync(h->loss_sync + 4, h->vf.loss_out, sizeof(double), cudaMemcpyDeviceToDevice, st), h->err);
        }
        // small batch on one rank: all the steps in one cooperative launch, bit-identical to the loop below (mlp.cu)
        const bool fused = c.train_vf_iters > 0 && !(c.flags & RRL_FLAG_NO_FUSED_ITERS) &&
                           value_iters_fused_supported(h->vf, tb, c.train_vf_iters, c.nranks);
        if (fused)
            RRL_TRY(value_iters_fused_launch(h->vf, tb, c.train_vf_iters, c.vf_lr, c.adam_beta1, c.adam_beta2, c.adam_eps,
                                             h->loss_sync + 3, h->loss_sync + 4, st, h->err, &h->launches));
        for (int i = 0; i < (fused ? 0 : c.train_vf_iters); ++i) {
            RRL_TRY(value_pass_launch(h->vf, tb, true, nullptr, st, h->err, &h->launches));
            if (i == 0)  // v_l_old: loss before the first step (REINFORCE.py:103)
                RRL_CUDA_OK(cudaMemcpyAsync(h->loss_sync + 3, h->vf.loss_out, sizeof(double), cudaMemcpyDeviceToDevice, st), h->err);
            if (i == c.train_vf_iters - 1)  // loss_v of the last iteration, computed before its step (REINFORCE.py:113)
                RRL_CUDA_OK(cudaMemcpyAsync(h->loss_sync + 4, h->vf.loss_out, sizeof(double), cudaMemcpyDeviceToDevice, st), h->err);
            RRL_TRY(allreduce_grad(h, h->vf));
            RRL_TRY(adam_launch(h->vf, c.vf_lr, c.adam_beta1, c.adam_beta2, c.adam_eps, st, h->err, &h->launches));
        }
    }
    RRL_CUDA_OK(cudaEventRecord(h->ev[4], st), h->err);
    if (c.nranks > 1) {
        int rc = h->nccl->AllReduce(h->loss_sync, h->loss_sync, 5, NCCL_FLOAT64, NCCL_SUM, h->comm, st);
        if (rc != 0) {
            h->err.set("ncclAllReduce(losses) failed: %s", h->nccl->GetErrorString(rc));
            return RRL_ENCCL;
        }
    }
    double ls[8], stats[6];
    RRL_CUDA_OK(cudaMemcpyAsync(ls, h->loss_sync, sizeof ls, cudaMemcpyDeviceToHost, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(stats, h->stats, sizeof stats, cudaMemcpyDeviceToHost, st), h->err);
    RRL_CUDA_OK(cudaStreamSynchronize(st), h->err);

    if (out) {
        memset(out, 0, sizeof *out);
        const double N = (double)h->n_global;
        out->loss_pi = -ls[0] / N;
        out->delta_loss_pi = 0.0;
        out->kl = ls[1] / N;
        out->entropy = -ls[2] / N;
        if (c.with_baseline) {
            out->loss_v_old = ls[3] / N;
            out->loss_v = ls[4] / N;
            out->delta_loss_v = out->loss_v - out->loss_v_old;
        }
        out->adv_mean = stats[3], out->adv_std = stats[4];
        out->n_transitions = h->n_global, out->n_episodes = h->ep_global;
        cudaEventElapsedTime(&out->ms_scan, h->ev[0], h->ev[1]);
        cudaEventElapsedTime(&out->ms_policy, h->ev[2], h->ev[3]);
        cudaEventElapsedTime(&out->ms_value, h->ev[3], h->ev[4]);
        out->ms_total = out->ms_scan + out->ms_policy + out->ms_value;
        out->kernel_launches = h->launches - launches0;
    }
    // ReplayBuffer.get() resets the pointers (replay_buffer.py:88-89); buffers keep their contents
    h->ptr = 0, h->offsets.assign(1, 0), h->last_vals.clear(), h->finished = false;
    return RRL_OK;
}

int32_t rrl_eval_grad(rrl_learner_t *h, int32_t which, float *grad, int64_t n, double out3[3])
{
    H_CHECK(h);
    if (which < 0 || which > 1 || (which == 1 && !h->cfg.with_baseline)) {
        h->err.set("which=%d: no such network", which);
        return RRL_EINVAL;
    }
    if (!h->finished) RRL_TRY(learner_finish(h));
    NetState &net = which == 0 ? h->pi : h->vf;
    if (grad && n != net.dims.n_params()) {
        h->err.set("gradient size %lld != %lld", (long long)n, (long long)net.dims.n_params());
        return RRL_EINVAL;
    }
    if (h->n_global <= 0) {
        h->err.set("rrl_eval_grad: empty buffer");
        return RRL_ESTATE;
    }
    const TrainBuffers tb = train_buffers(h);
    if (which == 0)
        RRL_TRY(policy_pass_launch(net, tb, h->cfg.logp_mode, true, nullptr, h->stream, h->err, &h->launches));
    else
        RRL_TRY(value_pass_launch(net, tb, true, nullptr, h->stream, h->err, &h->launches));
    RRL_TRY(allreduce_grad(h, net));
    double ls[4];
    if (grad) RRL_CUDA_OK(cudaMemcpyAsync(grad, net.grad, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(ls, net.loss_out, sizeof ls, cudaMemcpyDeviceToHost, h->stream), h->err);
    RRL_CUDA_OK(cudaStreamSynchronize(h->stream), h->err);
    if (out3) {
        const double N = (double)h->n_global;  // local sums over the global N (single rank: exact)
        if (which == 0)
            out3[0] = -ls[0] / N, out3[1] = ls[1] / N, out3[2] = -ls[2] / N;
        else
            out3[0] = ls[0] / N, out3[1] = 0.0, out3[2] = 0.0;
    }
    return RRL_OK;
}

// ---- stateless device operators ------------------------------------------------------------
int32_t rrl_scan_workspace_bytes(int64_t n, size_t *bytes)
{
    if (n < 0 || !bytes) return RRL_EINVAL;
    *bytes = scan_workspace_bytes(n);
    return RRL_OK;
}

int32_t rrl_scan_device(const float *rew, const float *val, const int64_t *offsets, const float *last_vals,
                        int64_t n_total, int64_t n_ep, double gamma, double lam, int32_t with_baseline, float *adv,
                        float *ret, double *stats_out, void *workspace, size_t workspace_bytes, void *stream)
{
    if (!rew || !adv || !ret || (with_baseline && !val) || (n_ep > 0 && (!offsets || !last_vals))) {
        rrl::g_create_err.set("rrl_scan_device: null pointer");
        return RRL_EINVAL;
    }
    return scan_launch(rew, val, offsets, last_vals, n_total, n_ep, gamma, lam, with_baseline != 0, adv, ret, stats_out,
                       workspace, workspace_bytes, true, -1, static_cast<cudaStream_t>(stream), rrl::g_create_err, nullptr);
}

int32_t rrl_normalise_device(const float *adv, int64_t n, const double *stats, float *out, void *stream)
{
    if (!adv || !stats || !out || n < 0) return RRL_EINVAL;
    return normalise_launch(adv, n, stats, out, static_cast<cudaStream_t>(stream), rrl::g_create_err);
}

}  // extern "C"
