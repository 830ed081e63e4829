// Internal definition of the opaque handles behind include/relayrl_b200.h.
#pragma once
#include <vector>

#include "common.cuh"
#include "nccl_dl.h"

namespace rrl {

// One pinned staging slot of the ingest pipeline (double-buffered).
struct StagingSlot {
    char *host = nullptr;   // pinned
    char *dev = nullptr;    // device landing zone for f64 payloads (converted by K1b)
    size_t bytes = 0;
    cudaEvent_t done = nullptr;  // recorded after the slot's copies were enqueued
    bool in_flight = false;
};

}  // namespace rrl

struct rrl_learner {
    rrl_config_t cfg;
    rrl::ErrBuf err;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[8] = {};
    int sms = 148;

    // replay buffer, SoA f32 (reference: replay_buffer.py:20-28)
    float *obs = nullptr, *act = nullptr, *mask = nullptr, *rew = nullptr, *val = nullptr, *logp = nullptr,
          *adv = nullptr, *ret = nullptr;
    float *tmp = nullptr;              // [capacity] scratch for the normalised-advantage tap
    int64_t ptr = 0;                   // ReplayBuffer.ptr
    int64_t high_water = 0;            // largest ptr a scan has covered: adv/ret hold stale values up to here (Q7)
    std::vector<int64_t> offsets;      // successive path_start_idx values, offsets[0] = 0
    std::vector<float> last_vals;      // one per closed path
    int64_t *d_offsets = nullptr;
    float *d_last_vals = nullptr;
    int64_t d_offsets_cap = 0;
    bool finished = false;             // scan + stats are current for [0, ptr)
    int64_t n_global = 0, ep_global = 0;

    // scan workspace + statistics
    void *scan_ws = nullptr;
    size_t scan_ws_bytes = 0;
    double *stats = nullptr;           // AdvStats (5 doubles) + gather area [nranks][3]
    double *stats_gather = nullptr;

    // ingest staging
    rrl::StagingSlot slot[2];
    int cur_slot = 0;
    int64_t slot_rows = 0;             // rows a slot can hold
    int64_t staged = 0;                // rows packed into the current slot
    int staged_obs_dtype = RRL_F32, staged_mask_dtype = RRL_F32;

    // networks
    rrl::NetState pi, vf;

    // data parallel
    rrl::NcclApi *nccl = nullptr;
    void *comm = nullptr;              // ncclComm_t
    double *loss_sync = nullptr;       // [8] device doubles all-reduced once per epoch

    int launches = 0;
};

struct rrl_actor {
    int d = 0, A = 0, H = 0, with_baseline = 0, device = 0;
    int64_t max_batch = 0;
    rrl::ErrBuf err;
    cudaStream_t stream = nullptr;
    float *pi_params = nullptr, *pi_w1p = nullptr, *vf_params = nullptr, *vf_w1p = nullptr;
    unsigned char *pi_pack = nullptr, *vf_pack = nullptr;  // bf16x3 planes for the tensor-core kernel (null: shape not supported)
    // device mirrors of the host-facing buffers of rrl_actor_step
    float *d_obs = nullptr, *d_mask = nullptr, *d_logp = nullptr, *d_v = nullptr, *d_probs = nullptr, *d_logits = nullptr;
    int64_t *d_act = nullptr;
    cudaEvent_t order_ev = nullptr;  // ordering against the caller's default stream (rrl_actor_step_device)
};

namespace rrl {
// Device-pointer entry points run on the handle's own non-blocking stream, which is NOT implicitly ordered against the
// stream the caller filled its buffers on.  These two calls order it against the legacy default stream (stream 0: where
// torch and most plain CUDA code work): the library's work starts after everything queued there, and whatever the caller
// queues there next -- a free, a reuse of the buffers -- starts after the library's reads.
inline cudaError_t order_after_default_stream(cudaStream_t lib, cudaEvent_t ev)
{
    cudaError_t e = cudaEventRecord(ev, cudaStreamLegacy);
    return e != cudaSuccess ? e : cudaStreamWaitEvent(lib, ev, 0);
}
inline cudaError_t order_default_stream_after(cudaStream_t lib, cudaEvent_t ev)
{
    cudaError_t e = cudaEventRecord(ev, lib);
    return e != cudaSuccess ? e : cudaStreamWaitEvent(cudaStreamLegacy, ev, 0);
}
extern thread_local ErrBuf g_create_err;  // failures without a handle
int32_t learner_flush_staging(rrl_learner *h);
// all-or-nothing appends (wire decoders): where the replay buffer stood, and back to it
struct LearnerMark {
    int64_t ptr;
    size_t n_offsets, n_last_vals;
    bool finished;
};
int32_t learner_mark(rrl_learner *h, LearnerMark &m);
void learner_rollback(rrl_learner *h, const LearnerMark &m);
int32_t learner_finish(rrl_learner *h);
int32_t net_alloc(NetState &net, int d, int H, int out, int grid, bool train, bool data_parallel, ErrBuf &err);
void net_free(NetState &net);
}  // namespace rrl
