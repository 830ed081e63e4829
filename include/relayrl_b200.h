/* relayrl_b200.h -- C ABI of librelayrl_b200.so: the B200-native (sm_100a) learner hot path of
 * RelayRL (trajectory ingest -> reward-to-go/GAE scan -> advantage normalisation -> policy/value
 * MLP forward/backward -> Adam -> parameter hand-off -> batched actor inference).
 *
 * Every entry point is `extern "C"`, takes plain pointers and sizes, returns an int32 status
 * (0 = RRL_OK, negative = RRL_E*), never throws, and never falls back to the CPU: without a
 * CUDA device the calls fail with RRL_ECUDA.  Citations are to the reference tree
 * (jrcalgo/RelayRL-prototype), paths relative to relayrl_framework/.
 *
 * Threading: calls on one handle must be serialised by the caller (the reference serialises
 * every algorithm call behind one asyncio.Lock, src/native/python/python_algorithm_reply.py:66,77,100)
 * but may come from different OS threads; every call selects the handle's device itself.
 * Different handles are independent.
 *
 * Data parallel (nranks > 1): rrl_finish, rrl_train_epoch, rrl_eval_grad and the parity taps that finish implicitly
 * (rrl_get_returns, rrl_get_advantages, rrl_get_adv_stats) are COLLECTIVE: every rank must make the same call in the same
 * order, like any NCCL program.  A rank that fails locally before the collective (RRL_ECAPACITY, RRL_ESTATE, a CUDA
 * error) returns its status at once and the other ranks would wait for it: treat any non-zero status of these calls as
 * fatal for the job and tear every rank down (the reference has no multi-process learner to mirror here).
 */
#ifndef RELAYRL_B200_H
#define RELAYRL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RRL_ABI_VERSION 1

/* status codes */
#define RRL_OK 0
#define RRL_EINVAL (-1)     /* bad argument / unsupported shape */
#define RRL_ECUDA (-2)      /* CUDA runtime error (see rrl_last_error) */
#define RRL_ECAPACITY (-3)  /* replay buffer full: reference `assert self.ptr < self.max_size`
                               (src/native/python/algorithms/REINFORCE/replay_buffer.py:38,87) */
#define RRL_ESTATE (-4)     /* call order violated (e.g. train before finish) */
#define RRL_ENCCL (-5)      /* NCCL unavailable or failed */
#define RRL_EFORMAT (-6)    /* malformed wire bytes (safetensors / pickle) */
#define RRL_ENOMEM (-7)     /* host memory exhausted (no exception ever crosses this ABI) */

/* element types of host payloads (wire dtypes, src/types/action.rs:92-101) */
#define RRL_F32 0
#define RRL_F64 1

/* how `logp` is defined: the reference gathers the RAW masked logit
 * (src/native/python/algorithms/REINFORCE/kernel.py:33-37); mode 1 is the textbook log-softmax. */
#define RRL_LOGP_REFERENCE 0
#define RRL_LOGP_LOG_SOFTMAX 1

/* rrl_config_t.flags.  RRL_FLAG_TC_ANY_SIZE: run the tensor-core learner kernel at every batch size; by default batches
 * under 32 768 transitions take the CUDA-core kernel (csrc/mlp_tc.cu, accuracy note).  For tests and diagnostics. */
#define RRL_FLAG_TC_ANY_SIZE 1
/* RRL_FLAG_NO_FUSED_ITERS: on one rank, batches of at most one 64-row tile per SM run all train_vf_iters value steps of an
 * epoch in one cooperative launch (bit-identical to one launch per step, ~3x less latency at CartPole sizes); this flag
 * keeps one launch per step.  For tests and diagnostics. */
#define RRL_FLAG_NO_FUSED_ITERS 2

#define RRL_NCCL_ID_BYTES 128

typedef struct rrl_learner rrl_learner_t;
typedef struct rrl_actor rrl_actor_t;
typedef struct rrl_rollout rrl_rollout_t;

/* Constructor arguments of the reference algorithm class
 * (REINFORCE.__init__, src/native/python/algorithms/REINFORCE/REINFORCE.py:17-62) plus the
 * hyper-parameters it reads through ConfigLoader (src/sys_utils/config_loader.rs:148-159;
 * floats are f32 in Rust and arrive as f32-rounded doubles, src/bindings/python/o3_config_loader.rs:76-84). */
typedef struct rrl_config {
    int32_t obs_dim;         /* d */
    int32_t act_dim;         /* A (discrete actions) */
    int32_t hidden;          /* H of the two hidden layers; reference hard-codes 128 (REINFORCE.py:46,49) */
    int32_t with_baseline;   /* with_vf_baseline */
    int32_t train_vf_iters;  /* value Adam steps per epoch (REINFORCE.py:111) */
    int32_t logp_mode;       /* RRL_LOGP_* */
    int32_t device;          /* CUDA ordinal */
    int32_t rank;            /* data-parallel rank; 0 when nranks == 1 */
    int32_t nranks;          /* 1 = single GPU */
    int32_t flags;           /* RRL_FLAG_*; 0 by default */
    int64_t capacity;        /* buf_size, transitions */
    double gamma, lam;       /* replay_buffer.py:30 */
    double pi_lr, vf_lr;     /* REINFORCE.py:47,50 */
    double adam_beta1, adam_beta2, adam_eps; /* torch.optim.Adam defaults 0.9, 0.999, 1e-8 */
    uint8_t nccl_id[RRL_NCCL_ID_BYTES]; /* from rrl_nccl_unique_id on rank 0, when nranks > 1 */
} rrl_config_t;

/* A run of actions as the reference's receive_trajectory sees them
 * (REINFORCE.py:70-95): row i is a transition (done[i] == 0) or a terminal marker
 * (done[i] != 0: nothing is stored, rew[i] becomes the path's last_val, REINFORCE.py:86).
 * Host pointers, struct-of-arrays, row-major.  Rows of terminal markers are ignored in
 * obs/act/mask/val/logp.  Wire dtypes: obs/mask F64 from Python agents, else F32. */
typedef struct rrl_batch_view {
    const void *obs;     /* [count, obs_dim], obs_dtype */
    const float *act;    /* [count] action index stored as f32 (replay_buffer.py:21) */
    const void *mask;    /* [count, act_dim], mask_dtype */
    const float *rew;    /* [count] */
    const float *val;    /* [count] data['v'];   may be NULL without baseline */
    const float *logp;   /* [count] data['logp_a'] */
    const uint8_t *done; /* [count] */
    int64_t count;
    int32_t obs_dtype, mask_dtype; /* RRL_F32 | RRL_F64 */
} rrl_batch_view_t;

/* What REINFORCE.train_model stores in its logger (REINFORCE.py:117-125) + buffer statistics. */
typedef struct rrl_epoch_stats {
    double loss_pi, delta_loss_pi, kl, entropy; /* delta_loss_pi is identically 0 */
    double loss_v, delta_loss_v, loss_v_old;
    double adv_mean, adv_std;                   /* statistics_scalar, BaseReplayBuffer.py:30-53 */
    int64_t n_transitions, n_episodes;          /* global over all ranks */
    /* device time of the stages of this call, CUDA events on the library stream, milliseconds */
    float ms_scan, ms_policy, ms_value, ms_total;
    int32_t kernel_launches;                    /* kernels launched by this call */
    int32_t reserved;
} rrl_epoch_stats_t;

/* ---- library ------------------------------------------------------------------------- */
int32_t rrl_abi_version(void);
const char *rrl_status_string(int32_t status);
/* Message of the last failure on this handle (owned by the handle), or of the last failed
 * create when h == NULL (thread-local). */
const char *rrl_last_error(const rrl_learner_t *h);
int32_t rrl_device_count(int32_t *count);
/* Pinned host memory for callers that want the ingest copy to be truly asynchronous. */
int32_t rrl_host_alloc(void **ptr, size_t bytes);
int32_t rrl_host_free(void *ptr);
int32_t rrl_nccl_unique_id(uint8_t out[RRL_NCCL_ID_BYTES]);
void rrl_config_defaults(rrl_config_t *cfg);

/* ---- learner: replaces REINFORCE + ReplayBuffer ------------------------------------------ */
/* REINFORCE.__init__ (REINFORCE.py:17-62).  Weights start at zero: call rrl_set_params. */
int32_t rrl_learner_create(const rrl_config_t *cfg, rrl_learner_t **out);
int32_t rrl_learner_destroy(rrl_learner_t *h);
/* The CUDA stream (cudaStream_t) every call on this handle is ordered on.  No counterpart in the reference
 * (its learner is synchronous CPU code); callers use it to order their own device work against the
 * learner's, or to time calls with CUDA events (bench.py does). */
int32_t rrl_learner_stream(const rrl_learner_t *h, void **stream);

/* Flat parameters in torch.nn.Linear order W1,b1,W2,b2,W3,b3, W row-major [out,in]
 * (kernel.py:15-21; BaseKernel.py:25-39).  which: 0 = policy.pi_network, 1 = baseline.v_network. */
int32_t rrl_param_count(const rrl_learner_t *h, int32_t which, int64_t *n);
int32_t rrl_set_params(rrl_learner_t *h, int32_t which, const float *flat, int64_t n);
int32_t rrl_get_params(rrl_learner_t *h, int32_t which, float *flat, int64_t n);
/* Adam moments + step count (torch.optim.Adam state: exp_avg, exp_avg_sq, step). */
int32_t rrl_get_opt_state(rrl_learner_t *h, int32_t which, float *m, float *v, int64_t n, int64_t *step);
int32_t rrl_set_opt_state(rrl_learner_t *h, int32_t which, const float *m, const float *v, int64_t n, int64_t step);

/* Checkpoint / resume of the training state (both networks' parameters, Adam moments and step counts, plus a caller tag
 * such as the epoch counter) in one file, written atomically (temp file + rename).  No counterpart in the reference: a
 * restarted server starts from fresh weights (src/network/server/training_zmq.rs:322-465), only the TorchScript policy is
 * ever saved (REINFORCE.py:64-68).  Take it between epochs (the replay buffer is not saved).  A run resumed from a
 * checkpoint continues bit-identically to the uninterrupted run.  Load fails with RRL_EFORMAT / RRL_EINVAL -- and leaves
 * the learner untouched -- when the file is damaged or was written for other dimensions. */
int32_t rrl_checkpoint_save(rrl_learner_t *h, const char *path, uint64_t user_tag);
int32_t rrl_checkpoint_load(rrl_learner_t *h, const char *path, uint64_t *user_tag);

/* receive_trajectory's per-action loop + ReplayBuffer.store (REINFORCE.py:70-87,
 * replay_buffer.py:34-46): packs rows into pinned staging, copies asynchronously into the flat
 * device buffers, records episode offsets (successive path_start_idx values) on the host. */
int32_t rrl_ingest_packed(rrl_learner_t *h, const rrl_batch_view_t *batch);
/* Same, with device-resident SoA f32 inputs already free of terminal markers.  Stream contract: the copies run on the
 * learner's own stream (rrl_learner_stream), ordered AFTER everything queued on the legacy default stream (stream 0)
 * at the time of the call, and work queued on stream 0 after the call waits for them -- so a caller that fills and
 * frees its buffers on the default stream (torch's default) needs no synchronisation.  A caller working on another
 * stream must make its inputs complete before the call (e.g. cudaStreamSynchronize, or an event the default stream
 * waits on) and keep them alive until rrl_learner_stream has been synchronised.  rrl_actor_step_device orders its
 * inputs the same way and returns with its outputs complete.  `offsets`
 * (n_ep + 1 entries relative to this batch, offsets[0] == 0) and `last_vals` (n_ep) are host arrays. */
int32_t rrl_ingest_device(rrl_learner_t *h, const float *obs, const float *act, const float *mask,
                          const float *rew, const float *val, const float *logp, int64_t count,
                          const int64_t *offsets, const float *last_vals, int64_t n_ep);
/* Wire formats: one safetensors blob per tensor field
 * (src/types/action.rs:196-218,287-418; src/sys_utils/grpc_utils.rs:31-66 for the proto fields). */
int32_t rrl_ingest_wire_action(rrl_learner_t *h, const uint8_t *obs_blob, size_t obs_len,
                               const uint8_t *act_blob, size_t act_len, const uint8_t *mask_blob,
                               size_t mask_len, float rew, const uint8_t *v_blob, size_t v_len,
                               const uint8_t *logp_blob, size_t logp_len, uint8_t done);
/* gRPC wire format: one serialized proto `Trajectory` (proto/relayrl_grpc.proto:5-18) as TrainingServerGrpc::
 * send_actions receives it (src/network/server/training_grpc.rs:585-599).  Replaces deserialize_action /
 * grpc_trajectory_to_relayrl_trajectory (src/sys_utils/grpc_utils.rs:83-158) and everything behind them
 * (JSON pipe, PyO3 getters, REINFORCE.receive_trajectory's per-action loop): obs/action/mask are safetensors
 * blobs, data["logp_a"] / data["v"] are serde_json RelayRLData values (action.rs:203-218), an action with
 * done = true is a terminal marker carrying only its reward.  Appends every action of the message. */
int32_t rrl_ingest_grpc_trajectory(rrl_learner_t *h, const uint8_t *msg, size_t len, int64_t *n_actions);
/* The decoder alone, host only (no GPU needed): fills caller-provided SoA f32 columns of `cap` rows in the
 * rrl_batch_view_t convention and returns the row count; errbuf (optional) receives the failure message. */
int32_t rrl_decode_grpc_trajectory(const uint8_t *msg, size_t len, int32_t obs_dim, int32_t act_dim,
                                   int32_t with_baseline, float *obs, float *act, float *mask, float *rew,
                                   float *val, float *logp, uint8_t *done, int64_t cap, int64_t *count,
                                   char *errbuf, size_t errcap);

/* ZMQ wire format: the bytes an agent PUSHes, serde_pickle::to_writer(&Vec<RelayRLAction>) (src/types/trajectory.rs:50-55),
 * as TrainingServerZmq::_start_training_loop receives them (src/network/server/training_zmq.rs:994-1001).  Replaces
 * pickle::from_slice + the JSON pipe + traj_from_json + receive_trajectory's per-action loop.  The decoder is a plain-data
 * pickle reader (protocols 2-4; no globals, no reduce) that accepts serde's data model for these types under both
 * encodings of Vec<u8> (list of ints / bytes), of unit variants (str / 1-tuple) and of newtype variants (1-entry dict /
 * 2-tuple); anything else is RRL_EFORMAT.  All of the message is appended, or -- on any error -- none of it (the same
 * holds for rrl_ingest_grpc_trajectory).  serde-pickle 1.2.0 itself is not in the reference tree: verified against
 * Python's pickler on the equivalent object graphs, not against a live Rust peer. */
int32_t rrl_ingest_pickle(rrl_learner_t *h, const uint8_t *msg, size_t len, int64_t *n_actions);
/* The decoder alone, host only (same convention as rrl_decode_grpc_trajectory). */
int32_t rrl_decode_pickle_trajectory(const uint8_t *msg, size_t len, int32_t obs_dim, int32_t act_dim,
                                     int32_t with_baseline, float *obs, float *act, float *mask, float *rew,
                                     float *val, float *logp, uint8_t *done, int64_t cap, int64_t *count,
                                     char *errbuf, size_t errcap);

/* ReplayBuffer.finish_path for every closed path of the epoch (replay_buffer.py:48-79) and
 * statistics_scalar over [0, ptr) (BaseReplayBuffer.py:30-53): one fused segmented reverse scan.
 * Returns this rank's counts. */
int32_t rrl_finish(rrl_learner_t *h, int64_t *n_transitions, int64_t *n_episodes);

/* Parity taps (host outputs).  Offsets are the successive path_start_idx values, offsets[0] = 0. */
int32_t rrl_get_offsets(rrl_learner_t *h, int64_t *out, int64_t cap, int64_t *n_written);
int32_t rrl_get_returns(rrl_learner_t *h, float *out, int64_t cap);
/* normalised != 0: (adv - mean) / std as ReplayBuffer.get returns it (replay_buffer.py:94-105). */
int32_t rrl_get_advantages(rrl_learner_t *h, float *out, int64_t cap, int32_t normalised);
int32_t rrl_get_adv_stats(rrl_learner_t *h, float *mean, float *std);

/* REINFORCE.train_model (REINFORCE.py:97-125): ReplayBuffer.get() semantics (pointers reset),
 * one policy Adam step, train_vf_iters value Adam steps.  Calls rrl_finish if it was not called. */
int32_t rrl_train_epoch(rrl_learner_t *h, rrl_epoch_stats_t *stats);
/* Pieces of the above for parity tests: losses and flat gradients at the current parameters on
 * the current (finished) buffer; no update. which: 0 policy (out3 = loss_pi, kl, entropy), 1 value (loss_v). */
int32_t rrl_eval_grad(rrl_learner_t *h, int32_t which, float *grad, int64_t n, double out3[3]);
int32_t rrl_reset_buffer(rrl_learner_t *h);

/* ---- actor: replaces PolicyWith[out]Baseline.step (kernel.py:99-107,128-135) as called from
 * RelayRLAgent*::request_for_action (src/network/client/agent_zmq.rs:458-571) ----------------- */
int32_t rrl_actor_create(int32_t obs_dim, int32_t act_dim, int32_t hidden, int32_t with_baseline,
                         int32_t device, int64_t max_batch, rrl_actor_t **out);
int32_t rrl_actor_destroy(rrl_actor_t *a);
/* The CUDA stream the actor's kernels run on (same role as rrl_learner_stream; no reference counterpart). */
int32_t rrl_actor_stream(const rrl_actor_t *a, void **stream);
int32_t rrl_actor_set_params(rrl_actor_t *a, int32_t which, const float *flat, int64_t n);
/* One batched step on host buffers: act int64 [batch], logp_a f32 [batch] (= logits[i, act_i]),
 * v f32 [batch] (NULL without baseline).  Sampling: inverse-CDF on softmax(logits) with one
 * Philox4x32-10 uniform per env keyed by (seed, offset + env index).  probs/logits may be NULL. */
int32_t rrl_actor_step(rrl_actor_t *a, const float *obs, const float *mask, int64_t batch,
                       uint64_t seed, uint64_t offset, int64_t *act, float *logp_a, float *v,
                       float *probs, float *logits);
/* Same with device pointers; force_act (device, may be NULL) replaces the sampled action. */
int32_t rrl_actor_step_device(rrl_actor_t *a, const float *obs, const float *mask, int64_t batch,
                              uint64_t seed, uint64_t offset, const int64_t *force_act,
                              int64_t *act, float *logp_a, float *v, float *probs, float *logits);
const char *rrl_actor_last_error(const rrl_actor_t *a);

/* ---- agent-side batched step service --------------------------------------------
 * request_for_action / flag_last_action (src/network/client/agent_zmq.rs:458-571,605-610; agent_grpc.rs:372-455) for
 * n_envs environments per call instead of one.  rrl_rollout_step runs the actor's batched step on host observations and
 * stores, for every environment, the RelayRLAction the reference would build (obs, sampled act, mask, the reward passed
 * with this call, data{logp_a, v}: agent_zmq.rs:536-552) as one packed row in device staging, environment-major, at
 * most max_len rows per environment.  rrl_rollout_flag_last closes the listed environments' episodes: the terminal
 * marker's reward becomes the path's last_val (REINFORCE.py:86) and the rows move into the learner's replay buffer
 * device-to-device -- no safetensors blobs, no pickle, no host copy of the transitions.  Actor and learner must live on
 * the same device with the same obs_dim / act_dim / baseline.  Calls on one rollout are serialised by the caller. */
int32_t rrl_rollout_create(rrl_actor_t *actor, rrl_learner_t *learner, int64_t n_envs, int64_t max_len, rrl_rollout_t **out);
int32_t rrl_rollout_destroy(rrl_rollout_t *r);
int32_t rrl_rollout_step(rrl_rollout_t *r, const float *obs, const float *mask, const float *reward, uint64_t seed,
                         uint64_t offset, int64_t *act, float *logp_a, float *v);
int32_t rrl_rollout_flag_last(rrl_rollout_t *r, const int64_t *env_ids, const float *rewards, int64_t count);
const char *rrl_rollout_last_error(const rrl_rollout_t *r);

/* ---- stateless device operators (for callers that own their device memory) ----------------- */
/* Scratch the scan needs for n transitions (tile descriptors for the single-pass look-back). */
int32_t rrl_scan_workspace_bytes(int64_t n, size_t *bytes);
/* finish_path over all closed paths + adv statistics, everything device-resident except scalars.
 * rew/val/adv/ret: [n_total] f32, 16-byte aligned; offsets: [n_ep+1] int64; last_vals: [n_ep] f32.
 * stats_out (device, 5 doubles): sum, sumsq, n, mean(f32 value), std(f32 value) over adv[0, n_total).
 * Elements >= offsets[n_ep] (unfinished tail) keep their adv/ret. stream = cudaStream_t. */
int32_t rrl_scan_device(const float *rew, const float *val, const int64_t *offsets,
                        const float *last_vals, int64_t n_total, int64_t n_ep, double gamma,
                        double lam, int32_t with_baseline, float *adv, float *ret,
                        double *stats_out, void *workspace, size_t workspace_bytes, void *stream);
/* Same operator for callers that cannot allocate device memory (no counterpart in the reference:
 * its learner never touches a GPU).  Meant for hosts whose device-memory budget is fixed at
 * start-up -- a locked memory pool, a quota with no head-room, a plug-in inside an application
 * that forbids allocation calls after initialisation -- and which therefore cannot supply the
 * workspace rrl_scan_device wants.  The look-back descriptors live in a 393 472-byte array
 * compiled into the library (resident on every GPU that loads it, until the process exits), so
 * n_total is limited to RRL_SCAN_NOALLOC_MAX transitions; larger calls fail with RRL_EINVAL.
 * The call itself performs no cudaMalloc/cudaFree and uses no allocating library routine.
 * The first call on each device resolves the array's device address (cudaGetSymbolAddress, with
 * the caller's current device): make that first call before starting any stream capture.
 * All calls on a device share the one array: two calls that overlap on the GPU (different
 * streams, or different threads) race on it -- serialise them.  The array keeps the previous
 * call's contents; the part that must start from a known state is reset on `stream` by every call.
 * Arguments and results are exactly those of rrl_scan_device, minus the workspace. */
#define RRL_SCAN_NOALLOC_MAX (16ll * 1024 * 1024)
int32_t rrl_scan_device_noalloc(const float *rew, const float *val, const int64_t *offsets,
                                const float *last_vals, int64_t n_total, int64_t n_ep, double gamma,
                                double lam, int32_t with_baseline, float *adv, float *ret,
                                double *stats_out, void *stream);
/* out[i] = (adv[i] - mean) / std in f32, mean/std read from stats (device, as written above). */
int32_t rrl_normalise_device(const float *adv, int64_t n, const double *stats, float *out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* RELAYRL_B200_H */
