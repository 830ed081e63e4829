// C ABI, part 1: library helpers, learner life-cycle, parameters, ingest (K1), finish (K2).
#include <stdarg.h>
#include <string.h>

#include <vector>

#include "wire.h"

namespace rrl {

void ErrBuf::set(const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    msg = buf;
}

thread_local ErrBuf g_create_err;

template <typename T>
static int32_t dmalloc(T **p, size_t count, ErrBuf &err, bool zero = true)
{
    *p = nullptr;
    if (count == 0) count = 1;
    RRL_CUDA_OK(cudaMalloc(reinterpret_cast<void **>(p), count * sizeof(T)), err);
    if (zero) RRL_CUDA_OK(cudaMemset(*p, 0, count * sizeof(T)), err);
    return RRL_OK;
}

int32_t net_alloc(NetState &net, int d, int H, int out, int grid, bool train, bool data_parallel, ErrBuf &err)
{
    net.dims = NetDims{d, H, out};
    net.step = 0;
    net.grid = grid;
    const size_t P = (size_t)net.dims.n_params();
    RRL_TRY(dmalloc(&net.params, P, err));
    net.grad = net.m = net.v = nullptr;
    net.grad64 = nullptr;
    net.fused_barrier = nullptr;
    net.partials64 = net.loss_partials = net.loss_out = nullptr;
    if (train) {
        RRL_TRY(dmalloc(&net.grad, P, err));
        RRL_TRY(dmalloc(&net.m, P, err));
        RRL_TRY(dmalloc(&net.v, P, err));
        RRL_TRY(dmalloc(&net.partials64, P * (size_t)grid, err));
        RRL_TRY(dmalloc(&net.loss_partials, 4 * (size_t)grid, err));
        RRL_TRY(dmalloc(&net.loss_out, 4, err));
        if (data_parallel) RRL_TRY(dmalloc(&net.grad64, P, err));
        int dev = 0, coop = 0;
        if (!data_parallel && cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev) == cudaSuccess && coop)
            RRL_TRY(dmalloc(&net.fused_barrier, 4, err));  // (without cooperative launch the fused small-batch path stays off)
    }
    return RRL_OK;
}

void net_free(NetState &net)
{
    cudaFree(net.params), cudaFree(net.grad), cudaFree(net.m), cudaFree(net.v);
    cudaFree(net.partials64), cudaFree(net.loss_partials), cudaFree(net.loss_out), cudaFree(net.grad64), cudaFree(net.fused_barrier);
    net = NetState{};
}

static size_t esz(int dtype) { return dtype == RRL_F64 ? 8 : 4; }

// Region offsets inside a staging slot sized for `R` rows of the widest wire types.
struct SlotMap {
    size_t obs, act, mask, rew, val, logp, total;
    size_t dev_obs, dev_mask, dev_total;
};
static SlotMap slot_map(int64_t R, int d, int A)
{
    SlotMap m;
    size_t o = 0;
    m.obs = o, o += (size_t)R * d * 8;
    m.mask = o, o += (size_t)R * A * 8;
    m.act = o, o += (size_t)R * 4;
    m.rew = o, o += (size_t)R * 4;
    m.val = o, o += (size_t)R * 4;
    m.logp = o, o += (size_t)R * 4;
    m.total = o;
    m.dev_obs = 0, m.dev_mask = (size_t)R * d * 8, m.dev_total = (size_t)R * (d + A) * 8;
    return m;
}

// Enqueue the H2D copies (and f64->f32 conversions) of the rows packed into the current slot.
int32_t learner_flush_staging(rrl_learner *h)
{
    if (h->staged == 0) return RRL_OK;
    const int d = h->cfg.obs_dim, A = h->cfg.act_dim;
    StagingSlot &s = h->slot[h->cur_slot];
    const SlotMap m = slot_map(h->slot_rows, d, A);
    const int64_t n = h->staged, base = h->ptr - h->staged;
    cudaStream_t st = h->stream;
    if (h->staged_obs_dtype == RRL_F64) {
        RRL_CUDA_OK(cudaMemcpyAsync(s.dev + m.dev_obs, s.host + m.obs, (size_t)n * d * 8, cudaMemcpyHostToDevice, st), h->err);
        RRL_TRY(cvt_f64_f32_launch(reinterpret_cast<const double *>(s.dev + m.dev_obs), h->obs + base * d, n * d, st, h->err));
        h->launches++;
    } else {
        RRL_CUDA_OK(cudaMemcpyAsync(h->obs + base * d, s.host + m.obs, (size_t)n * d * 4, cudaMemcpyHostToDevice, st), h->err);
    }
    if (h->staged_mask_dtype == RRL_F64) {
        RRL_CUDA_OK(cudaMemcpyAsync(s.dev + m.dev_mask, s.host + m.mask, (size_t)n * A * 8, cudaMemcpyHostToDevice, st), h->err);
        RRL_TRY(cvt_f64_f32_launch(reinterpret_cast<const double *>(s.dev + m.dev_mask), h->mask + base * A, n * A, st, h->err));
        h->launches++;
    } else {
        RRL_CUDA_OK(cudaMemcpyAsync(h->mask + base * A, s.host + m.mask, (size_t)n * A * 4, cudaMemcpyHostToDevice, st), h->err);
    }
    RRL_CUDA_OK(cudaMemcpyAsync(h->act + base, s.host + m.act, (size_t)n * 4, cudaMemcpyHostToDevice, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->rew + base, s.host + m.rew, (size_t)n * 4, cudaMemcpyHostToDevice, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->logp + base, s.host + m.logp, (size_t)n * 4, cudaMemcpyHostToDevice, st), h->err);
    if (h->cfg.with_baseline)
        RRL_CUDA_OK(cudaMemcpyAsync(h->val + base, s.host + m.val, (size_t)n * 4, cudaMemcpyHostToDevice, st), h->err);
    RRL_CUDA_OK(cudaEventRecord(s.done, st), h->err);
    s.in_flight = true;
    h->staged = 0;
    h->cur_slot ^= 1;
    StagingSlot &nx = h->slot[h->cur_slot];
    if (nx.in_flight) {  // the other slot's copies must have left the pinned buffer before it is refilled
        RRL_CUDA_OK(cudaEventSynchronize(nx.done), h->err);
        nx.in_flight = false;
    }
    return RRL_OK;
}

int32_t learner_mark(rrl_learner *h, LearnerMark &m)
{
    RRL_TRY(learner_flush_staging(h));  // nothing half-staged below the mark
    m.ptr = h->ptr, m.n_offsets = h->offsets.size(), m.n_last_vals = h->last_vals.size(), m.finished = h->finished;
    return RRL_OK;
}

void learner_rollback(rrl_learner *h, const LearnerMark &m)
{
    learner_flush_staging(h);  // rows staged since the mark land beyond it, where the buffer's contents mean nothing
    h->ptr = m.ptr, h->offsets.resize(m.n_offsets), h->last_vals.resize(m.n_last_vals), h->finished = m.finished;
}

// Pack rows [i, j) of the batch (no terminal markers among them) into staging.
static int32_t stage_rows(rrl_learner *h, const rrl_batch_view_t *b, int64_t i, int64_t j)
{
    const int d = h->cfg.obs_dim, A = h->cfg.act_dim;
    const size_t eo = esz(b->obs_dtype), em = esz(b->mask_dtype);
    if (h->staged > 0 && (h->staged_obs_dtype != b->obs_dtype || h->staged_mask_dtype != b->mask_dtype))
        RRL_TRY(learner_flush_staging(h));
    h->staged_obs_dtype = b->obs_dtype, h->staged_mask_dtype = b->mask_dtype;
    const SlotMap m = slot_map(h->slot_rows, d, A);
    while (i < j) {
        if (h->staged == h->slot_rows) RRL_TRY(learner_flush_staging(h));
        StagingSlot &s = h->slot[h->cur_slot];
        const int64_t k = std::min<int64_t>(j - i, h->slot_rows - h->staged), at = h->staged;
        memcpy(s.host + m.obs + (size_t)at * d * eo, static_cast<const char *>(b->obs) + (size_t)i * d * eo, (size_t)k * d * eo);
        memcpy(s.host + m.mask + (size_t)at * A * em, static_cast<const char *>(b->mask) + (size_t)i * A * em, (size_t)k * A * em);
        memcpy(s.host + m.act + (size_t)at * 4, b->act + i, (size_t)k * 4);
        memcpy(s.host + m.rew + (size_t)at * 4, b->rew + i, (size_t)k * 4);
        memcpy(s.host + m.logp + (size_t)at * 4, b->logp + i, (size_t)k * 4);
        if (h->cfg.with_baseline) memcpy(s.host + m.val + (size_t)at * 4, b->val + i, (size_t)k * 4);
        h->staged += k, h->ptr += k, i += k;
    }
    return RRL_OK;
}

// Large f32 runs skip the packing pass: one asynchronous copy per column straight from the caller's
// buffer (truly asynchronous when that buffer is pinned, e.g. from rrl_host_alloc).
static int32_t copy_rows_direct(rrl_learner *h, const rrl_batch_view_t *b, int64_t i, int64_t j)
{
    RRL_TRY(learner_flush_staging(h));
    const int d = h->cfg.obs_dim, A = h->cfg.act_dim;
    const int64_t n = j - i, base = h->ptr;
    cudaStream_t st = h->stream;
    RRL_CUDA_OK(cudaMemcpyAsync(h->obs + base * d, static_cast<const float *>(b->obs) + i * d, (size_t)n * d * 4, cudaMemcpyHostToDevice, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->mask + base * A, static_cast<const float *>(b->mask) + i * A, (size_t)n * A * 4, cudaMemcpyHostToDevice, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->act + base, b->act + i, (size_t)n * 4, cudaMemcpyHostToDevice, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->rew + base, b->rew + i, (size_t)n * 4, cudaMemcpyHostToDevice, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->logp + base, b->logp + i, (size_t)n * 4, cudaMemcpyHostToDevice, st), h->err);
    if (h->cfg.with_baseline)
        RRL_CUDA_OK(cudaMemcpyAsync(h->val + base, b->val + i, (size_t)n * 4, cudaMemcpyHostToDevice, st), h->err);
    h->ptr += n;
    return RRL_OK;
}

static inline int64_t next_marker(const uint8_t *done, int64_t i, int64_t n)
{
    while (i < n && (reinterpret_cast<uintptr_t>(done + i) & 7)) {
        if (done[i]) return i;
        ++i;
    }
    while (i + 8 <= n) {
        uint64_t w;
        memcpy(&w, done + i, 8);
        if (w) break;
        i += 8;
    }
    while (i < n && !done[i]) ++i;
    return i;
}

int32_t learner_finish(rrl_learner *h)
{
    RRL_TRY(learner_flush_staging(h));
    const int64_t n_ep = (int64_t)h->offsets.size() - 1;
    cudaStream_t st = h->stream;
    if (n_ep + 1 > h->d_offsets_cap) {
        int64_t cap = std::max<int64_t>(1024, 2 * (n_ep + 1));
        cudaFree(h->d_offsets), cudaFree(h->d_last_vals);
        h->d_offsets = nullptr, h->d_last_vals = nullptr, h->d_offsets_cap = 0;
        RRL_CUDA_OK(cudaMalloc(reinterpret_cast<void **>(&h->d_offsets), cap * sizeof(int64_t)), h->err);
        RRL_CUDA_OK(cudaMalloc(reinterpret_cast<void **>(&h->d_last_vals), cap * sizeof(float)), h->err);
        h->d_offsets_cap = cap;
    }
    RRL_CUDA_OK(cudaMemcpyAsync(h->d_offsets, h->offsets.data(), (n_ep + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st), h->err);
    if (n_ep > 0)
        RRL_CUDA_OK(cudaMemcpyAsync(h->d_last_vals, h->last_vals.data(), n_ep * sizeof(float), cudaMemcpyHostToDevice, st), h->err);
    RRL_CUDA_OK(cudaEventRecord(h->ev[0], st), h->err);
    const bool single = h->cfg.nranks <= 1;
    RRL_TRY(scan_launch(h->rew, h->val, h->d_offsets, h->d_last_vals, h->ptr, n_ep, h->cfg.gamma, h->cfg.lam,
                        h->cfg.with_baseline != 0, h->adv, h->ret, h->stats, h->scan_ws, h->scan_ws_bytes, single,
                        h->offsets.back(), st, h->err, &h->launches));
    h->n_global = h->ptr, h->ep_global = n_ep;
    h->high_water = std::max(h->high_water, h->ptr);
    if (!single) {
        // (sum, sumsq, n, n_ep) of every rank, merged in rank order on every rank => identical mean/std
        const double nep = (double)n_ep;
        RRL_CUDA_OK(cudaMemcpyAsync(h->stats + 3, &nep, sizeof(double), cudaMemcpyHostToDevice, st), h->err);
        RRL_CUDA_OK(cudaStreamSynchronize(st), h->err);  // `nep` is a stack variable
        int rc = h->nccl->AllGather(h->stats, h->stats_gather, 4, NCCL_FLOAT64, h->comm, st);
        if (rc != 0) {
            h->err.set("ncclAllGather(stats) failed: %s", h->nccl->GetErrorString(rc));
            return RRL_ENCCL;
        }
        RRL_TRY(stats_merge_launch(h->stats_gather, h->cfg.nranks, h->stats, st, h->err));
        h->launches++;
        double tot[6];
        RRL_CUDA_OK(cudaMemcpyAsync(tot, h->stats, sizeof tot, cudaMemcpyDeviceToHost, st), h->err);
        RRL_CUDA_OK(cudaStreamSynchronize(st), h->err);
        h->n_global = (int64_t)tot[2], h->ep_global = (int64_t)tot[5];
    }
    RRL_CUDA_OK(cudaEventRecord(h->ev[1], st), h->err);
    h->finished = true;
    return RRL_OK;
}

}  // namespace rrl

using namespace rrl;

#define H_CHECK(h)                       \
    if (!(h)) return RRL_EINVAL;         \
    if (cudaSetDevice((h)->cfg.device) != cudaSuccess) { (h)->err.set("cudaSetDevice(%d) failed", (h)->cfg.device); return RRL_ECUDA; }

extern "C" {

int32_t rrl_abi_version(void) { return RRL_ABI_VERSION; }

const char *rrl_status_string(int32_t s)
{
    switch (s) {
        case RRL_OK: return "ok";
        case RRL_EINVAL: return "invalid argument";
        case RRL_ECUDA: return "CUDA error";
        case RRL_ECAPACITY: return "replay buffer capacity exceeded";
        case RRL_ESTATE: return "invalid call order";
        case RRL_ENCCL: return "NCCL error";
        case RRL_EFORMAT: return "malformed wire data";
        case RRL_ENOMEM: return "out of host memory";
        default: return "unknown status";
    }
}

const char *rrl_last_error(const rrl_learner_t *h) { return h ? h->err.msg.c_str() : g_create_err.msg.c_str(); }

int32_t rrl_device_count(int32_t *count)
{
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) {
        if (count) *count = 0;
        g_create_err.set("cudaGetDeviceCount failed: no CUDA device / driver");
        return RRL_ECUDA;
    }
    if (count) *count = c;
    return RRL_OK;
}

int32_t rrl_host_alloc(void **ptr, size_t bytes)
{
    if (!ptr) return RRL_EINVAL;
    if (cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
        g_create_err.set("cudaHostAlloc(%zu) failed", bytes);
        return RRL_ECUDA;
    }
    return RRL_OK;
}

int32_t rrl_host_free(void *ptr) { return cudaFreeHost(ptr) == cudaSuccess ? RRL_OK : RRL_ECUDA; }

int32_t rrl_nccl_unique_id(uint8_t out[RRL_NCCL_ID_BYTES])
{
    const char *why = nullptr;
    NcclApi *api = nccl_api(&why);
    if (!api) {
        g_create_err.set("%s", why);
        return RRL_ENCCL;
    }
    NcclUniqueId id;
    int rc = api->GetUniqueId(&id);
    if (rc != 0) {
        g_create_err.set("ncclGetUniqueId failed: %s", api->GetErrorString(rc));
        return RRL_ENCCL;
    }
    memcpy(out, id.internal, RRL_NCCL_ID_BYTES);
    return RRL_OK;
}

void rrl_config_defaults(rrl_config_t *c)
{
    memset(c, 0, sizeof *c);
    c->obs_dim = 4, c->act_dim = 2, c->hidden = 128;  // REINFORCE.py:46,49
    c->with_baseline = 1, c->train_vf_iters = 80;     // config_loader.rs:68-78
    c->logp_mode = RRL_LOGP_REFERENCE;
    c->nranks = 1;
    c->capacity = 1 << 20;
    c->gamma = (double)0.98f, c->lam = (double)0.97f;  // f32-rounded (config_loader.rs:148-159)
    c->pi_lr = (double)3e-4f, c->vf_lr = (double)1e-3f;
    c->adam_beta1 = 0.9, c->adam_beta2 = 0.999, c->adam_eps = 1e-8;
}

int32_t rrl_learner_destroy(rrl_learner_t *h)
{
    if (!h) return RRL_OK;
    cudaSetDevice(h->cfg.device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->comm && h->nccl) h->nccl->CommDestroy(h->comm);
    cudaFree(h->obs), cudaFree(h->act), cudaFree(h->mask), cudaFree(h->rew), cudaFree(h->val), cudaFree(h->logp);
    cudaFree(h->adv), cudaFree(h->ret), cudaFree(h->tmp), cudaFree(h->d_offsets), cudaFree(h->d_last_vals);
    cudaFree(h->scan_ws), cudaFree(h->stats), cudaFree(h->stats_gather), cudaFree(h->loss_sync);
    for (auto &s : h->slot) {
        if (s.host) cudaFreeHost(s.host);
        cudaFree(s.dev);
        if (s.done) cudaEventDestroy(s.done);
    }
    net_free(h->pi), net_free(h->vf);
    for (auto &e : h->ev)
        if (e) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return RRL_OK;
}

int32_t rrl_learner_create(const rrl_config_t *cfg, rrl_learner_t **out)
{
    if (!cfg || !out) return RRL_EINVAL;
    *out = nullptr;
    ErrBuf &ce = g_create_err;
    if (cfg->obs_dim < 1 || cfg->act_dim < 1 || cfg->capacity < 1 || cfg->nranks < 1 || cfg->rank < 0 ||
        cfg->rank >= cfg->nranks || cfg->train_vf_iters < 0) {
        ce.set("rrl_learner_create: bad config (obs_dim=%d act_dim=%d capacity=%lld rank=%d/%d)", cfg->obs_dim,
               cfg->act_dim, (long long)cfg->capacity, cfg->rank, cfg->nranks);
        return RRL_EINVAL;
    }
    RRL_TRY(mlp_check_dims(cfg->obs_dim, cfg->hidden, cfg->act_dim, ce));
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || cfg->device < 0 || cfg->device >= ndev) {
        ce.set("rrl_learner_create: CUDA device %d unavailable (%d visible); this library has no CPU path", cfg->device, ndev);
        return RRL_ECUDA;
    }
    RRL_CUDA_OK(cudaSetDevice(cfg->device), ce);
    rrl_learner *h = new rrl_learner();
    h->cfg = *cfg;
    auto fail = [&](int32_t st) {
        ce.msg = h->err.msg;
        rrl_learner_destroy(h);
        return st;
    };
#define C_TRY(expr)                            \
    do {                                       \
        int32_t _s = (expr);                   \
        if (_s != RRL_OK) return fail(_s);     \
    } while (0)
#define C_CUDA(expr)                                                                    \
    do {                                                                                \
        cudaError_t _e = (expr);                                                        \
        if (_e != cudaSuccess) {                                                        \
            h->err.set("%s failed: %s", #expr, cudaGetErrorString(_e));                 \
            return fail(RRL_ECUDA);                                                     \
        }                                                                               \
    } while (0)
    cudaDeviceGetAttribute(&h->sms, cudaDevAttrMultiProcessorCount, cfg->device);
    C_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    for (auto &e : h->ev) C_CUDA(cudaEventCreate(&e));
    const size_t cap = (size_t)cfg->capacity, d = cfg->obs_dim, A = cfg->act_dim;
    // adv/ret start at zero and are never re-zeroed, like the reference's numpy buffers
    C_TRY(dmalloc(&h->obs, cap * d, h->err));
    C_TRY(dmalloc(&h->act, cap, h->err));
    C_TRY(dmalloc(&h->mask, cap * A, h->err));
    C_TRY(dmalloc(&h->rew, cap, h->err));
    C_TRY(dmalloc(&h->val, cap, h->err));
    C_TRY(dmalloc(&h->logp, cap, h->err));
    C_TRY(dmalloc(&h->adv, cap, h->err));
    C_TRY(dmalloc(&h->ret, cap, h->err));
    C_TRY(dmalloc(&h->tmp, cap, h->err));
    h->scan_ws_bytes = scan_workspace_bytes(cfg->capacity);
    C_CUDA(cudaMalloc(&h->scan_ws, h->scan_ws_bytes));
    C_TRY(dmalloc(&h->stats, 8, h->err));
    C_TRY(dmalloc(&h->stats_gather, 4 * (size_t)cfg->nranks, h->err));
    C_TRY(dmalloc(&h->loss_sync, 8, h->err));
    h->offsets.assign(1, 0);
    // staging: two pinned slots of <= 32 MiB
    const size_t row_bytes = 8 * (d + A) + 16;
    h->slot_rows = (int64_t)std::max<size_t>(1024, std::min<size_t>((32u << 20) / row_bytes, 1u << 20));
    h->slot_rows = std::min<int64_t>(h->slot_rows, std::max<int64_t>(cfg->capacity, 1));
    const SlotMap sm = slot_map(h->slot_rows, (int)d, (int)A);
    for (auto &s : h->slot) {
        C_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&s.host), sm.total, cudaHostAllocDefault));
        C_CUDA(cudaMalloc(reinterpret_cast<void **>(&s.dev), sm.dev_total));
        C_CUDA(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
        s.bytes = sm.total;
    }
    const int grid = mlp_grid_for_device(cfg->device, cfg->hidden);
    C_TRY(net_alloc(h->pi, cfg->obs_dim, cfg->hidden, cfg->act_dim, grid, true, cfg->nranks > 1, h->err));
    if (cfg->with_baseline) C_TRY(net_alloc(h->vf, cfg->obs_dim, cfg->hidden, 1, grid, true, cfg->nranks > 1, h->err));
    if (cfg->nranks > 1) {
        const char *why = nullptr;
        h->nccl = nccl_api(&why);
        if (!h->nccl) {
            h->err.set("nranks=%d needs NCCL: %s", cfg->nranks, why);
            return fail(RRL_ENCCL);
        }
        NcclUniqueId id;
        memcpy(id.internal, cfg->nccl_id, RRL_NCCL_ID_BYTES);
        int rc = h->nccl->CommInitRank(&h->comm, cfg->nranks, id, cfg->rank);
        if (rc != 0) {
            h->err.set("ncclCommInitRank failed: %s", h->nccl->GetErrorString(rc));
            h->comm = nullptr;
            return fail(RRL_ENCCL);
        }
    }
    C_CUDA(cudaDeviceSynchronize());
#undef C_TRY
#undef C_CUDA
    *out = h;
    return RRL_OK;
}

int32_t rrl_learner_stream(const rrl_learner_t *h, void **stream)
{
    if (!h || !stream) return RRL_EINVAL;
    *stream = static_cast<void *>(h->stream);
    return RRL_OK;
}

int32_t rrl_param_count(const rrl_learner_t *h, int32_t which, int64_t *n)
{
    if (!h || !n || which < 0 || which > 1 || (which == 1 && !h->cfg.with_baseline)) return RRL_EINVAL;
    *n = (which == 0 ? h->pi : h->vf).dims.n_params();
    return RRL_OK;
}

static NetState *pick_net(rrl_learner_t *h, int32_t which, int64_t n)
{
    if (which < 0 || which > 1 || (which == 1 && !h->cfg.with_baseline)) {
        h->err.set("which=%d: no such network", which);
        return nullptr;
    }
    NetState *net = which == 0 ? &h->pi : &h->vf;
    if (n != net->dims.n_params()) {
        h->err.set("parameter count %lld != %lld", (long long)n, (long long)net->dims.n_params());
        return nullptr;
    }
    return net;
}

int32_t rrl_set_params(rrl_learner_t *h, int32_t which, const float *flat, int64_t n)
{
    H_CHECK(h);
    NetState *net = pick_net(h, which, n);
    if (!net || !flat) return RRL_EINVAL;
    RRL_CUDA_OK(cudaMemcpyAsync(net->params, flat, n * sizeof(float), cudaMemcpyHostToDevice, h->stream), h->err);
    RRL_CUDA_OK(cudaStreamSynchronize(h->stream), h->err);
    return RRL_OK;
}

int32_t rrl_get_params(rrl_learner_t *h, int32_t which, float *flat, int64_t n)
{
    H_CHECK(h);
    NetState *net = pick_net(h, which, n);
    if (!net || !flat) return RRL_EINVAL;
    RRL_CUDA_OK(cudaMemcpyAsync(flat, net->params, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream), h->err);
    RRL_CUDA_OK(cudaStreamSynchronize(h->stream), h->err);
    return RRL_OK;
}

int32_t rrl_get_opt_state(rrl_learner_t *h, int32_t which, float *m, float *v, int64_t n, int64_t *step)
{
    H_CHECK(h);
    NetState *net = pick_net(h, which, n);
    if (!net) return RRL_EINVAL;
    if (m) RRL_CUDA_OK(cudaMemcpyAsync(m, net->m, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream), h->err);
    if (v) RRL_CUDA_OK(cudaMemcpyAsync(v, net->v, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream), h->err);
    RRL_CUDA_OK(cudaStreamSynchronize(h->stream), h->err);
    if (step) *step = net->step;
    return RRL_OK;
}

int32_t rrl_set_opt_state(rrl_learner_t *h, int32_t which, const float *m, const float *v, int64_t n, int64_t step)
{
    H_CHECK(h);
    NetState *net = pick_net(h, which, n);
    if (!net || !m || !v || step < 0) return RRL_EINVAL;
    RRL_CUDA_OK(cudaMemcpyAsync(net->m, m, n * sizeof(float), cudaMemcpyHostToDevice, h->stream), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(net->v, v, n * sizeof(float), cudaMemcpyHostToDevice, h->stream), h->err);
    RRL_CUDA_OK(cudaStreamSynchronize(h->stream), h->err);
    net->step = step;
    return RRL_OK;
}

int32_t rrl_ingest_packed(rrl_learner_t *h, const rrl_batch_view_t *b)
{
    H_CHECK(h);
    if (!b || b->count < 0 || (b->count > 0 && (!b->obs || !b->act || !b->mask || !b->rew || !b->logp || !b->done)) ||
        (h->cfg.with_baseline && b->count > 0 && !b->val) || (b->obs_dtype != RRL_F32 && b->obs_dtype != RRL_F64) ||
        (b->mask_dtype != RRL_F32 && b->mask_dtype != RRL_F64)) {
        h->err.set("rrl_ingest_packed: null column or bad dtype");
        return RRL_EINVAL;
    }
    h->finished = false;
    int64_t i = 0;
    const int64_t n = b->count;
    while (i < n) {
        if (b->done[i]) {
            // terminal marker: nothing stored; path_start_idx <- ptr; its reward is last_val (REINFORCE.py:86)
            h->offsets.push_back(h->ptr);
            h->last_vals.push_back(b->rew[i]);
            ++i;
            continue;
        }
        const int64_t j = next_marker(b->done, i, n);
        if (h->ptr + (j - i) > h->cfg.capacity) {
            h->err.set("replay buffer full: ptr=%lld + %lld rows > capacity %lld (replay_buffer.py:38)", (long long)h->ptr,
                       (long long)(j - i), (long long)h->cfg.capacity);
            return RRL_ECAPACITY;
        }
        if (j - i >= 65536 && b->obs_dtype == RRL_F32 && b->mask_dtype == RRL_F32)
            RRL_TRY(copy_rows_direct(h, b, i, j));
        else
            RRL_TRY(stage_rows(h, b, i, j));
        i = j;
    }
    return RRL_OK;
}

int32_t rrl_ingest_device(rrl_learner_t *h, const float *obs, const float *act, const float *mask, const float *rew,
                          const float *val, const float *logp, int64_t count, const int64_t *offsets,
                          const float *last_vals, int64_t n_ep)
{
    H_CHECK(h);
    if (count < 0 || n_ep < 0 || (count > 0 && (!obs || !act || !mask || !rew || !logp)) ||
        (h->cfg.with_baseline && count > 0 && !val) || (n_ep > 0 && (!offsets || !last_vals))) {
        h->err.set("rrl_ingest_device: null column");
        return RRL_EINVAL;
    }
    if (n_ep > 0 && (offsets[0] != 0 || offsets[n_ep] > count)) {
        h->err.set("rrl_ingest_device: offsets must start at 0 and end <= count");
        return RRL_EINVAL;
    }
    if (h->ptr + count > h->cfg.capacity) {
        h->err.set("replay buffer full: ptr=%lld + %lld rows > capacity %lld", (long long)h->ptr, (long long)count,
                   (long long)h->cfg.capacity);
        return RRL_ECAPACITY;
    }
    RRL_TRY(learner_flush_staging(h));
    h->finished = false;
    const int d = h->cfg.obs_dim, A = h->cfg.act_dim;
    const int64_t base = h->ptr;
    cudaStream_t st = h->stream;
    // the caller's buffers live on the default stream's timeline: the copies start after what is queued there ...
    RRL_CUDA_OK(order_after_default_stream(st, h->ev[5]), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->obs + base * d, obs, (size_t)count * d * 4, cudaMemcpyDeviceToDevice, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->mask + base * A, mask, (size_t)count * A * 4, cudaMemcpyDeviceToDevice, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->act + base, act, (size_t)count * 4, cudaMemcpyDeviceToDevice, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->rew + base, rew, (size_t)count * 4, cudaMemcpyDeviceToDevice, st), h->err);
    RRL_CUDA_OK(cudaMemcpyAsync(h->logp + base, logp, (size_t)count * 4, cudaMemcpyDeviceToDevice, st), h->err);
    if (h->cfg.with_baseline)
        RRL_CUDA_OK(cudaMemcpyAsync(h->val + base, val, (size_t)count * 4, cudaMemcpyDeviceToDevice, st), h->err);
    // ... and whatever the caller queues there next (a free, a refill of these buffers) waits for the copies
    RRL_CUDA_OK(order_default_stream_after(st, h->ev[6]), h->err);
    for (int64_t e = 1; e <= n_ep; ++e) {
        if (offsets[e] < offsets[e - 1]) {
            h->err.set("rrl_ingest_device: offsets not monotone at %lld", (long long)e);
            return RRL_EINVAL;
        }
        h->offsets.push_back(base + offsets[e]);
        h->last_vals.push_back(last_vals[e - 1]);
    }
    h->ptr += count;
    return RRL_OK;
}

int32_t rrl_finish(rrl_learner_t *h, int64_t *n_transitions, int64_t *n_episodes)
{
    H_CHECK(h);
    RRL_TRY(learner_finish(h));
    if (n_transitions) *n_transitions = h->ptr;
    if (n_episodes) *n_episodes = (int64_t)h->offsets.size() - 1;
    return RRL_OK;
}

int32_t rrl_reset_buffer(rrl_learner_t *h)
{
    H_CHECK(h);
    RRL_TRY(learner_flush_staging(h));
    h->ptr = 0, h->offsets.assign(1, 0), h->last_vals.clear(), h->finished = false;
    return RRL_OK;
}

int32_t rrl_get_offsets(rrl_learner_t *h, int64_t *out, int64_t cap, int64_t *n_written)
{
    if (!h || !out) return RRL_EINVAL;
    const int64_t n = (int64_t)h->offsets.size();
    if (cap < n) {
        h->err.set("rrl_get_offsets: need %lld entries", (long long)n);
        return RRL_EINVAL;
    }
    memcpy(out, h->offsets.data(), n * sizeof(int64_t));
    if (n_written) *n_written = n;
    return RRL_OK;
}

static int32_t tap(rrl_learner_t *h, const float *src, float *out, int64_t cap)
{
    if (!out || cap < h->ptr) {
        h->err.set("output buffer too small: %lld < %lld", (long long)cap, (long long)h->ptr);
        return RRL_EINVAL;
    }
    RRL_CUDA_OK(cudaMemcpyAsync(out, src, h->ptr * sizeof(float), cudaMemcpyDeviceToHost, h->stream), h->err);
    RRL_CUDA_OK(cudaStreamSynchronize(h->stream), h->err);
    return RRL_OK;
}

int32_t rrl_get_returns(rrl_learner_t *h, float *out, int64_t cap)
{
    H_CHECK(h);
    if (!h->finished) RRL_TRY(learner_finish(h));
    return tap(h, h->ret, out, cap);
}

int32_t rrl_get_advantages(rrl_learner_t *h, float *out, int64_t cap, int32_t normalised)
{
    H_CHECK(h);
    if (!h->finished) RRL_TRY(learner_finish(h));
    if (!normalised) return tap(h, h->adv, out, cap);
    RRL_TRY(normalise_launch(h->adv, h->ptr, h->stats, h->tmp, h->stream, h->err));
    h->launches++;
    return tap(h, h->tmp, out, cap);
}

int32_t rrl_get_adv_stats(rrl_learner_t *h, float *mean, float *std)
{
    H_CHECK(h);
    if (!h->finished) RRL_TRY(learner_finish(h));
    double s[5];
    RRL_CUDA_OK(cudaMemcpyAsync(s, h->stats, sizeof s, cudaMemcpyDeviceToHost, h->stream), h->err);
    RRL_CUDA_OK(cudaStreamSynchronize(h->stream), h->err);
    if (mean) *mean = (float)s[3];
    if (std) *std = (float)s[4];
    return RRL_OK;
}

}  // extern "C"

// ---- checkpoint / resume of the training state --------------------------------------
// The reference keeps nothing across a restart: disable/enable/restart_server re-spawn the worker with a fresh
// algorithm (src/network/server/training_zmq.rs:322-465) and only the TorchScript policy is ever written
// (REINFORCE.py:64-68; utils/logger.py:200-269 save_state is never called).  One flat little-endian file:
//   "RRLCKPT1" | i32 obs_dim, act_dim, hidden, with_baseline | u64 user_tag (the caller's epoch counter)
//   per network (policy, then value when with_baseline): i64 n_params, i64 adam_step, f32 params[n], m[n], v[n]
//   i64 high_water, f32 adv[high_water], ret[high_water]
// A checkpoint is taken between epochs, where the replay buffer is empty (ReplayBuffer.get() resets it) -- except for
// one thing the reference's buffer carries across epochs: adv/ret are never re-zeroed, and transitions stored after an
// epoch's last terminal marker train on whatever those slots held before.  So the stale adv/ret contents
// up to the largest fill level seen are part of the state; without them a resumed run with unfinished tails diverges.
namespace {
const char CKPT_MAGIC[8] = {'R', 'R', 'L', 'C', 'K', 'P', 'T', '1'};
}

extern "C" int32_t rrl_checkpoint_save(rrl_learner_t *h, const char *path, uint64_t user_tag)
{
    H_CHECK(h);
    if (!path) return RRL_EINVAL;
    return abi_guard(h->err, "rrl_checkpoint_save", [&]() -> int32_t {
        const std::string tmp = std::string(path) + ".tmp";
        FILE *f = fopen(tmp.c_str(), "wb");
        if (!f) {
            h->err.set("rrl_checkpoint_save: cannot open %s", tmp.c_str());
            return RRL_EINVAL;
        }
        bool ok = fwrite(CKPT_MAGIC, 1, 8, f) == 8;
        const int32_t dims[4] = {h->cfg.obs_dim, h->cfg.act_dim, h->cfg.hidden, h->cfg.with_baseline};
        ok = ok && fwrite(dims, sizeof dims, 1, f) == 1 && fwrite(&user_tag, 8, 1, f) == 1;
        int32_t st = RRL_OK;
        for (int which = 0; which < (h->cfg.with_baseline ? 2 : 1) && ok && st == RRL_OK; ++which) {
            NetState &net = which == 0 ? h->pi : h->vf;
            const int64_t n = net.dims.n_params(), step = net.step;
            std::vector<float> buf((size_t)n * 3);
            st = rrl_get_params(h, which, buf.data(), n);
            if (st == RRL_OK) st = rrl_get_opt_state(h, which, buf.data() + n, buf.data() + 2 * n, n, nullptr);
            ok = ok && fwrite(&n, 8, 1, f) == 1 && fwrite(&step, 8, 1, f) == 1 && fwrite(buf.data(), 4, buf.size(), f) == buf.size();
        }
        if (ok && st == RRL_OK) {
            const int64_t hw = h->high_water;
            std::vector<float> buf((size_t)hw * 2 + 1);
            cudaError_t ce = cudaMemcpyAsync(buf.data(), h->adv, (size_t)hw * 4, cudaMemcpyDeviceToHost, h->stream);
            if (ce == cudaSuccess) ce = cudaMemcpyAsync(buf.data() + hw, h->ret, (size_t)hw * 4, cudaMemcpyDeviceToHost, h->stream);
            if (ce == cudaSuccess) ce = cudaStreamSynchronize(h->stream);
            if (ce != cudaSuccess) {
                h->err.set("rrl_checkpoint_save: reading adv/ret failed: %s", cudaGetErrorString(ce));
                st = RRL_ECUDA;
            }
            ok = fwrite(&hw, 8, 1, f) == 1 && fwrite(buf.data(), 4, (size_t)hw * 2, f) == (size_t)hw * 2;
        }
        ok = (fclose(f) == 0) && ok;
        if (st != RRL_OK || !ok || rename(tmp.c_str(), path) != 0) {  // written beside, renamed over: never a torn file
            remove(tmp.c_str());
            if (st == RRL_OK) h->err.set("rrl_checkpoint_save: write to %s failed", path);
            return st != RRL_OK ? st : RRL_EINVAL;
        }
        return RRL_OK;
    });
}

extern "C" int32_t rrl_checkpoint_load(rrl_learner_t *h, const char *path, uint64_t *user_tag)
{
    H_CHECK(h);
    if (!path) return RRL_EINVAL;
    return abi_guard(h->err, "rrl_checkpoint_load", [&]() -> int32_t {
        FILE *f = fopen(path, "rb");
        if (!f) {
            h->err.set("rrl_checkpoint_load: cannot open %s", path);
            return RRL_EINVAL;
        }
        char magic[8];
        int32_t dims[4];
        uint64_t tag = 0;
        bool ok = fread(magic, 1, 8, f) == 8 && memcmp(magic, CKPT_MAGIC, 8) == 0 && fread(dims, sizeof dims, 1, f) == 1 &&
                  fread(&tag, 8, 1, f) == 1;
        if (ok && (dims[0] != h->cfg.obs_dim || dims[1] != h->cfg.act_dim || dims[2] != h->cfg.hidden ||
                   (dims[3] != 0) != (h->cfg.with_baseline != 0))) {
            fclose(f);
            h->err.set("rrl_checkpoint_load: %s holds obs_dim=%d act_dim=%d hidden=%d baseline=%d, this learner %d/%d/%d/%d", path,
                       dims[0], dims[1], dims[2], dims[3], h->cfg.obs_dim, h->cfg.act_dim, h->cfg.hidden, h->cfg.with_baseline);
            return RRL_EINVAL;
        }
        // read everything first, apply only when the whole file is sound
        std::vector<float> bufs[2];
        int64_t steps[2] = {0, 0};
        const int nets = h->cfg.with_baseline ? 2 : 1;
        for (int which = 0; which < nets && ok; ++which) {
            const int64_t want = (which == 0 ? h->pi : h->vf).dims.n_params();
            int64_t n = 0;
            ok = fread(&n, 8, 1, f) == 1 && fread(&steps[which], 8, 1, f) == 1 && n == want && steps[which] >= 0;
            if (ok) {
                bufs[which].resize((size_t)n * 3);
                ok = fread(bufs[which].data(), 4, bufs[which].size(), f) == bufs[which].size();
            }
        }
        int64_t hw = 0;
        std::vector<float> stale;
        ok = ok && fread(&hw, 8, 1, f) == 1 && hw >= 0 && hw <= h->cfg.capacity;
        if (ok) {
            stale.resize((size_t)hw * 2 + 1);
            ok = fread(stale.data(), 4, (size_t)hw * 2, f) == (size_t)hw * 2;
        }
        char extra;
        ok = ok && fread(&extra, 1, 1, f) == 0;  // nothing may follow
        fclose(f);
        if (!ok) {
            h->err.set("rrl_checkpoint_load: %s is not a checkpoint of this library (bad magic, sizes or truncated)", path);
            return RRL_EFORMAT;
        }
        for (int which = 0; which < nets; ++which) {
            const int64_t n = (int64_t)bufs[which].size() / 3;
            RRL_TRY(rrl_set_params(h, which, bufs[which].data(), n));
            RRL_TRY(rrl_set_opt_state(h, which, bufs[which].data() + n, bufs[which].data() + 2 * n, n, steps[which]));
        }
        RRL_CUDA_OK(cudaMemcpyAsync(h->adv, stale.data(), (size_t)hw * 4, cudaMemcpyHostToDevice, h->stream), h->err);
        RRL_CUDA_OK(cudaMemcpyAsync(h->ret, stale.data() + hw, (size_t)hw * 4, cudaMemcpyHostToDevice, h->stream), h->err);
        RRL_CUDA_OK(cudaStreamSynchronize(h->stream), h->err);
        h->high_water = hw;
        if (user_tag) *user_tag = tag;
        return RRL_OK;
    });
}
