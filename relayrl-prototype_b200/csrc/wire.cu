// Wire-format ingest: one RelayRLAction whose tensor fields are safetensors blobs, exactly as the
// reference ships them (TensorData.data, relayrl_framework/src/types/action.rs:196-201, encoded by
// safetensors::serialize in action.rs:345,372-381; the proto fields obs/action/mask carry the same
// blob bytes, src/sys_utils/grpc_utils.rs:31-66).
//
// safetensors 0.5 layout: u64 little-endian header length | JSON header (space padded) | payload,
// header = {"<name>":{"dtype":"F64","shape":[4],"data_offsets":[0,32]}} with one tensor.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <thread>
#include <vector>

#include "wire.h"

namespace rrl {

// bounded helpers: the header is network bytes without a terminator
static const char *find_key(const char *hdr, size_t n, const char *key)
{
    const size_t kl = strlen(key);
    for (size_t i = 0; i + kl <= n; ++i)
        if (memcmp(hdr + i, key, kl) == 0) return hdr + i + kl;
    return nullptr;
}
static bool starts_with(const char *p, const char *end, const char *lit)
{
    const size_t n = strlen(lit);
    return (size_t)(end - p) >= n && memcmp(p, lit, n) == 0;
}
// unsigned decimal at p (at least one digit, no overflow past 2^63); advances p
static bool rd_uint(const char *&p, const char *end, unsigned long long &v)
{
    const char *q = p;
    v = 0;
    while (q < end && *q >= '0' && *q <= '9') {
        if (v > (~0ull >> 1) / 10) return false;
        v = v * 10 + (unsigned)(*q++ - '0');
    }
    if (q == p) return false;
    p = q;
    return true;
}

int32_t parse_blob(const uint8_t *blob, size_t len, BlobView &out, ErrBuf &err)
{
    if (!blob || len < 8) {
        err.set("safetensors blob shorter than its length prefix");
        return RRL_EFORMAT;
    }
    uint64_t hl;
    memcpy(&hl, blob, 8);
    if (hl > len - 8 || hl < 2) {
        err.set("safetensors header length %llu exceeds blob (%zu bytes)", (unsigned long long)hl, len);
        return RRL_EFORMAT;
    }
    const char *hdr = reinterpret_cast<const char *>(blob + 8);
    const char *end = hdr + hl;
    const char *dt = find_key(hdr, hl, "\"dtype\":\"");
    const char *off = find_key(hdr, hl, "\"data_offsets\":[");
    const char *shp = find_key(hdr, hl, "\"shape\":[");
    if (!dt || !off || !shp) {
        err.set("safetensors header lacks dtype / shape / data_offsets");
        return RRL_EFORMAT;
    }
    size_t esz;
    if (starts_with(dt, end, "F64\"")) out.dtype = 1, esz = 8;
    else if (starts_with(dt, end, "F32\"")) out.dtype = 0, esz = 4;
    else if (starts_with(dt, end, "I64\"")) out.dtype = 2, esz = 8;
    else if (starts_with(dt, end, "I32\"")) out.dtype = 3, esz = 4;
    else if (starts_with(dt, end, "I16\"")) out.dtype = 5, esz = 2;
    else if (starts_with(dt, end, "U8\"") || starts_with(dt, end, "BOOL\"")) out.dtype = 4, esz = 1;
    else {
        err.set("safetensors dtype %.*s unsupported", (int)std::min<size_t>(4, (size_t)(end - dt)), dt);
        return RRL_EFORMAT;
    }
    unsigned long long b0 = 0, b1 = 0;
    if (!rd_uint(off, end, b0) || off >= end || *off != ',' || !rd_uint(++off, end, b1) || off >= end || *off != ']') {
        err.set("safetensors data_offsets malformed");
        return RRL_EFORMAT;
    }
    // shape: [] (a scalar, one element) or [n0, n1, ...]
    unsigned long long numel = 1;
    while (shp < end && *shp == ' ') ++shp;
    if (shp < end && *shp != ']') {
        while (true) {
            unsigned long long dim = 0;
            while (shp < end && *shp == ' ') ++shp;
            if (!rd_uint(shp, end, dim) || (dim != 0 && numel > (1ull << 40) / dim)) {
                err.set("safetensors shape malformed");
                return RRL_EFORMAT;
            }
            numel *= dim;
            while (shp < end && *shp == ' ') ++shp;
            if (shp < end && *shp == ',') {
                ++shp;
                continue;
            }
            break;
        }
    }
    if (shp >= end || *shp != ']') {
        err.set("safetensors shape malformed");
        return RRL_EFORMAT;
    }
    const size_t payload = len - 8 - hl;
    if (b1 < b0 || b1 > payload || (b1 - b0) % esz) {
        err.set("safetensors data_offsets [%llu,%llu] outside payload of %zu bytes", b0, b1, payload);
        return RRL_EFORMAT;
    }
    if (numel * esz != b1 - b0) {
        err.set("safetensors shape holds %llu elements but data_offsets span %llu bytes", numel, b1 - b0);
        return RRL_EFORMAT;
    }
    out.data = blob + 8 + hl + b0;
    out.nbytes = b1 - b0;
    out.count = out.nbytes / esz;
    return RRL_OK;
}

float blob_at(const BlobView &b, size_t i)
{
    switch (b.dtype) {
        case 0: { float v; memcpy(&v, b.data + 4 * i, 4); return v; }
        case 1: { double v; memcpy(&v, b.data + 8 * i, 8); return (float)v; }  // f64 -> f32 as numpy's store does
        case 2: { int64_t v; memcpy(&v, b.data + 8 * i, 8); return (float)v; }
        case 3: { int32_t v; memcpy(&v, b.data + 4 * i, 4); return (float)v; }
        case 5: { int16_t v; memcpy(&v, b.data + 2 * i, 2); return (float)v; }
        default: return (float)b.data[i];
    }
}


// ---- proto3 wire format (varint / fixed32 / length-delimited), just enough for Trajectory ---------------
// message RelayRLAction { bytes obs = 1; bytes action = 2; bytes mask = 3; float reward = 4;
//                         map<string, bytes> data = 5; bool done = 6; bool reward_update_flag = 7; }
// message Trajectory    { repeated RelayRLAction actions = 1; }        (proto/relayrl_grpc.proto:5-18)
// proto3 leaves out fields that hold their default (empty bytes, 0.0, false).
struct Cur {
    const uint8_t *p, *end;
};
static bool rd_varint(Cur &c, uint64_t &v)
{
    v = 0;
    for (int shift = 0; shift < 64 && c.p < c.end; shift += 7) {
        const uint8_t b = *c.p++;
        v |= (uint64_t)(b & 0x7f) << shift;
        if (!(b & 0x80)) return true;
    }
    return false;
}
static bool rd_len(Cur &c, Cur &sub)
{
    uint64_t n;
    if (!rd_varint(c, n) || n > (uint64_t)(c.end - c.p)) return false;
    sub.p = c.p, sub.end = c.p + n, c.p += n;
    return true;
}
static bool skip_field(Cur &c, unsigned wire_type)
{
    uint64_t v;
    Cur sub;
    switch (wire_type) {
        case 0: return rd_varint(c, v);
        case 1: if (c.end - c.p < 8) return false; c.p += 8; return true;
        case 2: return rd_len(c, sub);
        case 5: if (c.end - c.p < 4) return false; c.p += 4; return true;
        default: return false;  // groups are not used by this schema
    }
}

// One value of the proto `data` map: serde_json of RelayRLData (grpc_utils.rs:45-55), an externally tagged
// enum (action.rs:203-218): {"Tensor":{"shape":[..],"dtype":"Float","data":[<safetensors blob as byte
// values>]}} or a scalar {"Float":x} / {"Double":x} / {"Int":n} / {"Long":n} / {"Short":n} / {"Byte":n}.
static const char *skip_ws(const char *p, const char *end)
{
    while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
    return p;
}
static int32_t parse_data_value(const char *js, size_t n, std::vector<uint8_t> &scratch, float &out, const char *what, ErrBuf &err)
{
    const char *end = js + n;
    const char *p = skip_ws(js, end);
    if (p >= end || *p != '{') {
        err.set("data['%s'] is not a JSON object", what);
        return RRL_EFORMAT;
    }
    p = skip_ws(p + 1, end);
    if (p >= end || *p != '"') {
        err.set("data['%s']: variant tag missing", what);
        return RRL_EFORMAT;
    }
    const char *tag = ++p;
    while (p < end && *p != '"') ++p;
    const size_t tl = (size_t)(p - tag);
    p = skip_ws(p + 1, end);
    if (p >= end || *p != ':') {
        err.set("data['%s']: malformed variant", what);
        return RRL_EFORMAT;
    }
    p = skip_ws(p + 1, end);
    if (tl == 6 && !memcmp(tag, "Tensor", 6)) {
        const char *d = find_key(p, (size_t)(end - p), "\"data\"");
        if (d) d = skip_ws(d, end);
        if (!d || d >= end || *d != ':') {
            err.set("data['%s']: Tensor without a data array", what);
            return RRL_EFORMAT;
        }
        d = skip_ws(d + 1, end);
        if (d >= end || *d != '[') {
            err.set("data['%s']: Tensor.data is not an array", what);
            return RRL_EFORMAT;
        }
        ++d;
        scratch.clear();
        while (true) {
            d = skip_ws(d, end);
            if (d >= end) {
                err.set("data['%s']: unterminated byte array", what);
                return RRL_EFORMAT;
            }
            if (*d == ']') break;
            unsigned v = 0;
            const char *q = d;
            while (q < end && *q >= '0' && *q <= '9') v = v * 10 + (unsigned)(*q++ - '0');
            if (q == d || v > 255) {
                err.set("data['%s']: byte array holds a non-byte value", what);
                return RRL_EFORMAT;
            }
            scratch.push_back((uint8_t)v);
            d = skip_ws(q, end);
            if (d < end && *d == ',') ++d;
        }
        BlobView b;
        RRL_TRY(parse_blob(scratch.data(), scratch.size(), b, err));
        if (b.count < 1) {
            err.set("data['%s'] is an empty tensor", what);
            return RRL_EFORMAT;
        }
        out = blob_at(b, 0);
        return RRL_OK;
    }
    const bool scalar = (tl == 5 && !memcmp(tag, "Float", 5)) || (tl == 6 && !memcmp(tag, "Double", 6)) ||
                        (tl == 3 && !memcmp(tag, "Int", 3)) || (tl == 4 && !memcmp(tag, "Long", 4)) ||
                        (tl == 5 && !memcmp(tag, "Short", 5)) || (tl == 4 && !memcmp(tag, "Byte", 4));
    if (!scalar) {
        err.set("data['%s']: variant %.*s carries no number", what, (int)tl, tag);
        return RRL_EFORMAT;
    }
    char buf[64];
    const size_t m = (size_t)(end - p) < sizeof(buf) - 1 ? (size_t)(end - p) : sizeof(buf) - 1;
    memcpy(buf, p, m);
    buf[m] = 0;
    char *q = nullptr;
    const double v = strtod(buf, &q);
    if (q == buf) {
        err.set("data['%s']: number expected", what);
        return RRL_EFORMAT;
    }
    out = (float)v;
    return RRL_OK;
}


// One action sub-message -> row r of the columns.  Returns RRL_OK or RRL_EFORMAT with `err` set.
static int32_t decode_grpc_action(Cur act, int64_t r, int d, int A, bool with_baseline, const HostColumns &o,
                                  std::vector<uint8_t> &scratch, ErrBuf &err)
{
    Cur f_obs{nullptr, nullptr}, f_act{nullptr, nullptr}, f_mask{nullptr, nullptr}, f_logp{nullptr, nullptr}, f_v{nullptr, nullptr};
    float rew = 0.f;
    bool done = false;
    while (act.p < act.end) {
        uint64_t k;
        if (!rd_varint(act, k)) goto truncated;
        {
            const unsigned field = (unsigned)(k >> 3), wt = (unsigned)(k & 7);
            if (field >= 1 && field <= 3 && wt == 2) {
                Cur b;
                if (!rd_len(act, b)) goto truncated;
                (field == 1 ? f_obs : field == 2 ? f_act : f_mask) = b;
            } else if (field == 4 && wt == 5) {
                if (act.end - act.p < 4) goto truncated;
                memcpy(&rew, act.p, 4);
                act.p += 4;
            } else if (field == 5 && wt == 2) {  // one map entry: { string key = 1; bytes value = 2; }
                Cur e, kstr{nullptr, nullptr}, val{nullptr, nullptr};
                if (!rd_len(act, e)) goto truncated;
                while (e.p < e.end) {
                    uint64_t ek;
                    if (!rd_varint(e, ek)) goto truncated;
                    if ((ek & 7) == 2 && (ek >> 3) == 1) {
                        if (!rd_len(e, kstr)) goto truncated;
                    } else if ((ek & 7) == 2 && (ek >> 3) == 2) {
                        if (!rd_len(e, val)) goto truncated;
                    } else if (!skip_field(e, (unsigned)(ek & 7))) {
                        goto truncated;
                    }
                }
                const size_t kl = (size_t)(kstr.end - kstr.p);
                if (kl == 6 && !memcmp(kstr.p, "logp_a", 6)) f_logp = val;
                else if (kl == 1 && kstr.p[0] == 'v') f_v = val;
            } else if (field == 6 && wt == 0) {
                uint64_t b;
                if (!rd_varint(act, b)) goto truncated;
                done = b != 0;
            } else if (!skip_field(act, wt)) {
                goto truncated;
            }
        }
    }
    o.rew[r] = rew, o.done[r] = done ? 1 : 0;
    o.act[r] = 0.f, o.val[r] = 0.f, o.logp[r] = 0.f;
    for (int i = 0; i < d; ++i) o.obs[r * d + i] = 0.f;
    for (int i = 0; i < A; ++i) o.mask[r * A + i] = 0.f;
    if (!done) {  // a terminal marker has obs = act = mask = data = None (agent_zmq.rs:607-608)
        BlobView bo, ba, bm;
        RRL_TRY(parse_blob(f_obs.p, (size_t)(f_obs.end - f_obs.p), bo, err));
        RRL_TRY(parse_blob(f_act.p, (size_t)(f_act.end - f_act.p), ba, err));
        RRL_TRY(parse_blob(f_mask.p, (size_t)(f_mask.end - f_mask.p), bm, err));
        if ((int)bo.count != d || (int)bm.count != A || ba.count < 1) {
            err.set("action %lld: obs has %zu values (want %d), mask %zu (want %d), act %zu", (long long)r, bo.count, d, bm.count,
                    A, ba.count);
            return RRL_EFORMAT;
        }
        for (int i = 0; i < d; ++i) o.obs[r * d + i] = blob_at(bo, i);
        for (int i = 0; i < A; ++i) o.mask[r * A + i] = blob_at(bm, i);
        o.act[r] = blob_at(ba, 0);
        if (!f_logp.p) {
            err.set("action %lld: data['logp_a'] missing", (long long)r);
            return RRL_EFORMAT;
        }
        RRL_TRY(parse_data_value(reinterpret_cast<const char *>(f_logp.p), (size_t)(f_logp.end - f_logp.p), scratch, o.logp[r],
                                 "logp_a", err));
        if (with_baseline) {
            if (!f_v.p) {
                err.set("action %lld: data['v'] missing", (long long)r);
                return RRL_EFORMAT;
            }
            RRL_TRY(parse_data_value(reinterpret_cast<const char *>(f_v.p), (size_t)(f_v.end - f_v.p), scratch, o.val[r], "v", err));
        }
    }
    return RRL_OK;
truncated:
    err.set("action %lld of the serialized Trajectory is truncated or not proto3 wire format", (long long)r);
    return RRL_EFORMAT;
}

// A serialized Trajectory is decoded in two passes.  index_grpc_actions walks the top-level fields once (cheap: one
// varint pair per action) and returns the action sub-messages -- so the caller knows the row count BEFORE it sizes
// anything.  decode_grpc_actions then decodes the actions -- ~420 B of safetensors blobs and ~500 B of JSON byte arrays
// apiece, the expensive part -- independently into SoA rows (terminal markers become rows with done = 1 that carry
// only their reward, as rrl_batch_view_t expects), on several host threads when the message is large.  On failure the
// error of the lowest-numbered bad action is reported, as a serial decoder would.
static int32_t index_grpc_actions(const uint8_t *msg, size_t len, std::vector<Cur> &actions, ErrBuf &err)
{
    actions.clear();
    Cur top{msg, msg + len};
    while (top.p < top.end) {
        uint64_t key;
        Cur sub;
        const bool ok = rd_varint(top, key) && (key == ((1u << 3) | 2u) ? rd_len(top, sub) : skip_field(top, (unsigned)(key & 7)));
        if (!ok) {
            err.set("serialized Trajectory is truncated or not proto3 wire format (at byte %zu of %zu)", (size_t)(top.p - msg), len);
            return RRL_EFORMAT;
        }
        if (key == ((1u << 3) | 2u)) actions.push_back(sub);
    }
    return RRL_OK;
}

static int32_t decode_grpc_actions(const std::vector<Cur> &actions, int d, int A, bool with_baseline, const HostColumns &o,
                                   int64_t &count, ErrBuf &err)
{
    count = 0;
    const int64_t n = (int64_t)actions.size();
    if (n > o.cap) {
        err.set("trajectory holds %lld actions, more than the %lld rows the caller provided", (long long)n, (long long)o.cap);
        return RRL_ECAPACITY;
    }
    unsigned hw = std::thread::hardware_concurrency();
    const int64_t per_thread_min = 2048;  // below this a thread costs more than it saves
    const int nthreads = (int)std::max<int64_t>(1, std::min<int64_t>({(int64_t)(hw ? hw : 1), 16, n / per_thread_min}));
    std::vector<int32_t> status(nthreads, RRL_OK);
    std::vector<ErrBuf> errs(nthreads);
    auto work = [&](int t) {
        try {
            std::vector<uint8_t> scratch;
            scratch.reserve(256);
            const int64_t lo = n * t / nthreads, hi = n * (t + 1) / nthreads;
            for (int64_t r = lo; r < hi; ++r) {
                const int32_t st = decode_grpc_action(actions[(size_t)r], r, d, A, with_baseline, o, scratch, errs[t]);
                if (st != RRL_OK) {
                    status[t] = st;
                    return;  // later actions of this range are not needed: the call fails
                }
            }
        } catch (...) {  // nothing may escape a worker thread (std::terminate) or the C ABI
            status[t] = RRL_ENOMEM;
            try {
                errs[t].set("gRPC trajectory decode: out of host memory");
            } catch (...) {
            }
        }
    };
    if (nthreads == 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        int started = 1;
        try {
            for (int t = 1; t < nthreads; ++t, ++started) pool.emplace_back(work, t);
        } catch (...) {  // the system refused another thread: the rest of the ranges run here
        }
        work(0);
        for (int t = started; t < nthreads; ++t) work(t);
        for (auto &th : pool) th.join();
    }
    for (int t = 0; t < nthreads; ++t)  // ranges are in action order: the first failing range holds the first bad action
        if (status[t] != RRL_OK) {
            err.msg = errs[t].msg;
            return status[t];
        }
    count = n;
    return RRL_OK;
}

}  // namespace rrl

using namespace rrl;

extern "C" int32_t rrl_ingest_wire_action(rrl_learner_t *h, const uint8_t *obs_blob, size_t obs_len,
                                          const uint8_t *act_blob, size_t act_len, const uint8_t *mask_blob,
                                          size_t mask_len, float rew, const uint8_t *v_blob, size_t v_len,
                                          const uint8_t *logp_blob, size_t logp_len, uint8_t done)
{
    if (!h) return RRL_EINVAL;
    return abi_guard(h->err, "rrl_ingest_wire_action", [&]() -> int32_t {
    rrl_batch_view_t b;
    memset(&b, 0, sizeof b);
    uint8_t dn = done ? 1 : 0;
    float act = 0.f, val = 0.f, logp = 0.f;
    const int d = h->cfg.obs_dim, A = h->cfg.act_dim;
    std::vector<float> obs(d, 0.f), mask(A, 0.f);
    if (!done) {
        // a terminal marker carries obs = act = mask = data = None (agent_zmq.rs:607-608)
        BlobView o, a, m, l, v;
        RRL_TRY(parse_blob(obs_blob, obs_len, o, h->err));
        RRL_TRY(parse_blob(act_blob, act_len, a, h->err));
        RRL_TRY(parse_blob(mask_blob, mask_len, m, h->err));
        RRL_TRY(parse_blob(logp_blob, logp_len, l, h->err));
        if ((int)o.count != d || (int)m.count != A || a.count < 1 || l.count < 1) {
            h->err.set("wire action: obs has %zu values (want %d), mask %zu (want %d), act %zu, logp_a %zu", o.count, d,
                       m.count, A, a.count, l.count);
            return RRL_EFORMAT;
        }
        for (int i = 0; i < d; ++i) obs[i] = blob_at(o, i);
        for (int i = 0; i < A; ++i) mask[i] = blob_at(m, i);
        act = blob_at(a, 0), logp = blob_at(l, 0);
        if (h->cfg.with_baseline) {
            RRL_TRY(parse_blob(v_blob, v_len, v, h->err));
            if (v.count < 1) {
                h->err.set("wire action: data['v'] is empty");
                return RRL_EFORMAT;
            }
            val = blob_at(v, 0);
        }
    }
    b.obs = obs.data(), b.act = &act, b.mask = mask.data(), b.rew = &rew, b.val = &val, b.logp = &logp, b.done = &dn;
    b.count = 1, b.obs_dtype = RRL_F32, b.mask_dtype = RRL_F32;
    return rrl_ingest_packed(h, &b);
    });
}

extern "C" int32_t rrl_decode_grpc_trajectory(const uint8_t *msg, size_t len, int32_t obs_dim, int32_t act_dim, int32_t with_baseline,
                                              float *obs, float *act, float *mask, float *rew, float *val, float *logp, uint8_t *done,
                                              int64_t cap, int64_t *count, char *errbuf, size_t errcap)
{
    if ((!msg && len) || obs_dim < 1 || act_dim < 1 || !obs || !act || !mask || !rew || !val || !logp || !done || cap < 0 || !count)
        return RRL_EINVAL;
    ErrBuf err;
    const HostColumns o{obs, act, mask, rew, val, logp, done, cap};
    const int32_t st = abi_guard(err, "rrl_decode_grpc_trajectory", [&]() -> int32_t {
        std::vector<Cur> actions;
        RRL_TRY(index_grpc_actions(msg, len, actions, err));
        return decode_grpc_actions(actions, obs_dim, act_dim, with_baseline != 0, o, *count, err);
    });
    if (st != RRL_OK && errbuf && errcap) {
        strncpy(errbuf, err.msg.c_str(), errcap - 1);
        errbuf[errcap - 1] = 0;
    }
    return st;
}

extern "C" int32_t rrl_ingest_grpc_trajectory(rrl_learner_t *h, const uint8_t *msg, size_t len, int64_t *n_actions)
{
    if (!h || (!msg && len)) return RRL_EINVAL;
    return abi_guard(h->err, "rrl_ingest_grpc_trajectory", [&]() -> int32_t {
        const int d = h->cfg.obs_dim, A = h->cfg.act_dim;
        // index pass first: the host rows are sized from the action count (never from the byte length), in chunks of
        // at most INGEST_CHUNK_ROWS so that a huge or hostile message cannot ask for gigabytes; a failure in a later
        // chunk rolls the buffer back to where it stood (all of the message is appended, or none of it)
        std::vector<Cur> actions;
        RRL_TRY(index_grpc_actions(msg, len, actions, h->err));
        const int64_t n = (int64_t)actions.size();
        if (n_actions) *n_actions = n;
        if (n == 0) return RRL_OK;
        LearnerMark mark;
        RRL_TRY(learner_mark(h, mark));
        OwnedColumns cols(std::min<int64_t>(n, INGEST_CHUNK_ROWS), d, A);
        std::vector<Cur> part;
        for (int64_t lo = 0; lo < n; lo += INGEST_CHUNK_ROWS) {
            const int64_t hi = std::min<int64_t>(n, lo + INGEST_CHUNK_ROWS);
            part.assign(actions.begin() + lo, actions.begin() + hi);
            int64_t count = 0;
            int32_t st = decode_grpc_actions(part, d, A, h->cfg.with_baseline != 0, cols.view, count, h->err);
            if (st == RRL_OK) {
                rrl_batch_view_t b;
                memset(&b, 0, sizeof b);
                b.obs = cols.view.obs, b.act = cols.view.act, b.mask = cols.view.mask, b.rew = cols.view.rew, b.val = cols.view.val;
                b.logp = cols.view.logp, b.done = cols.view.done, b.count = count, b.obs_dtype = RRL_F32, b.mask_dtype = RRL_F32;
                st = rrl_ingest_packed(h, &b);
            }
            if (st != RRL_OK) {
                const std::string why = h->err.msg;
                learner_rollback(h, mark);
                h->err.msg = why;
                return st;
            }
        }
        return RRL_OK;
    });
}
