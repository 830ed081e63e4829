// Wire-format ingest: one RelayRLAction whose tensor fields are safetensors blobs, exactly as the
// reference ships them (TensorData.data, relayrl_framework/src/types/action.rs:196-201, encoded by
// safetensors::serialize in action.rs:345,372-381; the proto fields obs/action/mask carry the same
// blob bytes, src/sys_utils/grpc_utils.rs:31-66).
//
// safetensors 0.5 layout: u64 little-endian header length | JSON header (space padded) | payload,
// header = {"<name>":{"dtype":"F64","shape":[4],"data_offsets":[0,32]}} with one tensor.
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "learner.cuh"

namespace rrl {

struct BlobView {
    int dtype;  // 0 F32, 1 F64, 2 I64, 3 I32, 4 U8/BOOL
    const uint8_t *data;
    size_t nbytes;
    size_t count;
};

static const char *find_key(const char *hdr, size_t n, const char *key)
{
    const size_t kl = strlen(key);
    for (size_t i = 0; i + kl <= n; ++i)
        if (memcmp(hdr + i, key, kl) == 0) return hdr + i + kl;
    return nullptr;
}

static int32_t parse_blob(const uint8_t *blob, size_t len, BlobView &out, ErrBuf &err)
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
    if (!dt || !off) {
        err.set("safetensors header lacks dtype / data_offsets");
        return RRL_EFORMAT;
    }
    size_t esz;
    if (!strncmp(dt, "F64", 3)) out.dtype = 1, esz = 8;
    else if (!strncmp(dt, "F32", 3)) out.dtype = 0, esz = 4;
    else if (!strncmp(dt, "I64", 3)) out.dtype = 2, esz = 8;
    else if (!strncmp(dt, "I32", 3)) out.dtype = 3, esz = 4;
    else if (!strncmp(dt, "U8", 2) || !strncmp(dt, "BOOL", 4)) out.dtype = 4, esz = 1;
    else {
        err.set("safetensors dtype %.4s unsupported", dt);
        return RRL_EFORMAT;
    }
    char *p = nullptr;
    const unsigned long long b0 = strtoull(off, &p, 10);
    if (!p || p >= end || *p != ',') {
        err.set("safetensors data_offsets malformed");
        return RRL_EFORMAT;
    }
    const unsigned long long b1 = strtoull(p + 1, &p, 10);
    const size_t payload = len - 8 - hl;
    if (b1 < b0 || b1 > payload || (b1 - b0) % esz) {
        err.set("safetensors data_offsets [%llu,%llu] outside payload of %zu bytes", b0, b1, payload);
        return RRL_EFORMAT;
    }
    out.data = blob + 8 + hl + b0;
    out.nbytes = b1 - b0;
    out.count = out.nbytes / esz;
    return RRL_OK;
}

static float blob_at(const BlobView &b, size_t i)
{
    switch (b.dtype) {
        case 0: { float v; memcpy(&v, b.data + 4 * i, 4); return v; }
        case 1: { double v; memcpy(&v, b.data + 8 * i, 8); return (float)v; }  // f64 -> f32 as numpy's store does
        case 2: { int64_t v; memcpy(&v, b.data + 8 * i, 8); return (float)v; }
        case 3: { int32_t v; memcpy(&v, b.data + 4 * i, 4); return (float)v; }
        default: return (float)b.data[i];
    }
}

}  // namespace rrl

using namespace rrl;

extern "C" int32_t rrl_ingest_wire_action(rrl_learner_t *h, const uint8_t *obs_blob, size_t obs_len,
                                          const uint8_t *act_blob, size_t act_len, const uint8_t *mask_blob,
                                          size_t mask_len, float rew, const uint8_t *v_blob, size_t v_len,
                                          const uint8_t *logp_blob, size_t logp_len, uint8_t done)
{
    if (!h) return RRL_EINVAL;
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
}
