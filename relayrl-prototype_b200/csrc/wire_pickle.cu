// serde-pickle ingest: the ZMQ transport's trajectory message, decoded natively into the packed columns.
//
// An agent ships `serde_pickle::to_writer(&Vec<RelayRLAction>, Default::default())`
// (relayrl_framework/src/types/trajectory.rs:50-55) over a PUSH socket and the training server decodes it with
// `pickle::from_slice::<Vec<RelayRLAction>>` (src/network/server/training_zmq.rs:994-1001), re-serialises every blob
// as JSON integers for the Python worker's pipe, which rebuilds numpy arrays one transition at a time.  This file
// replaces all of that: bytes in, SoA rows out.
//
// The encoder is the third-party crate serde-pickle 1.2.0 (Cargo.toml:48), whose sources are not in the reference
// tree, so its exact opcode stream cannot be confirmed here (no Rust toolchain, no network).  What IS fixed is the
// published pickle protocol and serde's data model for these types (src/types/action.rs:92-101,196-218,428-437):
//   Vec<T> -> list; struct -> dict keyed by field name; Option::None -> None; bool, f32 -> float, integers;
//   String -> str; Vec<u8> -> list of ints (serde's default) or bytes (serde_bytes); unit enum variant (DType) ->
//   str, or a 1-tuple under the crate's "compat" representation; newtype variant (RelayRLData::Tensor(..),
//   ::Float(..)) -> 1-entry dict {variant: value}, or a 2-tuple (variant, value).
// So the decoder is a small pickle virtual machine for protocols 2-4 (the opcodes that build None / bool / int /
// float / str / bytes / list / tuple / dict, MARK-batched or one at a time, with the memo) followed by a walk of
// the object graph that accepts every one of those shapes.  Anything else -- an unknown opcode, a global/reduce
// (code execution in a real unpickler), a graph of another shape -- is RRL_EFORMAT.  Checked against Python's own
// pickler on the equivalent graphs (tests/test_wire_pickle.py); a live Rust peer remains unverified.
#include <string.h>

#include <vector>

#include "wire.h"

namespace rrl {

namespace {

enum Kind : uint8_t { K_NONE, K_BOOL, K_INT, K_FLOAT, K_STR, K_BYTES, K_LIST, K_TUPLE, K_DICT, K_MARK };

struct PVal {
    Kind kind = K_NONE;
    union {
        int64_t i;
        double f;
        struct {
            const uint8_t *p;
            uint32_t n;
        } s;           // K_STR / K_BYTES: a view into the message
        uint32_t idx;  // K_LIST / K_TUPLE / K_DICT: index into Graph::nodes
    };
    PVal() : i(0) {}
};

// Containers do not own heap memory: their elements live in two arenas of the graph (one malloc per action was the cost
// of the first version: 25 containers per action).  A container that grows while it is not the last thing in its arena is
// moved to the end first; pickle builds containers in batches, so that is rare.
struct Node {
    uint32_t start = 0, len = 0;    // list / tuple elements, or dict key, value, key, value, ... : Graph::pool[start .. start+len)
    uint32_t bstart = 0, blen = 0;  // a list that so far holds only integers 0..255 (a Vec<u8> under serde's default): Graph::bytes
    bool all_bytes = true;
};

struct Graph {
    std::vector<Node> nodes;
    std::vector<PVal> pool;
    std::vector<uint8_t> bytes;
    std::vector<PVal> stack, memo;
    PVal top;
    const PVal *items(const Node &n) const { return pool.data() + n.start; }
    const uint8_t *byte_data(const Node &n) const { return bytes.data() + n.bstart; }
    void push_item(Node &n, const PVal &v)
    {
        if ((size_t)n.start + n.len != pool.size()) {  // not at the end of the arena: move
            const uint32_t at = (uint32_t)pool.size();
            pool.resize(pool.size() + n.len);
            if (n.len) memmove(pool.data() + at, pool.data() + n.start, n.len * sizeof(PVal));
            n.start = at;
        }
        pool.push_back(v);
        ++n.len;
    }
    uint8_t *grow_bytes(Node &n, size_t cnt)  // room for cnt more bytes of n; returns where to write them
    {
        if ((size_t)n.bstart + n.blen != bytes.size()) {
            const uint32_t at = (uint32_t)bytes.size();
            bytes.resize(bytes.size() + n.blen);
            if (n.blen) memmove(bytes.data() + at, bytes.data() + n.bstart, n.blen);
            n.bstart = at;
        }
        bytes.resize(bytes.size() + cnt);
        uint8_t *dst = bytes.data() + n.bstart + n.blen;
        n.blen += (uint32_t)cnt;
        return dst;
    }
};

bool key_is(const PVal &k, const char *name)
{
    const size_t n = strlen(name);
    return (k.kind == K_STR || k.kind == K_BYTES) && k.s.n == n && memcmp(k.s.p, name, n) == 0;
}

const PVal *dict_get(const Graph &g, const PVal &d, const char *name)
{
    if (d.kind != K_DICT) return nullptr;
    const Node &n = g.nodes[d.idx];
    const PVal *it = g.items(n);
    for (uint32_t i = 0; i + 1 < n.len; i += 2)
        if (key_is(it[i], name)) return &it[i + 1];
    return nullptr;
}

void list_append(Graph &g, Node &n, const PVal &v)
{
    if (n.all_bytes && v.kind == K_INT && v.i >= 0 && v.i <= 255) {
        *g.grow_bytes(n, 1) = (uint8_t)v.i;
        return;
    }
    if (n.all_bytes) {  // first non-byte element: fall back to generic items
        n.all_bytes = false;
        for (uint32_t i = 0; i < n.blen; ++i) {
            PVal x;
            x.kind = K_INT, x.i = g.bytes[n.bstart + i];
            g.push_item(n, x);
        }
        n.blen = 0;
    }
    g.push_item(n, v);
}

// ---- the virtual machine ------------------------------------------------------------------------------------
int32_t run_vm(const uint8_t *msg, size_t len, Graph &g, ErrBuf &err)
{
    const uint8_t *p = msg, *end = msg + len;
    g.nodes.reserve(len / 24 + 16), g.pool.reserve(len / 8 + 16), g.bytes.reserve(len / 2 + 16), g.stack.reserve(1024);
    auto need = [&](size_t n) { return (size_t)(end - p) >= n; };
    auto bad = [&](const char *what) {
        err.set("pickle stream: %s at byte %zu of %zu", what, (size_t)(p - msg), len);
        return RRL_EFORMAT;
    };
    auto rd_u32 = [&](uint32_t &v) {
        if (!need(4)) return false;
        memcpy(&v, p, 4), p += 4;
        return true;
    };
    auto new_node = [&](Kind k) {
        g.nodes.emplace_back();
        PVal v;
        v.kind = k, v.idx = (uint32_t)(g.nodes.size() - 1);
        return v;
    };
    auto find_mark = [&](size_t &at) {
        for (size_t i = g.stack.size(); i-- > 0;)
            if (g.stack[i].kind == K_MARK) {
                at = i;
                return true;
            }
        return false;
    };
    auto push_view = [&](Kind k, uint64_t n) {
        if (n > (uint64_t)(end - p) || n > 0xffffffffull) return false;
        PVal v;
        v.kind = k, v.s.p = p, v.s.n = (uint32_t)n;
        g.stack.push_back(v);
        p += n;
        return true;
    };
    while (p < end) {
        const uint8_t op = *p++;
        switch (op) {
            case 0x80:  // PROTO
                if (!need(1)) return bad("truncated PROTO");
                if (*p < 2 || *p > 5) return bad("unsupported pickle protocol");
                ++p;
                break;
            case 0x95:  // FRAME (protocol 4): a length hint only
                if (!need(8)) return bad("truncated FRAME");
                p += 8;
                break;
            case '.':  // STOP
                if (g.stack.size() != 1 || g.stack[0].kind == K_MARK) return bad("STOP with a stack that does not hold one object");
                g.top = g.stack[0];
                return RRL_OK;
            case 'N': g.stack.emplace_back(); break;
            case 0x88:
            case 0x89: {
                PVal v;
                v.kind = K_BOOL, v.i = op == 0x88;
                g.stack.push_back(v);
                break;
            }
            case 'K':
            case 'M':
            case 'J': {
                const size_t n = op == 'K' ? 1 : op == 'M' ? 2 : 4;
                if (!need(n)) return bad("truncated integer");
                PVal v;
                v.kind = K_INT;
                if (op == 'K') v.i = p[0];
                else if (op == 'M') v.i = p[0] | (p[1] << 8);
                else {
                    int32_t x;
                    memcpy(&x, p, 4);
                    v.i = x;
                }
                p += n;
                g.stack.push_back(v);
                break;
            }
            case 0x8a: {  // LONG1: little-endian two's complement
                if (!need(1)) return bad("truncated LONG1");
                const size_t n = *p++;
                if (n > 8 || !need(n)) return bad("LONG1 wider than 64 bits");
                uint64_t u = 0;
                for (size_t i = 0; i < n; ++i) u |= (uint64_t)p[i] << (8 * i);
                if (n > 0 && n < 8 && (p[n - 1] & 0x80)) u |= ~0ull << (8 * n);
                p += n;
                PVal v;
                v.kind = K_INT, v.i = (int64_t)u;
                g.stack.push_back(v);
                break;
            }
            case 'G': {  // BINFLOAT: big-endian IEEE double
                if (!need(8)) return bad("truncated BINFLOAT");
                uint64_t u = 0;
                for (int i = 0; i < 8; ++i) u = (u << 8) | p[i];
                p += 8;
                PVal v;
                v.kind = K_FLOAT;
                memcpy(&v.f, &u, 8);
                g.stack.push_back(v);
                break;
            }
            case 'X':
            case 'B':
            case 'T': {  // BINUNICODE / BINBYTES / BINSTRING: u32 length
                uint32_t n;
                if (!rd_u32(n) || !push_view(op == 'X' ? K_STR : K_BYTES, n)) return bad("truncated string / bytes");
                break;
            }
            case 0x8c:
            case 'C':
            case 'U': {  // SHORT_BINUNICODE / SHORT_BINBYTES / SHORT_BINSTRING: u8 length
                if (!need(1)) return bad("truncated short string");
                const uint8_t n = *p++;
                if (!push_view(op == 0x8c ? K_STR : K_BYTES, n)) return bad("truncated short string");
                break;
            }
            case 0x8d:
            case 0x8e: {  // BINUNICODE8 / BINBYTES8
                if (!need(8)) return bad("truncated 8-byte length");
                uint64_t n;
                memcpy(&n, p, 8), p += 8;
                if (!push_view(op == 0x8d ? K_STR : K_BYTES, n)) return bad("truncated long string / bytes");
                break;
            }
            case ']': g.stack.push_back(new_node(K_LIST)); break;
            case '}': g.stack.push_back(new_node(K_DICT)); break;
            case ')': g.stack.push_back(new_node(K_TUPLE)); break;
            case '(': {
                // fast path: a batch of one-byte integers appended to a byte list, "( K b K b ... e" -- the body of every
                // Vec<u8> under serde's default representation, i.e. most of the message
                if (!g.stack.empty() && g.stack.back().kind == K_LIST && g.nodes[g.stack.back().idx].all_bytes &&
                    g.nodes[g.stack.back().idx].len == 0) {
                    const uint8_t *q = p;
                    while (end - q >= 2 && q[0] == 'K') q += 2;
                    if (q < end && *q == 'e' && q > p) {
                        const size_t cnt = (size_t)(q - p) / 2;
                        uint8_t *dst = g.grow_bytes(g.nodes[g.stack.back().idx], cnt);
                        for (size_t i = 0; i < cnt; ++i) dst[i] = p[2 * i + 1];
                        p = q + 1;
                        break;
                    }
                }
                PVal v;
                v.kind = K_MARK;
                g.stack.push_back(v);
                break;
            }
            case 'a': {  // APPEND
                if (g.stack.size() < 2 || g.stack[g.stack.size() - 2].kind != K_LIST) return bad("APPEND without a list");
                const PVal v = g.stack.back();
                g.stack.pop_back();
                if (v.kind == K_MARK) return bad("APPEND of a mark");
                list_append(g, g.nodes[g.stack.back().idx], v);
                break;
            }
            case 'e': {  // APPENDS
                size_t m;
                if (!find_mark(m) || m == 0 || g.stack[m - 1].kind != K_LIST) return bad("APPENDS without mark / list");
                Node &n = g.nodes[g.stack[m - 1].idx];
                for (size_t i = m + 1; i < g.stack.size(); ++i) list_append(g, n, g.stack[i]);
                g.stack.resize(m);
                break;
            }
            case 's': {  // SETITEM
                if (g.stack.size() < 3 || g.stack[g.stack.size() - 3].kind != K_DICT) return bad("SETITEM without a dict");
                const PVal v = g.stack.back(), k = g.stack[g.stack.size() - 2];
                if (v.kind == K_MARK || k.kind == K_MARK) return bad("SETITEM of a mark");
                g.stack.resize(g.stack.size() - 2);
                Node &n = g.nodes[g.stack.back().idx];
                g.push_item(n, k), g.push_item(n, v);
                break;
            }
            case 'u': {  // SETITEMS
                size_t m;
                if (!find_mark(m) || m == 0 || g.stack[m - 1].kind != K_DICT || ((g.stack.size() - m - 1) & 1))
                    return bad("SETITEMS without mark / dict / pairs");
                Node &n = g.nodes[g.stack[m - 1].idx];
                for (size_t i = m + 1; i < g.stack.size(); ++i) g.push_item(n, g.stack[i]);
                g.stack.resize(m);
                break;
            }
            case 't':
            case 'l':
            case 'd': {  // TUPLE / LIST / DICT from the topmost mark
                size_t m;
                if (!find_mark(m)) return bad("container opcode without a mark");
                if (op == 'd' && ((g.stack.size() - m - 1) & 1)) return bad("DICT with an odd number of items");
                PVal c = new_node(op == 't' ? K_TUPLE : op == 'l' ? K_LIST : K_DICT);
                Node &n = g.nodes[c.idx];
                if (op == 'l')
                    for (size_t i = m + 1; i < g.stack.size(); ++i) list_append(g, n, g.stack[i]);
                else
                    for (size_t i = m + 1; i < g.stack.size(); ++i) g.push_item(n, g.stack[i]);
                g.stack.resize(m);
                g.stack.push_back(c);
                break;
            }
            case 0x85:
            case 0x86:
            case 0x87: {  // TUPLE1 / TUPLE2 / TUPLE3
                const size_t k = op - 0x84;
                if (g.stack.size() < k) return bad("TUPLEn on a short stack");
                for (size_t i = g.stack.size() - k; i < g.stack.size(); ++i)
                    if (g.stack[i].kind == K_MARK) return bad("TUPLEn over a mark");
                PVal c = new_node(K_TUPLE);
                for (size_t i = g.stack.size() - k; i < g.stack.size(); ++i) g.push_item(g.nodes[c.idx], g.stack[i]);
                g.stack.resize(g.stack.size() - k);
                g.stack.push_back(c);
                break;
            }
            case 'q':
            case 'r':
            case 0x94: {  // BINPUT / LONG_BINPUT / MEMOIZE
                uint32_t at = (uint32_t)g.memo.size();
                if (op == 'q') {
                    if (!need(1)) return bad("truncated BINPUT");
                    at = *p++;
                } else if (op == 'r') {
                    if (!rd_u32(at)) return bad("truncated LONG_BINPUT");
                }
                if (g.stack.empty() || g.stack.back().kind == K_MARK) return bad("memo of nothing");
                if (at > len) return bad("memo index beyond the message size");
                if (at >= g.memo.size()) g.memo.resize((size_t)at + 1);
                g.memo[at] = g.stack.back();
                break;
            }
            case 'h':
            case 'j': {  // BINGET / LONG_BINGET
                uint32_t at;
                if (op == 'h') {
                    if (!need(1)) return bad("truncated BINGET");
                    at = *p++;
                } else if (!rd_u32(at)) {
                    return bad("truncated LONG_BINGET");
                }
                if (at >= g.memo.size()) return bad("BINGET of an unset memo slot");
                g.stack.push_back(g.memo[at]);
                break;
            }
            default:
                // GLOBAL, REDUCE, BUILD, NEWOBJ, PERSID, EXT*, text-mode opcodes: not data, not accepted
                err.set("pickle stream: opcode 0x%02x at byte %zu is not part of a plain-data pickle", op, (size_t)(p - 1 - msg));
                return RRL_EFORMAT;
        }
    }
    return bad("no STOP opcode");
}

// ---- the object graph -> rows -----------------------------------------------------------------------------------
// TensorData {shape, dtype, data}: only `data` (the safetensors blob, which states dtype and shape itself) is needed
int32_t tensor_blob(const Graph &g, const PVal &t, BlobView &b, const char *what, int64_t r, ErrBuf &err)
{
    const PVal *d = dict_get(g, t, "data");
    if (!d) {
        err.set("action %lld: %s is not a TensorData (no 'data' field)", (long long)r, what);
        return RRL_EFORMAT;
    }
    if (d->kind == K_BYTES) return parse_blob(d->s.p, d->s.n, b, err);
    if (d->kind == K_LIST && g.nodes[d->idx].all_bytes) return parse_blob(g.byte_data(g.nodes[d->idx]), g.nodes[d->idx].blen, b, err);
    err.set("action %lld: %s.data is neither bytes nor a list of byte values", (long long)r, what);
    return RRL_EFORMAT;
}

bool number(const PVal &v, float &out)
{
    if (v.kind == K_FLOAT) return out = (float)v.f, true;
    if (v.kind == K_INT || v.kind == K_BOOL) return out = (float)v.i, true;
    return false;
}

// RelayRLData (action.rs:203-218): {variant: payload} | (variant, payload) | a bare number | a bare TensorData
int32_t data_value(const Graph &g, const PVal &v, float &out, const char *what, int64_t r, ErrBuf &err)
{
    const PVal *payload = &v;
    if (v.kind == K_DICT && g.nodes[v.idx].len == 2 && !dict_get(g, v, "data"))
        payload = g.items(g.nodes[v.idx]) + 1;
    else if (v.kind == K_TUPLE && g.nodes[v.idx].len == 2 && g.items(g.nodes[v.idx])[0].kind == K_STR)
        payload = g.items(g.nodes[v.idx]) + 1;
    if (number(*payload, out)) return RRL_OK;
    if (payload->kind == K_DICT) {
        BlobView b;
        RRL_TRY(tensor_blob(g, *payload, b, what, r, err));
        if (b.count < 1) {
            err.set("action %lld: data['%s'] is an empty tensor", (long long)r, what);
            return RRL_EFORMAT;
        }
        out = blob_at(b, 0);
        return RRL_OK;
    }
    err.set("action %lld: data['%s'] carries neither a number nor a tensor", (long long)r, what);
    return RRL_EFORMAT;
}

int32_t action_row(const Graph &g, const PVal &a, int64_t r, int d, int A, bool with_baseline, const HostColumns &o, ErrBuf &err)
{
    if (a.kind != K_DICT) {
        err.set("action %lld is not a struct (dict)", (long long)r);
        return RRL_EFORMAT;
    }
    const PVal *obs = dict_get(g, a, "obs"), *act = dict_get(g, a, "act"), *mask = dict_get(g, a, "mask");
    const PVal *rew = dict_get(g, a, "rew"), *data = dict_get(g, a, "data"), *done = dict_get(g, a, "done");
    float rw = 0.f;
    if (!rew || !number(*rew, rw) || !done || (done->kind != K_BOOL && done->kind != K_INT)) {
        err.set("action %lld lacks rew / done (src/types/action.rs:428-437)", (long long)r);
        return RRL_EFORMAT;
    }
    const bool dn = done->i != 0;
    o.rew[r] = rw, o.done[r] = dn ? 1 : 0;
    o.act[r] = 0.f, o.val[r] = 0.f, o.logp[r] = 0.f;
    for (int i = 0; i < d; ++i) o.obs[r * d + i] = 0.f;
    for (int i = 0; i < A; ++i) o.mask[r * A + i] = 0.f;
    if (dn) return RRL_OK;  // a terminal marker has obs = act = mask = data = None (agent_zmq.rs:607-608)
    if (!obs || !act || !mask || obs->kind == K_NONE || act->kind == K_NONE || mask->kind == K_NONE) {
        err.set("action %lld: a transition without obs / act / mask", (long long)r);
        return RRL_EFORMAT;
    }
    BlobView bo, ba, bm;
    RRL_TRY(tensor_blob(g, *obs, bo, "obs", r, err));
    RRL_TRY(tensor_blob(g, *act, ba, "act", r, err));
    RRL_TRY(tensor_blob(g, *mask, bm, "mask", r, err));
    if ((int)bo.count != d || (int)bm.count != A || ba.count < 1) {
        err.set("action %lld: obs has %zu values (want %d), mask %zu (want %d), act %zu", (long long)r, bo.count, d, bm.count, A,
                ba.count);
        return RRL_EFORMAT;
    }
    for (int i = 0; i < d; ++i) o.obs[r * d + i] = blob_at(bo, i);
    for (int i = 0; i < A; ++i) o.mask[r * A + i] = blob_at(bm, i);
    o.act[r] = blob_at(ba, 0);
    const PVal *lp = data ? dict_get(g, *data, "logp_a") : nullptr;
    if (!lp) {
        err.set("action %lld: data['logp_a'] missing", (long long)r);
        return RRL_EFORMAT;
    }
    RRL_TRY(data_value(g, *lp, o.logp[r], "logp_a", r, err));
    if (with_baseline) {
        const PVal *v = dict_get(g, *data, "v");
        if (!v) {
            err.set("action %lld: data['v'] missing", (long long)r);
            return RRL_EFORMAT;
        }
        RRL_TRY(data_value(g, *v, o.val[r], "v", r, err));
    }
    return RRL_OK;
}

// the message's top object: the Vec<RelayRLAction> itself (a list; a tuple is tolerated)
int32_t top_actions(const Graph &g, const PVal *&items, int64_t &n, ErrBuf &err)
{
    if ((g.top.kind != K_LIST && g.top.kind != K_TUPLE) || (g.top.kind == K_LIST && g.nodes[g.top.idx].blen != 0)) {
        err.set("pickle message is not a sequence of actions (Vec<RelayRLAction>, src/types/trajectory.rs:50-55)");
        return RRL_EFORMAT;
    }
    items = g.items(g.nodes[g.top.idx]);
    n = g.nodes[g.top.idx].len;
    return RRL_OK;
}

}  // namespace

}  // namespace rrl

using namespace rrl;

extern "C" int32_t rrl_decode_pickle_trajectory(const uint8_t *msg, size_t len, int32_t obs_dim, int32_t act_dim, int32_t with_baseline,
                                                float *obs, float *act, float *mask, float *rew, float *val, float *logp,
                                                uint8_t *done, int64_t cap, int64_t *count, char *errbuf, size_t errcap)
{
    if ((!msg && len) || obs_dim < 1 || act_dim < 1 || !obs || !act || !mask || !rew || !val || !logp || !done || cap < 0 || !count)
        return RRL_EINVAL;
    ErrBuf err;
    const HostColumns o{obs, act, mask, rew, val, logp, done, cap};
    const int32_t st = abi_guard(err, "rrl_decode_pickle_trajectory", [&]() -> int32_t {
        Graph g;
        *count = 0;
        RRL_TRY(run_vm(msg, len, g, err));
        const PVal *items = nullptr;
        int64_t n = 0;
        RRL_TRY(top_actions(g, items, n, err));
        if (n > cap) {
            err.set("trajectory holds %lld actions, more than the %lld rows the caller provided", (long long)n, (long long)cap);
            return RRL_ECAPACITY;
        }
        for (int64_t r = 0; r < n; ++r) RRL_TRY(action_row(g, items[r], r, obs_dim, act_dim, with_baseline != 0, o, err));
        *count = n;
        return RRL_OK;
    });
    if (st != RRL_OK && errbuf && errcap) {
        strncpy(errbuf, err.msg.c_str(), errcap - 1);
        errbuf[errcap - 1] = 0;
    }
    return st;
}

extern "C" int32_t rrl_ingest_pickle(rrl_learner_t *h, const uint8_t *msg, size_t len, int64_t *n_actions)
{
    if (!h || (!msg && len)) return RRL_EINVAL;
    return abi_guard(h->err, "rrl_ingest_pickle", [&]() -> int32_t {
        const int d = h->cfg.obs_dim, A = h->cfg.act_dim;
        Graph g;
        RRL_TRY(run_vm(msg, len, g, h->err));
        const PVal *items = nullptr;
        int64_t n = 0;
        RRL_TRY(top_actions(g, items, n, h->err));
        if (n_actions) *n_actions = n;
        if (n == 0) return RRL_OK;
        // rows in chunks, all of the message or none of it (see rrl_ingest_grpc_trajectory)
        LearnerMark mark;
        RRL_TRY(learner_mark(h, mark));
        OwnedColumns cols(std::min<int64_t>(n, INGEST_CHUNK_ROWS), d, A);
        for (int64_t lo = 0; lo < n; lo += INGEST_CHUNK_ROWS) {
            const int64_t hi = std::min<int64_t>(n, lo + INGEST_CHUNK_ROWS);
            int32_t st = RRL_OK;
            for (int64_t r = lo; r < hi && st == RRL_OK; ++r) {
                st = action_row(g, items[r], r - lo, d, A, h->cfg.with_baseline != 0, cols.view, h->err);
                if (st != RRL_OK && lo > 0) {  // row numbers in messages count from the start of the trajectory
                    const std::string m = h->err.msg;
                    h->err.set("(chunk starting at action %lld) %s", (long long)lo, m.c_str());
                }
            }
            if (st == RRL_OK) {
                rrl_batch_view_t b;
                memset(&b, 0, sizeof b);
                b.obs = cols.view.obs, b.act = cols.view.act, b.mask = cols.view.mask, b.rew = cols.view.rew, b.val = cols.view.val;
                b.logp = cols.view.logp, b.done = cols.view.done, b.count = hi - lo, b.obs_dtype = RRL_F32, b.mask_dtype = RRL_F32;
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
