// Shared between the two actor kernels (actor.cu: CUDA cores, any shape; actor_tc.cu: tcgen05, wide layers): the
// argument block, the Philox draw and the per-environment tail -- masked softmax, sampling, outputs
// (kernel.py:23-37,99-107 of the reference's REINFORCE model).
#pragma once
#include "common.cuh"

namespace rrl {

struct ActorNet {
    const float *params;  // flat torch order
    const float *w1p;     // W1 padded to [H][dK] (dK = roundup(d, 4)), zero filled  (CUDA-core kernel)
    const unsigned char *pack;  // bf16x3 planes in the order the tensor-core kernel streams them (actor_tc.cu)
};

struct ActorArgs {
    ActorNet pi, vf;
    int d, dK, A, with_v;
    const float *obs, *mask;
    int64_t batch;
    uint64_t seed, offset;
    const int64_t *force_act;
    int64_t *act;
    float *logp_a, *v, *probs, *logits;
};

// Philox4x32-10 (Salmon et al. 2011), counter = (ctr, 0), key = seed; returns the first word.
__device__ __forceinline__ unsigned philox_first(uint64_t seed, uint64_t ctr)
{
    unsigned c0 = (unsigned)ctr, c1 = (unsigned)(ctr >> 32), c2 = 0u, c3 = 0u;
    unsigned k0 = (unsigned)seed, k1 = (unsigned)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const unsigned hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const unsigned n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0, c1 = n1, c2 = n2, c3 = n3;
        k0 += 0x9E3779B9u, k1 += 0xBB67AE85u;
    }
    return c0;
}

// One environment's tail.  head[0..A) holds x + b3 (the unmasked logits) in shared memory, element q at head[q * hs],
// and is overwritten with the masked logits.  Writes act, logp_a (= the raw masked logit of the action) and,
// when asked for, probs and logits.
__device__ __forceinline__ void actor_policy_tail(const ActorArgs &a, float *head, int hs, int64_t row)
{
    const int A = a.A;
    float mx = -INFINITY;
    for (int q = 0; q < A; ++q) {
        const float lg = head[q * hs] + (__ldcs(a.mask + row * A + q) - 1.f) * 1e8f;  // kernel.py:23-26
        head[q * hs] = lg;
        mx = fmaxf(mx, lg);
    }
    float se = 0.f;
    for (int q = 0; q < A; ++q) se += expf(head[q * hs] - mx);
    int at;
    if (a.force_act) {
        at = (int)a.force_act[row];
        at = min(max(at, 0), A - 1);
    } else {
        // inverse CDF with one Philox uniform in [0,1); a masked action has p == 0 exactly
        const float u = (float)(philox_first(a.seed, a.offset + (uint64_t)row) >> 8) * (1.0f / 16777216.0f);
        float c = 0.f;
        at = -1;
        int last_pos = 0;
        for (int q = 0; q < A; ++q) {
            const float p = expf(head[q * hs] - mx) / se;
            if (p > 0.f) last_pos = q;
            c += p;
            if (at < 0 && u < c) at = q;
        }
        if (at < 0) at = last_pos;
    }
    for (int q = 0; q < A; ++q) {
        const float lg = head[q * hs];
        if (a.logits) a.logits[row * A + q] = lg;
        if (a.probs) a.probs[row * A + q] = expf(lg - mx) / se;
    }
    a.act[row] = at;
    a.logp_a[row] = head[at * hs];
}

// actor_tc.cu
bool actor_tc_supported(int d, int H, int A, int64_t batch);
size_t actor_tc_pack_bytes(int d, int H);
int32_t actor_tc_pack_launch(const float *params, int d, int H, int out, unsigned char *pack, cudaStream_t stream, ErrBuf &err);
int32_t actor_tc_forward_launch(const ActorArgs &a, int H, cudaStream_t stream, ErrBuf &err);

}  // namespace rrl
