/* TEST INFRASTRUCTURE ONLY -- plain-C restatement of the reference's per-episode scan and
 * advantage statistics, used as the checker at sizes where the Python oracle is too slow
 * (64 M transitions) and as a timed CPU baseline.  Never linked into the product library.
 *
 * Parity is pinned: tests/test_oracle.py checks this file bit-for-bit against
 * oracle/learner_oracle.py (which is itself checked against the unmodified reference) and
 * against the tests/golden npz files generated from the reference.
 *
 * Build: see oracle/Makefile (-ffp-contract=off: no FMA contraction, so the recurrence
 * rounds exactly like scipy's lfilter does).
 *
 * Reference (relative to relayrl_framework/src/native/python/):
 *   _common/_algorithms/BaseReplayBuffer.py:12-27   discount_cumsum = lfilter([1],[1,-g], x[::-1])[::-1]
 *       direct-form-II-transposed, order 1:  y[n] = x[n] + z;  z = g * y[n]   (all float64)
 *   algorithms/REINFORCE/replay_buffer.py:48-79      finish_path
 *   _common/_algorithms/BaseReplayBuffer.py:30-53    statistics_scalar (float32, two-pass)
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>

/* One closed path [s, t).  adv/ret are float32 buffers, arithmetic is float64. */
static void finish_one(const float *rew, const float *val, int64_t s, int64_t t, float last_val,
                       double gamma, double lam, int with_baseline, float *adv, float *ret)
{
    const double gl = gamma * lam; /* Python evaluates gamma * lam once, in double */
    if (with_baseline) {
        /* rews = append(rew, lv); vals = append(val, lv)  (replay_buffer.py:66-67)
         * deltas = rews[:-1] + gamma * vals[1:] - vals[:-1]
         * adv = cumsum(deltas, gamma*lam); ret = cumsum(rews, gamma)[:-1] */
        double z_adv = 0.0;
        double y_ret = (double)last_val;      /* cumsum(rews)[last] = lv + 0 */
        double z_ret = gamma * y_ret;
        double v_next = (double)last_val;
        for (int64_t i = t - 1; i >= s; --i) {
            double r = (double)rew[i], v = (double)val[i];
            double delta = (r + gamma * v_next) - v;
            double y = delta + z_adv;
            z_adv = gl * y;
            adv[i] = (float)y;
            double yr = r + z_ret;
            z_ret = gamma * yr;
            ret[i] = (float)yr;
            v_next = v;
        }
    } else {
        /* last_val ignored (replay_buffer.py:74-77) */
        double z_adv = 0.0, z_ret = 0.0;
        for (int64_t i = t - 1; i >= s; --i) {
            double r = (double)rew[i];
            double y = r + z_adv;
            z_adv = gl * y;
            adv[i] = (float)y;
            double yr = r + z_ret;
            z_ret = gamma * yr;
            ret[i] = (float)yr;
        }
    }
}

/* All closed paths of an epoch: offsets has n_ep + 1 entries. */
void rrl_oracle_finish_paths(const float *rew, const float *val, const int64_t *offsets,
                             const float *last_vals, int64_t n_ep, double gamma, double lam,
                             int with_baseline, float *adv, float *ret)
{
    for (int64_t e = 0; e < n_ep; ++e)
        finish_one(rew, val, offsets[e], offsets[e + 1], last_vals ? last_vals[e] : 0.0f, gamma,
                   lam, with_baseline, adv, ret);
}

/* receive_trajectory's walk over a flat action stream (REINFORCE.py:70-95):
 * returns the number of closed paths; offsets must hold n+1, last_vals n, keep n entries. */
int64_t rrl_oracle_segment(const uint8_t *done, const float *rew, int64_t n, int64_t start_ptr,
                           int64_t *offsets, float *last_vals, int64_t *keep, int64_t *n_keep)
{
    int64_t ptr = start_ptr, ne = 0, nk = 0;
    offsets[0] = start_ptr;
    for (int64_t i = 0; i < n; ++i) {
        if (!done[i]) {
            keep[nk++] = i;
            ++ptr;
        } else {
            last_vals[ne] = rew[i];
            offsets[++ne] = ptr;
        }
    }
    *n_keep = nk;
    return ne;
}

/* numpy's pairwise float32 summation (numpy/core/src/umath/loops_utils.h.src, pairwise_sum):
 * blocks of <= 128 are summed with 8 interleaved accumulators, larger inputs are halved at a
 * multiple of 8.  Restated so that mean/std agree with np.sum bit-for-bit. */
static float pairwise_sum_f32(const float *a, int64_t n)
{
    if (n < 8) {
        float res = 0.f;
        for (int64_t i = 0; i < n; ++i) res += a[i];
        return res;
    } else if (n <= 128) {
        float r[8];
        for (int k = 0; k < 8; ++k) r[k] = a[k];
        int64_t i;
        for (i = 8; i < n - (n % 8); i += 8)
            for (int k = 0; k < 8; ++k) r[k] += a[i + k];
        float res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    } else {
        int64_t n2 = n / 2;
        n2 -= n2 % 8;
        return pairwise_sum_f32(a, n2) + pairwise_sum_f32(a + n2, n - n2);
    }
}

/* statistics_scalar + the normalisation in ReplayBuffer.get (replay_buffer.py:94,105).
 * tmp must hold n floats.  out may alias adv. */
void rrl_oracle_normalise(const float *adv, int64_t n, float *tmp, float *out, float *mean_out,
                          float *std_out)
{
    float mean = pairwise_sum_f32(adv, n) / (float)n;
    for (int64_t i = 0; i < n; ++i) {
        float d = adv[i] - mean;
        tmp[i] = d * d;
    }
    float std = sqrtf(pairwise_sum_f32(tmp, n) / (float)n);
    for (int64_t i = 0; i < n; ++i) out[i] = (adv[i] - mean) / std;
    *mean_out = mean;
    *std_out = std;
}
