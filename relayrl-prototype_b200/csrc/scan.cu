// K2/K3: reward-to-go + GAE-lambda as ONE segmented reverse scan, fused with the advantage
// statistics, and the (optional) materialised normalisation.
//
// Replaces, per epoch, E calls of ReplayBuffer.finish_path (two scipy.signal.lfilter calls each;
// relayrl_framework/src/native/python/algorithms/REINFORCE/replay_buffer.py:48-79,
// _common/_algorithms/BaseReplayBuffer.py:12-27) and statistics_scalar + the normalisation in
// ReplayBuffer.get (BaseReplayBuffer.py:30-53, replay_buffer.py:94,105).
//
// Data layout: rew/val/adv/ret are flat f32 [N] (the reference's SoA numpy buffers,
// replay_buffer.py:20-28); offsets int64 [E+1] are the successive path_start_idx values;
// last_vals f32 [E] are the terminal markers' rewards (REINFORCE.py:86).
//
// Algorithm: both recurrences y[t] = x[t] + c*y[t+1] are affine maps of the value carried in
// from the right; maps compose associatively, and an episode end resets the carry (the map's
// slope becomes 0).  One pass over HBM (16 B/transition with baseline, 12 B without):
//   persistent CTAs (cooperative launch), 2048-transition tiles taken right-to-left in a fixed
//   stride; a producer warp prefetches the next tile while 8 consumer warps scan the current one
//     -> cp.async 16-byte copies into swizzled shared memory, blocked 8-per-thread reads
//     -> episode ends of the tile found by a 32-ary warp search in `offsets`
//     -> thread-serial map, warp-shuffle suffix scan, 8 warp maps through shared memory
//     -> single-pass decoupled look-back across tiles (aggregate / inclusive descriptors whose
//        64-bit words carry their own validity: no flags, no fences on the critical path)
//     -> second thread-serial pass seeded with the carry, results staged through shared
//        memory and written with coalesced 128-bit stores
//     -> per-tile (sum, sum of squares) of the f32-rounded advantages in f64.
// Arithmetic is float64 with float32 storage, as in the reference.
#include <type_traits>

#include "common.cuh"

namespace rrl {

namespace {

#ifndef SCAN_CONSUMER_THREADS
#define SCAN_CONSUMER_THREADS 256
#endif
#ifndef SCAN_MIN_CTAS
#define SCAN_MIN_CTAS 2
#endif
constexpr int SCAN_THREADS = SCAN_CONSUMER_THREADS;
constexpr int SCAN_VPT = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_VPT;  // 2048 transitions per CTA
constexpr int SCAN_WARPS = SCAN_THREADS / 32;
constexpr int SCAN_SMEM_FLOATS = SCAN_TILE + SCAN_TILE / 8 + 8;

// 4 floats of padding after every 32: blocked 8-per-thread LDS.128 reads become conflict-free
__device__ __forceinline__ int sidx(int j) { return j + ((j >> 5) << 2); }

struct Map {  // ret: y = br + ar * carry ; adv: y = ba + aa * carry
    double ar, br, aa, ba;
};

// left∘right: `right` is applied first (it is nearer the end of the episode)
__device__ __forceinline__ Map compose(const Map &l, const Map &r)
{
    Map o;
    o.ar = l.ar * r.ar;
    o.br = fma(l.ar, r.br, l.br);
    o.aa = l.aa * r.aa;
    o.ba = fma(l.aa, r.ba, l.ba);
    return o;
}

__device__ __forceinline__ Map shfl_down_map(const Map &m, int d)
{
    Map o;
    o.ar = __shfl_down_sync(0xffffffffu, m.ar, d);
    o.br = __shfl_down_sync(0xffffffffu, m.br, d);
    o.aa = __shfl_down_sync(0xffffffffu, m.aa, d);
    o.ba = __shfl_down_sync(0xffffffffu, m.ba, d);
    return o;
}
__device__ __forceinline__ Map shfl_up_map(const Map &m, int d)
{
    Map o;
    o.ar = __shfl_up_sync(0xffffffffu, m.ar, d);
    o.br = __shfl_up_sync(0xffffffffu, m.br, d);
    o.aa = __shfl_up_sync(0xffffffffu, m.aa, d);
    o.ba = __shfl_up_sync(0xffffffffu, m.ba, d);
    return o;
}
__device__ __forceinline__ Map shfl_map(const Map &m, int src)
{
    Map o;
    o.ar = __shfl_sync(0xffffffffu, m.ar, src);
    o.br = __shfl_sync(0xffffffffu, m.br, src);
    o.aa = __shfl_sync(0xffffffffu, m.aa, src);
    o.ba = __shfl_sync(0xffffffffu, m.ba, src);
    return o;
}

// Tile descriptors carry their own validity: every 64-bit word starts as the sentinel 0xFFFF...F
// (memset before the launch) and is written once with a single 64-bit store.  A reader polls the
// words themselves (L2-coherent relaxed loads) -- no flag, no fence, no MEMBAR on the critical path.
// The sentinel is a NaN payload the arithmetic here cannot produce.
__device__ __forceinline__ bool is_unset(double x) { return __double_as_longlong(x) == -1ll; }
__device__ __forceinline__ double2 ld_relaxed2(const double2 *p)
{
    double2 v;
    asm volatile("ld.relaxed.gpu.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed2(double2 *p, double x, double y)
{
    asm volatile("st.relaxed.gpu.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(x), "d"(y) : "memory");
}

// f32 <-> f64 by bit manipulation: the hardware conversion instructions (F2F.F64.F32 and back) issue
// on the quarter-rate XU pipe, which the first version of this kernel saturated (95 % busy at 19 % of
// HBM peak, profiles/r1_summary.md).  Zero, denormal, Inf and NaN take the slow instruction.
// out of line on purpose: inlined, the compiler if-converts the rare branch and issues the slow
// conversion for every element
__device__ __noinline__ double f32_to_f64_slow(float x) { return (double)x; }
__device__ __noinline__ float f64_to_f32_slow(double d) { return __double2float_rn(d); }

__device__ __forceinline__ double f32_to_f64(float x)
{
    const unsigned u = __float_as_uint(x);
    const unsigned a = u & 0x7fffffffu;
    if (a - 0x00800000u >= 0x7f000000u && a != 0u) return f32_to_f64_slow(x);  // denormal / Inf / NaN (rare)
    const unsigned hi = (u & 0x80000000u) | ((a >> 3) + 0x38000000u);
    return __hiloint2double(a ? hi : u, u << 29);
}
// The same without the branch: right for zero and normal numbers; f32_needs_slow() says when it is not.  The scan
// tests all 8 (16) inputs of a thread at once and takes ONE branch to a slow copy of the loop, instead of one
// convergence barrier (BSSY/BRA/BSYNC) per conversion -- those were 16 % of the instruction stream.
__device__ __forceinline__ bool f32_needs_slow(float x)
{
    const unsigned a = __float_as_uint(x) & 0x7fffffffu;
    return a - 0x00800000u >= 0x7f000000u && a != 0u;
}
__device__ __forceinline__ double f32_to_f64_fast(float x)
{
    const unsigned u = __float_as_uint(x);
    const unsigned a = u & 0x7fffffffu;
    const unsigned hi = (u & 0x80000000u) | ((a >> 3) + 0x38000000u);
    return __hiloint2double(a ? hi : u, u << 29);
}
// d rounded to 24 significant bits, still a double (Veltkamp split with 2^29 + 1)
__device__ __forceinline__ double round_to_f32_precision(double d)
{
    // explicit _rn intrinsics: the compiler must not contract c - d into an FMA
    const double c = __dmul_rn(d, 536870913.0);
    return __dsub_rn(c, __dsub_rn(c, d));
}
// h = round_to_f32_precision(d): exact narrowing by bits when h is a normal f32, else the slow instruction on d
__device__ __forceinline__ float narrow_f32(double h, double d)
{
    const unsigned hi = (unsigned)__double2hiint(h), lo = (unsigned)__double2loint(h);
    const unsigned t = hi - 0x38000000u;             // exponent rebased from 1023 to 127 (sign untouched)
    if (((t << 1) >> 21) - 1u >= 254u) return f64_to_f32_slow(d);  // zero, f32-denormal range, overflow, Inf, NaN
    return __uint_as_float((hi & 0x80000000u) | __funnelshift_l(lo, t, 3));
}
// branch-free form + its validity test (same convention as above; an exact zero is handled in the fast form)
__device__ __forceinline__ bool narrow_needs_slow(double h)
{
    const unsigned hi = (unsigned)__double2hiint(h), lo = (unsigned)__double2loint(h);
    const unsigned t = hi - 0x38000000u;
    return ((t << 1) >> 21) - 1u >= 254u && ((hi << 1) | lo) != 0u;
}
__device__ __forceinline__ float narrow_f32_fast(double h)
{
    const unsigned hi = (unsigned)__double2hiint(h), lo = (unsigned)__double2loint(h);
    const unsigned bits = (hi & 0x80000000u) | __funnelshift_l(lo, hi - 0x38000000u, 3);
    return __uint_as_float(((hi << 1) | lo) != 0u ? bits : (hi & 0x80000000u));
}

struct ScanArgs {
    const float *rew, *val;
    const int64_t *offsets;
    const float *last_vals;
    int64_t n_ep;    // closed paths; the scan covers [0, offsets[n_ep]) (read on the device)
    double gamma, gl;
    double gamma_p, gl_p;  // gamma^SCAN_VPT, gl^SCAN_VPT: the slope of a thread's map when no path ends in its run
    float *adv, *ret;
    // workspace
    int *ticket;
    int *flags;      // (reserved)
    Map *agg;        // tile map (carry-in from the right -> value at the tile's first element)
    double2 *incl;   // final (ret, adv) at the tile's first element
    double2 *partials;
};

// tile states published for the look-back

// ---- persistent, warp-specialised tile pipeline ------------------------------------------------
// A CTA is 8 consumer warps + 1 producer warp.  CTA c owns the logical tiles c, c+G, c+2G, ...
// (logical tile j = physical tile num_tiles-1-j: right-most first); the launch is cooperative, so
// all G CTAs are co-resident and the look-back can never wait on a CTA that is not running.
// While the consumers work on tile j, the producer warp prepares tile j+G in the other buffer:
// cp.async of rew/val into swizzled shared memory (the bytes in flight that keep HBM busy), the
// 32-ary search for the tile's first path boundary, and the end-of-path bit mask + last_vals.
constexpr int SCAN_CTA_THREADS = SCAN_THREADS + 32;

struct TileBuf {
    float r[SCAN_SMEM_FLOATS];   // rew in, adv out
    float v[SCAN_SMEM_FLOATS];   // val in, ret out
    float lv[SCAN_TILE];         // last_val at the positions where a path ends
    unsigned end[SCAN_TILE / 32];
    int tile;                    // physical tile index, -1 = no more work
    int pad[3];
};
static_assert(sizeof(TileBuf) % 16 == 0, "TileBuf must keep 16-byte alignment");

__device__ __forceinline__ void cp_async16_zfill(void *smem_dst, const void *gmem_src, int src_bytes)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem_src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(SCAN_THREADS) : "memory"); }

// producer warp: fill `B` for physical tile `tile` (or mark it empty)
template <bool WB>
__device__ __forceinline__ void scan_prepare(const ScanArgs &a, TileBuf &B, int tile, int64_t n_scan, int lane)
{
    if (lane == 0) B.tile = tile;
    if (tile < 0) return;
    for (int w = lane; w < SCAN_TILE / 32; w += 32) B.end[w] = 0u;
    const int64_t t0 = (int64_t)tile * SCAN_TILE;
    const int cnt = (int)min((int64_t)SCAN_TILE, n_scan - t0);
    // 512 16-byte chunks per array, 16 per lane; out-of-range bytes are zero-filled
#pragma unroll 4
    for (int c = lane; c < SCAN_TILE / 4; c += 32) {
        const int valid = min(max(cnt - 4 * c, 0), 4) * 4;
        const int64_t g = valid ? t0 + 4 * c : 0;
        cp_async16_zfill(&B.r[sidx(4 * c)], a.rew + g, valid);
        if (WB) cp_async16_zfill(&B.v[sidx(4 * c)], a.val + g, valid);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (WB && lane == 0) B.v[sidx(SCAN_TILE)] = (t0 + SCAN_TILE < n_scan) ? a.val[t0 + SCAN_TILE] : 0.f;

    // first path boundary beyond t0: smallest e in [1, E] with offsets[e] > t0
    long long lo = 1, hi = a.n_ep;
    while (lo < hi) {
        const long long span = hi - lo;
        const long long p = lo + (span * (lane + 1)) / 33;
        const bool gt = a.offsets[p] > t0;
        const unsigned b = __ballot_sync(0xffffffffu, gt);
        if (b == 0u) {
            lo = lo + (span * 32) / 33 + 1;
        } else {
            const int f = __ffs(b) - 1;
            hi = lo + (span * (f + 1)) / 33;
            if (f > 0) lo = lo + (span * f) / 33 + 1;
        }
    }
    __syncwarp();  // the zeroed mask words are visible to the whole warp before the atomics
    const int64_t t1 = t0 + cnt;
    for (long long k0 = lo; k0 <= a.n_ep; k0 += 128) {
        // four independent probes per lane per round keep the dependent-load count low when
        // paths are very short
        bool more = false;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const long long k = k0 + u * 32 + lane;
            if (k <= a.n_ep) {
                const int64_t b = a.offsets[k];
                if (b <= t1) {
                    if (u == 3) more = true;
                    if (b > a.offsets[k - 1]) {  // a non-empty path ends at b-1
                        const int pos = (int)(b - 1 - t0);
                        atomicOr(&B.end[pos >> 5], 1u << (pos & 31));
                        B.lv[pos] = a.last_vals[k - 1];
                    }
                }
            }
        }
        if (!__any_sync(0xffffffffu, more)) break;  // offsets are sorted: nothing further can fall in the tile
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
}

// consumer warps (threads 0..255): scan the tile in `B`, publish / look back, write the outputs
template <bool WB>
__device__ __forceinline__ void scan_consume(const ScanArgs &a, TileBuf &B, int tile, int64_t n_scan,
                                             int num_tiles, Map *wm_s, double *carry_s, double (*red_s)[2],
                                             int tid)
{
    const int lane = tid & 31, warp = tid >> 5;
    float *r_s = B.r, *v_s = B.v;
    const int64_t t0 = (int64_t)tile * SCAN_TILE;
    const int cnt = (int)min((int64_t)SCAN_TILE, n_scan - t0);

    // ---- blocked reads: SCAN_VPT consecutive transitions per thread ---------------------------
    const int j0 = tid * SCAN_VPT;
    float r[SCAN_VPT], v[SCAN_VPT];
    {
        const float4 ra = *reinterpret_cast<const float4 *>(&r_s[sidx(j0)]);
        const float4 rb = *reinterpret_cast<const float4 *>(&r_s[sidx(j0) + 4]);
        r[0] = ra.x, r[1] = ra.y, r[2] = ra.z, r[3] = ra.w, r[4] = rb.x, r[5] = rb.y, r[6] = rb.z, r[7] = rb.w;
        if (WB) {
            const float4 va = *reinterpret_cast<const float4 *>(&v_s[sidx(j0)]);
            const float4 vb = *reinterpret_cast<const float4 *>(&v_s[sidx(j0) + 4]);
            v[0] = va.x, v[1] = va.y, v[2] = va.z, v[3] = va.w, v[4] = vb.x, v[5] = vb.y, v[6] = vb.z, v[7] = vb.w;
        }
    }
    const float v_edge = WB ? v_s[sidx(j0 + SCAN_VPT)] : 0.f;
    const unsigned eb = (B.end[j0 >> 5] >> (j0 & 31)) & 0xffu;
    const bool open_right = ((B.end[(cnt - 1) >> 5] >> ((cnt - 1) & 31)) & 1u) == 0u;

    // ---- pass 1: this thread's map; inputs are widened once and kept (rr = reward, dl = the
    // GAE delta, both pure data) so that pass 2 is two DFMAs per transition ----------------------
    // The widening is done branch-free; ONE test over all the thread's inputs sends a thread whose run holds a
    // denormal, Inf or NaN to a copy of the loop that uses the exact (slow) conversion.
    Map m;
    double rr[SCAN_VPT], dl[WB ? SCAN_VPT : 1];
    bool in_slow = WB && f32_needs_slow(v_edge);
#pragma unroll
    for (int k = 0; k < SCAN_VPT; ++k) in_slow |= f32_needs_slow(r[k]) || (WB && f32_needs_slow(v[k]));
    auto pass1 = [&](auto slow_tag) {
        constexpr bool SLOW = decltype(slow_tag)::value;
        auto widen = [](float x) { return SLOW ? f32_to_f64(x) : f32_to_f64_fast(x); };
        double ret_n = 0.0, adv_n = 0.0, v_n = WB ? widen(v_edge) : 0.0;
        if (eb == 0u) {  // no path ends inside this thread's run (the common case)
#pragma unroll
            for (int k = SCAN_VPT - 1; k >= 0; --k) {
                rr[k] = widen(r[k]);
                if (WB) {
                    const double vk = widen(v[k]);
                    dl[k] = fma(a.gamma, v_n, rr[k]) - vk;
                    v_n = vk;
                }
                adv_n = fma(a.gl, adv_n, WB ? dl[k] : rr[k]);
                ret_n = fma(a.gamma, ret_n, rr[k]);
            }
            m.ar = a.gamma_p, m.aa = a.gl_p;
        } else {
            double ar = 1.0, aa = 1.0;
#pragma unroll
            for (int k = SCAN_VPT - 1; k >= 0; --k) {
                if ((eb >> k) & 1u) {
                    // with baseline the marker's reward bootstraps both recurrences
                    // (replay_buffer.py:66-73); without it is ignored (replay_buffer.py:75-77)
                    const double lvk = f32_to_f64(B.lv[j0 + k]);
                    ret_n = WB ? lvk : 0.0;
                    adv_n = 0.0;
                    v_n = lvk;
                    ar = 0.0;
                    aa = 0.0;
                }
                rr[k] = widen(r[k]);
                if (WB) {
                    const double vk = widen(v[k]);
                    dl[k] = fma(a.gamma, v_n, rr[k]) - vk;
                    v_n = vk;
                }
                adv_n = fma(a.gl, adv_n, WB ? dl[k] : rr[k]);
                ret_n = fma(a.gamma, ret_n, rr[k]);
                ar *= a.gamma;
                aa *= a.gl;
            }
            m.ar = ar, m.aa = aa;
        }
        m.br = ret_n, m.ba = adv_n;
    };
    if (in_slow)
        pass1(std::true_type{});
    else
        pass1(std::false_type{});

    // ---- warp suffix scan of the maps (lane l: M_l∘M_{l+1}∘...∘M_31) ----------------------------
    Map S = m;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const Map o = shfl_down_map(S, d);
        if (lane + d < 32) S = compose(S, o);
    }
    Map Ex = shfl_down_map(S, 1);  // maps of the lanes to the right of this one
    if (lane == 31) Ex.ar = 1.0, Ex.br = 0.0, Ex.aa = 1.0, Ex.ba = 0.0;
    if (lane == 0) wm_s[warp] = S;
    consumer_sync();  // (also: every consumer has finished reading r_s / v_s)

    Map Wsuf;  // maps of the warps to the right of this one
    Wsuf.ar = 1.0, Wsuf.br = 0.0, Wsuf.aa = 1.0, Wsuf.ba = 0.0;
    for (int w = warp + 1; w < SCAN_WARPS; ++w) Wsuf = compose(Wsuf, wm_s[w]);

    // ---- publish the tile, fetch the carry from the tiles to the right -----------------------
    if (warp == 0) {
        const Map T = compose(wm_s[0], Wsuf);
        const bool closed = (T.ar == 0.0 && T.aa == 0.0);  // value at the tile's first element is final
        if (lane == 0) {
            if (closed) {
                st_relaxed2(&a.incl[tile], T.br, T.ba);
            } else {
                double2 *ap = reinterpret_cast<double2 *>(&a.agg[tile]);
                st_relaxed2(ap, T.ar, T.br);
                st_relaxed2(ap + 1, T.aa, T.ba);
            }
        }
        double cr = 0.0, ca = 0.0;
        if (open_right) {  // the right-most path of this tile continues in the next tile
            Map acc;
            acc.ar = 1.0, acc.br = 0.0, acc.aa = 1.0, acc.ba = 0.0;
            int base = tile + 1;
            while (true) {
                const int k = base + lane;
                Map q;
                q.ar = 0.0, q.br = 0.0, q.aa = 0.0, q.ba = 0.0;
                if (k < num_tiles) {
                    const double2 *ap = reinterpret_cast<const double2 *>(&a.agg[k]);
                    while (true) {
                        const double2 iv = ld_relaxed2(&a.incl[k]);
                        if (!is_unset(iv.x) && !is_unset(iv.y)) {  // final value of that tile's first element
                            q.br = iv.x, q.ba = iv.y;
                            break;
                        }
                        const double2 lo2 = ld_relaxed2(ap), hi2 = ld_relaxed2(ap + 1);
                        if (!is_unset(lo2.x) && !is_unset(lo2.y) && !is_unset(hi2.x) && !is_unset(hi2.y)) {
                            q.ar = lo2.x, q.br = lo2.y, q.aa = hi2.x, q.ba = hi2.y;
                            break;
                        }
                    }
                }
                __syncwarp();
                // ordered prefix over the window: P_l = q_0∘q_1∘...∘q_l
                Map P = q;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const Map o = shfl_up_map(P, d);
                    if (lane >= d) P = compose(o, P);
                }
                const unsigned z = __ballot_sync(0xffffffffu, P.ar == 0.0 && P.aa == 0.0);
                if (z != 0u) {
                    const Map Ps = shfl_map(P, __ffs(z) - 1);
                    const Map R = compose(acc, Ps);
                    cr = R.br, ca = R.ba;
                    break;
                }
                acc = compose(acc, shfl_map(P, 31));
                base += 32;
            }
        }
        if (lane == 0) {
            if (!closed) st_relaxed2(&a.incl[tile], fma(T.ar, cr, T.br), fma(T.aa, ca, T.ba));
            carry_s[0] = cr, carry_s[1] = ca;
        }
    }
    consumer_sync();

    // ---- pass 2: final values, seeded with the carry ------------------------------------------
    // Same scheme for the narrowing: branch-free, one test over the thread's 16 results; a thread with a result
    // in the f32-denormal range, an overflow, Inf or NaN redoes its run with the exact (slow) conversion.
    double s1 = 0.0, s2 = 0.0;
    {
        const double wr = fma(Wsuf.ar, carry_s[0], Wsuf.br), wa = fma(Wsuf.aa, carry_s[1], Wsuf.ba);
        const double ret0 = fma(Ex.ar, wr, Ex.br), adv0 = fma(Ex.aa, wa, Ex.ba);
        float ao[SCAN_VPT], ro[SCAN_VPT];
        auto pass2 = [&](auto slow_tag) -> bool {
            constexpr bool SLOW = decltype(slow_tag)::value;
            double ret_n = ret0, adv_n = adv0, t1 = 0.0, t2 = 0.0;
            bool bad = false;
#pragma unroll
            for (int k = SCAN_VPT - 1; k >= 0; --k) {
                if (eb != 0u && ((eb >> k) & 1u)) {
                    ret_n = WB ? f32_to_f64(B.lv[j0 + k]) : 0.0;
                    adv_n = 0.0;
                }
                adv_n = fma(a.gl, adv_n, WB ? dl[k] : rr[k]);
                ret_n = fma(a.gamma, ret_n, rr[k]);
                const double ha = round_to_f32_precision(adv_n), hr = round_to_f32_precision(ret_n);
                double x;  // the f32-rounded advantage, as the reference's buffer holds it
                if (SLOW) {
                    ao[k] = narrow_f32(ha, adv_n), ro[k] = narrow_f32(hr, ret_n);
                    x = f32_to_f64(ao[k]);
                } else {
                    bad |= narrow_needs_slow(ha) || narrow_needs_slow(hr);
                    ao[k] = narrow_f32_fast(ha), ro[k] = narrow_f32_fast(hr);
                    x = ha;  // ha is exactly that f32 value, widened
                }
                if (j0 + k < cnt) {
                    t1 += x;
                    t2 = fma(x, x, t2);
                }
            }
            s1 = t1, s2 = t2;
            return bad;
        };
        if (pass2(std::false_type{})) pass2(std::true_type{});
        *reinterpret_cast<float4 *>(&r_s[sidx(j0)]) = make_float4(ao[0], ao[1], ao[2], ao[3]);
        *reinterpret_cast<float4 *>(&r_s[sidx(j0) + 4]) = make_float4(ao[4], ao[5], ao[6], ao[7]);
        *reinterpret_cast<float4 *>(&v_s[sidx(j0)]) = make_float4(ro[0], ro[1], ro[2], ro[3]);
        *reinterpret_cast<float4 *>(&v_s[sidx(j0) + 4]) = make_float4(ro[4], ro[5], ro[6], ro[7]);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        s1 += __shfl_xor_sync(0xffffffffu, s1, d);
        s2 += __shfl_xor_sync(0xffffffffu, s2, d);
    }
    if (lane == 0) red_s[warp][0] = s1, red_s[warp][1] = s2;
    consumer_sync();

    // ---- coalesced 128-bit stores -------------------------------------------------------------
#pragma unroll
    for (int c = tid; c < SCAN_TILE / 4; c += SCAN_THREADS) {
        const int64_t g = t0 + 4 * c;
        const float4 a4 = *reinterpret_cast<const float4 *>(&r_s[sidx(4 * c)]);
        const float4 r4 = *reinterpret_cast<const float4 *>(&v_s[sidx(4 * c)]);
        if (g + 3 < n_scan) {
            __stcs(reinterpret_cast<float4 *>(a.adv + g), a4);
            __stcs(reinterpret_cast<float4 *>(a.ret + g), r4);
        } else {
            const float *ap = &a4.x, *rp = &r4.x;
            for (int k = 0; k < 4; ++k)
                if (g + k < n_scan) {
                    a.adv[g + k] = ap[k];
                    a.ret[g + k] = rp[k];
                }
        }
    }
    if (tid == 0) {
        double t1 = 0.0, t2 = 0.0;
        for (int w = 0; w < SCAN_WARPS; ++w) t1 += red_s[w][0], t2 += red_s[w][1];
        a.partials[tile] = make_double2(t1, t2);
    }
}

template <bool WB>
__global__ void __launch_bounds__(SCAN_CTA_THREADS, SCAN_MIN_CTAS) scan_kernel(ScanArgs a)
{
    extern __shared__ __align__(16) unsigned char scan_dyn[];
    TileBuf *buf = reinterpret_cast<TileBuf *>(scan_dyn);
    __shared__ Map wm_s[SCAN_WARPS];
    __shared__ double carry_s[2];
    __shared__ double red_s[SCAN_WARPS][2];

    const int tid = threadIdx.x, lane = tid & 31;
    const bool producer = tid >= SCAN_THREADS;
    const int64_t n_scan = a.offsets[a.n_ep];  // end of the last closed path
    const int num_tiles = (int)((n_scan + SCAN_TILE - 1) / SCAN_TILE);
    const int G = gridDim.x;
    int j = blockIdx.x;  // logical tile; physical = num_tiles - 1 - j (right-most first)

    if (producer) scan_prepare<WB>(a, buf[0], j < num_tiles ? num_tiles - 1 - j : -1, n_scan, lane);
    __syncthreads();
    int b = 0;
    while (true) {
        const int tile = buf[b].tile;
        if (tile < 0) break;
        if (producer) {
            const int jn = j + G;
            scan_prepare<WB>(a, buf[b ^ 1], jn < num_tiles ? num_tiles - 1 - jn : -1, n_scan, lane);
        } else {
            scan_consume<WB>(a, buf[b], tile, n_scan, num_tiles, wm_s, carry_s, red_s, tid);
        }
        __syncthreads();
        b ^= 1;
        j += G;
    }
}

// mean/std exactly as statistics_scalar defines them (BaseReplayBuffer.py:41-47): f32 mean,
// population variance about that f32 mean, f32 std, no epsilon.
__device__ void finalize_stats(double *stats)
{
    const double sum = stats[0], sumsq = stats[1], n = stats[2];
    const float mean_f = (float)(sum / n);
    const double mm = (double)mean_f;
    double m2 = sumsq - 2.0 * mm * sum + n * mm * mm;
    if (m2 < 0.0) m2 = 0.0;
    const float std_f = (float)sqrt(m2 / n);
    stats[3] = (double)mean_f;
    stats[4] = (double)std_f;
}

// Sums the per-tile partials in a fixed order (deterministic), adds the unfinished tail
// [n_scan, n_total) which keeps whatever adv it held.
__global__ void __launch_bounds__(1024) stats_reduce_kernel(const double2 *partials, const int64_t *offsets,
                                                            int64_t n_ep, const float *adv,
                                                            int64_t n_total, double *stats, int finalize)
{
    const int64_t n_scan = n_ep > 0 ? offsets[n_ep] : 0;
    const int num_tiles = (int)((n_scan + SCAN_TILE - 1) / SCAN_TILE);
    __shared__ double s_a[1024], s_b[1024];
    const int tid = threadIdx.x;
    double a = 0.0, b = 0.0;
    for (int i = tid; i < num_tiles; i += 1024) {
        const double2 p = partials[i];
        a += p.x, b += p.y;
    }
    for (int64_t i = n_scan + tid; i < n_total; i += 1024) {
        const double x = (double)adv[i];
        a += x, b = fma(x, x, b);
    }
    s_a[tid] = a, s_b[tid] = b;
    __syncthreads();
    for (int s = 512; s > 0; s >>= 1) {
        if (tid < s) s_a[tid] += s_a[tid + s], s_b[tid] += s_b[tid + s];
        __syncthreads();
    }
    if (tid == 0) {
        stats[0] = s_a[0], stats[1] = s_b[0], stats[2] = (double)n_total;
        if (finalize) finalize_stats(stats);
    }
}

__global__ void stats_finalize_kernel(double *stats) { finalize_stats(stats); }

// gathered: [nranks][4] (sum, sumsq, n, n_episodes) in rank order -> identical result on every rank
__global__ void stats_merge_kernel(const double *gathered, int nranks, double *stats)
{
    double a = 0.0, b = 0.0, n = 0.0, e = 0.0;
    for (int r = 0; r < nranks; ++r)
        a += gathered[4 * r], b += gathered[4 * r + 1], n += gathered[4 * r + 2], e += gathered[4 * r + 3];
    stats[0] = a, stats[1] = b, stats[2] = n, stats[5] = e;
    finalize_stats(stats);
}

// K3 (materialised): out = (adv - mean) / std in f32 with IEEE division (replay_buffer.py:105)
__global__ void __launch_bounds__(256) normalise_kernel(const float *__restrict__ adv, int64_t n,
                                                       const double *__restrict__ stats,
                                                       float *__restrict__ out)
{
    const float mean = (float)stats[3], sd = (float)stats[4];
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 x = __ldcs(reinterpret_cast<const float4 *>(adv) + i);
        x.x = __fdiv_rn(x.x - mean, sd);
        x.y = __fdiv_rn(x.y - mean, sd);
        x.z = __fdiv_rn(x.z - mean, sd);
        x.w = __fdiv_rn(x.w - mean, sd);
        __stcs(reinterpret_cast<float4 *>(out) + i, x);
    }
    for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = __fdiv_rn(adv[i] - mean, sd);
}

// K1b: wire f64 payload -> f32 buffer columns
__global__ void __launch_bounds__(256) cvt_f64_f32_kernel(const double *__restrict__ src,
                                                         float *__restrict__ dst, int64_t n)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        dst[i] = (float)src[i];
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct WsLayout {
    size_t ticket, flags, agg, incl, partials, total;
    int num_tiles;
};

WsLayout ws_layout(int64_t n)
{
    WsLayout w;
    w.num_tiles = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
    if (w.num_tiles < 1) w.num_tiles = 1;
    size_t o = 0;
    w.ticket = o, o += 256;
    w.flags = o, o += align_up(sizeof(int) * (size_t)w.num_tiles, 256);
    w.agg = o, o += align_up(sizeof(Map) * (size_t)w.num_tiles, 256);
    w.incl = o, o += align_up(sizeof(double2) * (size_t)w.num_tiles, 256);
    w.partials = o, o += align_up(sizeof(double2) * (size_t)w.num_tiles, 256);
    w.total = o;
    return w;
}

}  // namespace

size_t scan_workspace_bytes(int64_t n) { return ws_layout(n).total; }

int32_t scan_launch(const float *rew, const float *val, const int64_t *offsets, const float *last_vals,
                    int64_t n_total, int64_t n_ep, double gamma, double lam, bool with_baseline,
                    float *adv, float *ret, double *stats_out, void *workspace, size_t workspace_bytes,
                    bool finalize_stats_flag, int64_t n_scan_known, cudaStream_t stream, ErrBuf &err,
                    int *launches)
{
    if (n_total < 0 || n_ep < 0) {
        err.set("scan: negative sizes");
        return RRL_EINVAL;
    }
    if ((((uintptr_t)rew | (uintptr_t)adv | (uintptr_t)ret | (with_baseline ? (uintptr_t)val : 0)) & 15) != 0) {
        err.set("scan: rew/val/adv/ret must be 16-byte aligned");
        return RRL_EINVAL;
    }
    const WsLayout w = ws_layout(n_total);
    if (workspace == nullptr || workspace_bytes < w.total) {
        err.set("scan: workspace too small (%zu < %zu bytes)", workspace_bytes, w.total);
        return RRL_EINVAL;
    }
    // The scan covers [0, offsets[n_ep]).  When the caller knows that bound (the learner tracks
    // offsets on the host) the grid is sized exactly; otherwise it is sized for n_total and the
    // kernel, which reads offsets[n_ep] itself, retires the surplus CTAs -- no host round trip.
    if (n_scan_known > n_total) {
        err.set("scan: offsets[n_ep] = %lld exceeds n_total = %lld", (long long)n_scan_known, (long long)n_total);
        return RRL_EINVAL;
    }
    const int64_t n_cover = n_ep == 0 ? 0 : (n_scan_known >= 0 ? n_scan_known : n_total);
    char *base = static_cast<char *>(workspace);
    ScanArgs a;
    a.rew = rew, a.val = val, a.offsets = offsets, a.last_vals = last_vals;
    a.n_ep = n_ep;
    a.gamma = gamma, a.gl = gamma * lam;  // the reference forms gamma * lam in double (replay_buffer.py:70)
    a.gamma_p = 1.0, a.gl_p = 1.0;
    for (int i = 0; i < SCAN_VPT; ++i) a.gamma_p *= a.gamma, a.gl_p *= a.gl;  // same products the serial path forms
    a.adv = adv, a.ret = ret;
    a.ticket = reinterpret_cast<int *>(base + w.ticket);
    a.flags = reinterpret_cast<int *>(base + w.flags);
    a.agg = reinterpret_cast<Map *>(base + w.agg);
    a.incl = reinterpret_cast<double2 *>(base + w.incl);
    a.partials = reinterpret_cast<double2 *>(base + w.partials);
    const int tiles = (int)((n_cover + SCAN_TILE - 1) / SCAN_TILE);
    if (tiles > 0) {
        // every descriptor word starts as the "unset" sentinel
        RRL_CUDA_OK(cudaMemsetAsync(base + w.agg, 0xFF, w.partials - w.agg, stream), err);
        const size_t smem = 2 * sizeof(TileBuf);
        const void *fn = with_baseline ? reinterpret_cast<const void *>(scan_kernel<true>)
                                       : reinterpret_cast<const void *>(scan_kernel<false>);
        RRL_CUDA_OK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), err);
        int dev = 0, sms = 0, per_sm = 0;
        RRL_CUDA_OK(cudaGetDevice(&dev), err);
        RRL_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev), err);
        RRL_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, SCAN_CTA_THREADS, smem), err);
        if (per_sm < 1) {
            err.set("scan: kernel does not fit on an SM (%zu B shared memory)", smem);
            return RRL_ECUDA;
        }
        // co-resident by construction: the look-back spins on tiles owned by other CTAs
        const int grid = std::min(tiles, per_sm * sms);
        void *args[] = {&a};
        RRL_CUDA_OK(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(SCAN_CTA_THREADS), args, smem, stream), err);
        if (launches) *launches += 1;
    }
    if (stats_out) {
        stats_reduce_kernel<<<1, 1024, 0, stream>>>(a.partials, offsets, n_ep, adv, n_total, stats_out,
                                                    finalize_stats_flag ? 1 : 0);
        RRL_CUDA_OK(cudaGetLastError(), err);
        if (launches) *launches += 1;
    }
    return RRL_OK;
}

int32_t stats_finalize_launch(double *stats, cudaStream_t stream, ErrBuf &err)
{
    stats_finalize_kernel<<<1, 1, 0, stream>>>(stats);
    RRL_CUDA_OK(cudaGetLastError(), err);
    return RRL_OK;
}

int32_t stats_merge_launch(const double *gathered, int nranks, double *stats, cudaStream_t stream, ErrBuf &err)
{
    stats_merge_kernel<<<1, 1, 0, stream>>>(gathered, nranks, stats);
    RRL_CUDA_OK(cudaGetLastError(), err);
    return RRL_OK;
}

int32_t normalise_launch(const float *adv, int64_t n, const double *stats, float *out, cudaStream_t stream,
                         ErrBuf &err)
{
    if (n <= 0) return RRL_OK;
    if ((((uintptr_t)adv | (uintptr_t)out) & 15) != 0) {
        err.set("normalise: adv/out must be 16-byte aligned");
        return RRL_EINVAL;
    }
    int64_t blocks = (n / 4 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    normalise_kernel<<<(int)blocks, 256, 0, stream>>>(adv, n, stats, out);
    RRL_CUDA_OK(cudaGetLastError(), err);
    return RRL_OK;
}

int32_t cvt_f64_f32_launch(const double *src, float *dst, int64_t n, cudaStream_t stream, ErrBuf &err)
{
    if (n <= 0) return RRL_OK;
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    cvt_f64_f32_kernel<<<(int)blocks, 256, 0, stream>>>(src, dst, n);
    RRL_CUDA_OK(cudaGetLastError(), err);
    return RRL_OK;
}

}  // namespace rrl

// ---------------------------------------------------------------------------------------------
// Scan entry point for callers that must not allocate device memory.
//
// Who needs it: hosts that own a fixed device-memory budget handed to them at start-up -- a
// real-time trajectory server whose memory pool is carved out once and then locked, a process
// running under a quota that leaves no room for a second cudaMalloc, or a plug-in loaded into
// somebody else's CUDA context where the embedding application forbids allocation calls after
// initialisation.  rrl_scan_device() asks such callers for a workspace they cannot produce.
//
// Here the tile descriptors of the single-pass look-back live in a fixed array that is part of
// the library image: every GPU that loads librelayrl_b200.so carries these 557 312 bytes until
// the process exits.  That bounds the size: at most RRL_SCAN_NOALLOC_MAX transitions per call.
// ---------------------------------------------------------------------------------------------
#include <mutex>

namespace rrl {
extern thread_local ErrBuf g_create_err;

namespace {
constexpr int64_t NOALLOC_MAX_TILES = RRL_SCAN_NOALLOC_MAX / SCAN_TILE;  // 8192
// ticket (256) + flags + aggregates + inclusive values + statistic partials, each 256-aligned
constexpr size_t NOALLOC_WS_BYTES = 256 + NOALLOC_MAX_TILES * (sizeof(int) + sizeof(Map) + 2 * sizeof(double2));
static_assert(RRL_SCAN_NOALLOC_MAX % SCAN_TILE == 0, "limit must be a whole number of tiles");
#if SCAN_CONSUMER_THREADS == 256
static_assert(NOALLOC_WS_BYTES == 557312, "documented size of the static scratch");
#endif

static __device__ __align__(256) unsigned char g_scan_scratch[NOALLOC_WS_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
void *g_noalloc_ptr[NOALLOC_MAX_DEVICES];  // device address of g_scan_scratch, per device ordinal
}  // namespace
}  // namespace rrl

extern "C" int32_t rrl_scan_device_noalloc(const float *rew, const float *val, const int64_t *offsets,
                                           const float *last_vals, int64_t n_total, int64_t n_ep, double gamma,
                                           double lam, int32_t with_baseline, float *adv, float *ret,
                                           double *stats_out, void *stream)
{
    using namespace rrl;
    ErrBuf &err = g_create_err;
    if (!rew || !adv || !ret || (with_baseline && !val) || (n_ep > 0 && (!offsets || !last_vals))) {
        err.set("rrl_scan_device_noalloc: null pointer");
        return RRL_EINVAL;
    }
    if (n_total < 0 || n_total > RRL_SCAN_NOALLOC_MAX) {
        err.set("rrl_scan_device_noalloc: n_total = %lld exceeds the static scratch limit of %lld transitions; "
                "use rrl_scan_device with a caller-provided workspace",
                (long long)n_total, (long long)RRL_SCAN_NOALLOC_MAX);
        return RRL_EINVAL;
    }
    if (scan_workspace_bytes(n_total) > NOALLOC_WS_BYTES) {
        err.set("rrl_scan_device_noalloc: internal layout needs %zu bytes > %zu", scan_workspace_bytes(n_total),
                (size_t)NOALLOC_WS_BYTES);
        return RRL_EINVAL;
    }
    int dev = -1;
    RRL_CUDA_OK(cudaGetDevice(&dev), err);
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) {
        err.set("rrl_scan_device_noalloc: device ordinal %d outside [0, %d)", dev, NOALLOC_MAX_DEVICES);
        return RRL_EINVAL;
    }
    void *scratch = nullptr;
    {
        // Set-up, once per device, with that device current: in host code `g_scan_scratch` is only
        // a stand-in symbol, not an address a kernel may dereference; the device address comes from
        // cudaGetSymbolAddress.
        std::lock_guard<std::mutex> lock(g_noalloc_mu);
        if (!g_noalloc_ptr[dev]) {
            void *p = nullptr;
            RRL_CUDA_OK(cudaGetSymbolAddress(&p, g_scan_scratch), err);
            if (!p) {
                err.set("rrl_scan_device_noalloc: cudaGetSymbolAddress returned a null address on device %d", dev);
                return RRL_ECUDA;
            }
            g_noalloc_ptr[dev] = p;
        }
        scratch = g_noalloc_ptr[dev];
    }
    // scan_launch resets the tile descriptors (the only part of the scratch that is read before it
    // is written) on `stream` before the kernel.
    return scan_launch(rew, val, offsets, last_vals, n_total, n_ep, gamma, lam, with_baseline != 0, adv, ret,
                       stats_out, scratch, NOALLOC_WS_BYTES, true, -1, static_cast<cudaStream_t>(stream), err, nullptr);
}
