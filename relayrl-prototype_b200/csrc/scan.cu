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
// Algorithm: both recurrences y[t] = x[t] + c*y[t+1] are affine in the value carried in from the right, and an
// episode end resets the carry.  Because c is a constant, the slope of any run of whole threads / warps / tiles
// is a constant power of c (or 0 when a path ends inside it): only VALUES travel through shuffles, shared
// memory and the tile descriptors.  One pass over HBM (16 B/transition with baseline, 12 B without):
//   persistent co-resident CTAs walking 2048-transition tiles right-to-left in a fixed stride; a producer warp
//   prefetches the next tile while 8 consumer warps scan the current one
//     -> cp.async 16-byte copies into swizzled shared memory, blocked 8-per-thread reads
//     -> episode ends of the tile found by a 32-ary warp search in `offsets`
//     -> pass 1: thread-serial recurrences for a zero carry-in, the 16 results stay in registers
//     -> warp-shuffle suffix scan of the values, 8 warp values through shared memory
//     -> single-pass decoupled look-back across tiles (aggregate / inclusive descriptors whose
//        64-bit words carry their own validity: no flags, no fences on the critical path)
//     -> pass 2: one FMA per result adds (power of c) x carry, rounds to f32, 128-bit streaming stores
//        straight from registers
//     -> per-tile (sum, sum of squares) of the f32-rounded advantages in f64, fixed summation order.
// Arithmetic is float64 with float32 storage, as in the reference.
#include <cmath>

#include "common.cuh"

namespace rrl {

namespace {

constexpr int SCAN_THREADS = 256;                   // consumer threads (8 warps) ...
constexpr int SCAN_CTA_THREADS = SCAN_THREADS + 32;  // ... + one producer warp
constexpr int SCAN_VPT = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_VPT;  // 2048 transitions per tile
constexpr int SCAN_WARPS = SCAN_THREADS / 32;
constexpr int SCAN_SMEM_FLOATS = SCAN_TILE + SCAN_TILE / 8 + 8;
#ifndef SCAN_MIN_CTAS
#define SCAN_MIN_CTAS 3
#endif

// 4 floats of padding after every 32: blocked 8-per-thread LDS.128 reads are conflict-free
__device__ __forceinline__ int sidx(int j) { return j + ((j >> 5) << 2); }

// Tile descriptors carry their own validity: every 64-bit word starts as the sentinel 0xFFFF...F
// (memset before the launch) and is written once with a single 64-bit store.  A reader polls the
// words themselves (L2-coherent relaxed loads) -- no flag, no fence, no MEMBAR on the critical path.
// The sentinel is a NaN payload the arithmetic here cannot produce.
__device__ __forceinline__ bool is_unset(double x) { return __double_as_longlong(x) == -1ll; }
__device__ __forceinline__ double2 ld_relaxed2(const double2 *p)
{
    double2 v;
#ifdef SCAN_POLL_CG  // developer experiment: L2-cached weak load instead of the strong (relaxed.gpu) one
    asm volatile("ld.volatile.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
#else
    asm volatile("ld.relaxed.gpu.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
#endif
    return v;
}
__device__ __forceinline__ void st_relaxed2(double2 *p, double x, double y)
{
#ifdef SCAN_POLL_CG
    asm volatile("st.volatile.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(x), "d"(y) : "memory");
#else
    asm volatile("st.relaxed.gpu.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(x), "d"(y) : "memory");
#endif
}
__device__ __forceinline__ void st_stream4(float *p, float a, float b, float c, float d)
{
    asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// f32 <-> f64 with the hardware conversions (F2F): exact for every input incl. denormals, Inf and NaN, and -- measured
// on B200 with tools/cvt_rate.cu, profiles/r2_cvt_rate.txt -- 15 per clock per SM for both directions together, against
// 9.8 for the integer-ALU widening the round-1 kernel used (7 instructions + the select for zero).  At five conversions
// per transition that is 0.07 ms of the XU pipe for 64 Mi transitions, in parallel with the FP64 pipe and the loads.
__device__ __forceinline__ double widen(float x) { return (double)x; }
// acc += m * x where p holds: one predicated DFMA (the compiler's own choice is an unconditional DFMA + two FSELs)
__device__ __forceinline__ void fma_if(double &acc, double m, double x, bool p)
{
    asm("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %3, 0;\n\t@q fma.rn.f64 %0, %1, %2, %0;\n\t}" : "+d"(acc) : "d"(m), "d"(x), "r"((int)p));
}
__device__ __forceinline__ float narrow(double x) { return __double2float_rn(x); }

struct ScanArgs {
    const float *rew, *val;
    const int64_t *offsets;
    const float *last_vals;
    int64_t n_ep;     // closed paths; the scan covers [0, offsets[n_ep]) (read on the device)
    int64_t n_total;  // size of the buffers: offsets[n_ep] beyond it is a caller error the kernel refuses to follow
    double gamma, gl;
    // powers of the two discounts (c = gamma for returns, gamma*lam for advantages), formed on the host by repeated
    // multiplication: pw[j] = c^j (j = 0..8: an element k places from the right edge of its thread's run takes the
    // run's carry times c^k); st[s] = c^(8 * 2^s) (slope of 2^s whole threads: the warp scan's steps); wp = c^256
    // (one warp), tp = c^2048 (one tile)
    double pw_r[SCAN_VPT + 1], pw_a[SCAN_VPT + 1];
    double st_r[5], st_a[5];
    double wp_r, wp_a, tp_r, tp_a;
    float *adv, *ret;
    // workspace
    int *ticket;         // next logical tile (starts at -1: the memset that resets the descriptors covers it)
    double2 *agg;        // open tile: value at its first element for a zero carry-in (its slope is tp)
    double2 *incl;       // final (ret, adv) at the tile's first element
    double2 *partials;   // per tile: (sum, sum of squares) of the f32-rounded advantages
};

// ---- persistent, warp-specialised tile pipeline ------------------------------------------------
// A CTA is 8 consumer warps + 1 producer warp.
// While the consumers work on one tile the producer warp prepares the next one in the other buffer: cp.async of
// rew/val into swizzled shared memory (the bytes in flight that keep HBM busy), the 32-ary search for the tile's
// first path boundary, and the end-of-path bit mask + last_vals.
struct TileBuf {
    float r[SCAN_SMEM_FLOATS];   // rew
    float v[SCAN_SMEM_FLOATS];   // val (+ the first value of the next tile)
    float lv[SCAN_TILE];         // last_val at the positions where a path ends
    unsigned end[SCAN_TILE / 32];
    int tile;                    // physical tile index, -1 = no more work
    int pad[3];
};
static_assert(sizeof(TileBuf) % 16 == 0, "TileBuf must keep 16-byte alignment");

struct ScanCtl {  // consumer hand-offs inside one tile
    double2 wagg[SCAN_WARPS];      // per warp: value at its first element for a zero carry-in
    double2 wcarry[SCAN_WARPS];    // per warp: the carry that enters it from the right (after the look-back)
    double2 wstat[2][SCAN_WARPS];  // per warp: (sum, sum of squares); double buffered by tile parity
    unsigned wclosed[SCAN_WARPS];  // per warp: a path ends inside it
    double2 chain[2];              // chunk mode: (ret, adv) at the first element of the tile this CTA did last (by parity)
    double2 tpow[32];              // tp^i (ret, adv), i = 0..31: the look-back's per-lane slopes (set once per CTA)
};
// Slope below which a carry is dropped.  The carried value is at most max|x| / (1 - c) (<= 1e5 max|x| up to c = 0.99999),
// so what is dropped is below 1e-25 of the largest input: nine decades under the rounding error the float64 recurrence
// itself commits on every sum (1e-16 of its terms), i.e. invisible even to the float64 oracle, let alone after f32 rounding.
constexpr double SCAN_CUT = 1e-30;
constexpr double SCAN_CUT_HOST = SCAN_CUT;

#ifdef SCAN_TIMING  // developer build only: clock64 spans of CTA 0 (warp 1 of the consumers, lane 0; the producer's lane 0)
__device__ unsigned long long g_scan_timing[16];
#define SCT_DECL long long sct_prev = clock64()
#define SCT(slot, cond)                                                                          \
    do {                                                                                         \
        const long long sct_now = clock64();                                                     \
        if (blockIdx.x == gridDim.x / 2 && (cond)) atomicAdd(&g_scan_timing[slot], (unsigned long long)(sct_now - sct_prev)); \
        sct_prev = sct_now;                                                                      \
    } while (0)
#else
#define SCT_DECL
#define SCT(slot, cond)
#endif

__device__ __forceinline__ void cp_async16_zfill(void *smem_dst, const void *gmem_src, int src_bytes)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem_src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(SCAN_THREADS) : "memory"); }

// producer warp: claim the next tile and fill `B` for it (or mark it empty)
// Chunk mode: a CTA's tiles follow each other to the left, so the boundaries it needs are consecutive entries of `offsets`
// just below the previous tile's.  The producer keeps a 256-entry window of offsets / last_vals in shared memory and
// refills it (one round trip) when a tile's neighbourhood leaves it, instead of two or three dependent L2 round trips
// per tile.
constexpr int SCAN_OC = 256;
struct OffCache {
    long long off[SCAN_OC];
    float lv[SCAN_OC];
};
struct ProdState {
    long long cursor = 0;  // the previous tile's first boundary index (0: none yet)
    long long c0 = -1;     // offsets index of the window's first entry (-1: empty)
    OffCache *win = nullptr;
};

template <bool WB>
__device__ __forceinline__ void scan_prepare(const ScanArgs &a, TileBuf &B, int64_t n_scan, int num_tiles, int lane, int j,
                                             ProdState *ps = nullptr)
{
    long long *cursor = ps ? &ps->cursor : nullptr;
    const long long c0 = ps ? ps->c0 : -1;
    auto off_at = [&](long long k) -> long long {
        return (c0 >= 0 && k >= c0 && k < c0 + SCAN_OC) ? ps->win->off[k - c0] : a.offsets[k];
    };
    auto lv_at = [&](long long k) -> float {
        return (c0 >= 0 && k >= c0 && k < c0 + SCAN_OC) ? ps->win->lv[k - c0] : a.last_vals[k];
    };
    // j: logical tile, 0 is the right-most
    SCT_DECL;
    const int tile = j < num_tiles ? num_tiles - 1 - j : -1;
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
    if (WB && lane == 0) {  // the first value of the tile to the right (4-byte copy in the same group; zero beyond the end)
        const bool in = t0 + SCAN_TILE < n_scan;
        const unsigned sdst = (unsigned)__cvta_generic_to_shared(&B.v[sidx(SCAN_TILE)]);
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(sdst), "l"(a.val + (in ? t0 + SCAN_TILE : 0)), "r"(in ? 4 : 0) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    SCT(8, lane == 0);  // producer: issue

#ifdef SCAN_DBG_NO_SEARCH  // timing experiment only (wrong results)
    asm volatile("cp.async.wait_all;" ::: "memory");
    return;
#endif
    // first path boundary beyond t0: smallest e in [1, E] with offsets[e] > t0
    long long lo = 1, hi = a.n_ep;
    if (cursor && *cursor > 0) {
        // chunk mode: this tile lies directly to the left of the previous one, whose answer is *cursor -- walk down from
        // it (one round of 32 probes covers the usual case; the 32-ary search below is for a CTA's first tile only)
        lo = *cursor;
        while (true) {
            const long long p = lo - 1 - lane;
            const bool gt = p >= 1 && off_at(p) > t0;
            const unsigned b = __ballot_sync(0xffffffffu, gt);
            const int run = b == 0xffffffffu ? 32 : __ffs(~b) - 1;  // offsets are sorted: the probes that pass form a prefix
            lo -= run;
            if (run < 32) break;
        }
        hi = lo;
    }
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
    if (cursor) *cursor = lo;
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
                const int64_t b = off_at(k);
                if (b <= t1) {
                    if (u == 3) more = true;
                    if (b > off_at(k - 1)) {  // a non-empty path ends at b-1
                        const int pos = (int)(b - 1 - t0);
                        atomicOr(&B.end[pos >> 5], 1u << (pos & 31));
                        B.lv[pos] = lv_at(k - 1);
                    }
                }
            }
        }
        if (!__any_sync(0xffffffffu, more)) break;  // offsets are sorted: nothing further can fall in the tile
    }
    if (ps && (c0 < 0 || lo - 33 < c0 || lo + 128 > c0 + SCAN_OC)) {
        // the NEXT tile (to the left) probes [lo - 32, lo) and then reads up to 128 entries from its own answer upwards:
        // re-centre the window so that both stay inside for the next ~90 boundaries
        const long long n0 = lo - 120 > 0 ? lo - 120 : 0;
        __syncwarp();
        for (int i = lane; i < SCAN_OC; i += 32) {
            const long long k = n0 + i;
            ps->win->off[i] = k <= a.n_ep ? a.offsets[k] : 0x7fffffffffffffffll;
            ps->win->lv[i] = k < a.n_ep ? a.last_vals[k] : 0.f;
        }
        ps->c0 = n0;
        __syncwarp();
    }
    SCT(9, lane == 0);  // producer: search + mask
#ifndef SCAN_DBG_NO_WAIT  // (timing experiment only, wrong results: is the producer's wait for its data exposed?  measured: no)
    asm volatile("cp.async.wait_all;" ::: "memory");
#endif
    SCT(10, lane == 0);  // producer: wait for the data
}

// consumer warps (threads 0..255): scan the tile in `B`, publish / look back, write the outputs.
// lane_r / lane_a: c^(8 * (31 - lane)), the slope between the warp's right edge and this lane's right edge.
// CHUNK: the CTA walks a contiguous run of tiles, so the carry into a tile is the previous tile's value, kept in
// shared memory (S.chain) -- no descriptors, no polling.  `warm`: a warm-up tile to the right of the CTA's run: only its
// value at its first element is wanted (no outputs, no statistics).
template <bool WB, bool CHUNK>
__device__ __forceinline__ void scan_consume(const ScanArgs &a, TileBuf &B, int tile, int64_t n_scan, int num_tiles,
                                             ScanCtl &S, int tid, double lane_r, double lane_a, int parity, bool warm,
                                             double &acc1, double &acc2)
{
    const int lane = tid & 31, warp = tid >> 5;
    SCT_DECL;
    const bool sct_me = tid == 32;
    const float *r_s = B.r, *v_s = B.v;
    const int64_t t0 = (int64_t)tile * SCAN_TILE;
    const int cnt = (int)min((int64_t)SCAN_TILE, n_scan - t0);
    const bool full = cnt == SCAN_TILE;

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
    const unsigned eb = (B.end[j0 >> 5] >> (j0 & 31)) & 0xffu;  // bit k: a path ends at this thread's element k
    const bool open_right = ((B.end[(cnt - 1) >> 5] >> ((cnt - 1) & 31)) & 1u) == 0u;
    const unsigned em = __ballot_sync(0xffffffffu, eb != 0u);  // lanes of this warp that hold a path end

    // ---- pass 1: the run's values for a zero carry-in, kept in registers ----------------------------------
    // (float64 recurrences as the reference's lfilter calls, replay_buffer.py:63-77; pass 2 only adds the carry)
    double lr[SCAN_VPT], la[SCAN_VPT];
    {
        double ret_n = 0.0, adv_n = 0.0, v_n = WB ? widen(v_edge) : 0.0;
        if (em == 0u) {  // no path ends inside this warp's 256 transitions (warp-uniform)
#pragma unroll
            for (int k = SCAN_VPT - 1; k >= 0; --k) {
                const double rr = widen(r[k]);
                double dl = rr;
                if (WB) {
                    const double vk = widen(v[k]);
                    dl = fma(a.gamma, v_n, rr) - vk;
                    v_n = vk;
                }
                adv_n = fma(a.gl, adv_n, dl);
                ret_n = fma(a.gamma, ret_n, rr);
                la[k] = adv_n, lr[k] = ret_n;
            }
        } else {
#pragma unroll
            for (int k = SCAN_VPT - 1; k >= 0; --k) {
                if ((eb >> k) & 1u) {
                    // with baseline the marker's reward bootstraps both recurrences
                    // (replay_buffer.py:66-73); without it is ignored (replay_buffer.py:75-77)
                    const double lvk = widen(B.lv[j0 + k]);
                    ret_n = WB ? lvk : 0.0;
                    adv_n = 0.0;
                    v_n = lvk;
                }
                const double rr = widen(r[k]);
                double dl = rr;
                if (WB) {
                    const double vk = widen(v[k]);
                    dl = fma(a.gamma, v_n, rr) - vk;
                    v_n = vk;
                }
                adv_n = fma(a.gl, adv_n, dl);
                ret_n = fma(a.gamma, ret_n, rr);
                la[k] = adv_n, lr[k] = ret_n;
            }
        }
    }

    SCT(0, sct_me);  // loads + pass 1
    // ---- warp suffix scan: sr/sa = value at this lane's first element for a zero carry into the warp ----------
    // A step over d lanes multiplies by c^(8d) unless a path ends among the d lanes it bridges: the slope of an open
    // run of whole threads is a constant, so only the two values travel (two 64-bit shuffles per step).
    const unsigned above = em >> lane;                          // bit 0: this lane
    const int dist = above ? __ffs(above) - 1 : 32;             // lanes from here to the first one holding a path end
    double sr = lr[0], sa = la[0];
#pragma unroll
    for (int s = 0; s < 5; ++s) {
        const int d = 1 << s;
        const double o_r = __shfl_down_sync(0xffffffffu, sr, d), o_a = __shfl_down_sync(0xffffffffu, sa, d);
        const bool on = lane + d < 32 && dist >= d;
        fma_if(sr, a.st_r[s], o_r, on), fma_if(sa, a.st_a[s], o_a, on);
    }
    double er = __shfl_down_sync(0xffffffffu, sr, 1), ea = __shfl_down_sync(0xffffffffu, sa, 1);  // the lanes to the right
    if (lane == 31) er = 0.0, ea = 0.0;
    const bool lane_open = (above >> 1) == 0u;  // no path ends to the right of this lane inside the warp
    if (lane == 0) S.wagg[warp] = make_double2(sr, sa), S.wclosed[warp] = em;
    SCT(1, sct_me);  // warp scan
    consumer_sync();
    SCT(2, sct_me);  // barrier 1

    // ---- the eight warps; look-back kernel: warp 0 alone, with the tile's descriptor and the look-back; chunk kernel:
    // every warp for itself (a dozen instructions), which saves the second barrier and the wait for warp 0 -------------
    double wc_r = 0.0, wc_a = 0.0;  // the carry that enters this warp from the right
    if (CHUNK) {
        // every warp walks the (at most seven) warps to its right, starting from the carry the previous tile left
        const double2 prev = S.chain[parity ^ 1];
        if (open_right) wc_r = prev.x, wc_a = prev.y;
        for (int w2 = SCAN_WARPS - 1; w2 > warp; --w2) {
            const double2 g = S.wagg[w2];
            const bool cl = S.wclosed[w2] != 0u;
            wc_r = cl ? g.x : fma(a.wp_r, wc_r, g.x), wc_a = cl ? g.y : fma(a.wp_a, wc_a, g.y);
        }
        if (warp == 0 && lane == 0)  // the value at this tile's first element: the next tile's carry
            S.chain[parity] = em ? make_double2(sr, sa) : make_double2(fma(a.wp_r, wc_r, sr), fma(a.wp_a, wc_a, sa));
    } else if (warp == 0) {
        double xr = 0.0, xa = 0.0;
        if (lane < SCAN_WARPS) xr = S.wagg[lane].x, xa = S.wagg[lane].y;
        const unsigned cm = __ballot_sync(0xffffffffu, lane < SCAN_WARPS && S.wclosed[lane] != 0u);  // warps with a path end
        const unsigned wabove = cm >> lane;
        const int wdist = wabove ? __ffs(wabove) - 1 : 32;
        double pr = a.wp_r, pa = a.wp_a;
#pragma unroll
        for (int d = 1; d < SCAN_WARPS; d <<= 1) {
            const double o_r = __shfl_down_sync(0xffffffffu, xr, d), o_a = __shfl_down_sync(0xffffffffu, xa, d);
            if (lane + d < SCAN_WARPS && wdist >= d) xr = fma(pr, o_r, xr), xa = fma(pa, o_a, xa);
            pr *= pr, pa *= pa;
        }
        // lane 0: the whole tile
        const bool closed = cm != 0u;  // the value at the tile's first element does not depend on the carry-in
        if (lane == 0) {
            if (closed)
                st_relaxed2(&a.incl[tile], xr, xa);
            else
                st_relaxed2(&a.agg[tile], xr, xa);
        }
        double cr = 0.0, ca = 0.0;  // carry into the tile from the tiles to the right
#ifdef SCAN_DBG_NO_LOOKBACK  // timing experiment only (wrong results)
        if (false) {
#else
        if (open_right) {          // the right-most path of this tile continues in the next tile
#endif
            const double lt_r = S.tpow[lane].x, lt_a = S.tpow[lane].y;  // tp^lane
            double sc_r = 1.0, sc_a = 1.0;                               // tp^(32 * windows done)
            int base = tile + 1;
            while (true) {
                const int k = base + lane;
                double qr = 0.0, qa = 0.0;
                bool final_value = true;  // beyond the last tile there is nothing to carry
                // A tile whose slope towards this one has fallen below SCAN_CUT (see there) cannot change the result.
                // Such lanes are not polled and end the look-back: with gamma = 0.98 (tp = 1e-18) two tiles are read
                // instead of every tile in flight to the right (up to ~440: the CTAs move in step, so their inclusive
                // values appear together, late).
                const bool needed = fmax(sc_r * lt_r, sc_a * lt_a) >= SCAN_CUT;
                if (k < num_tiles && needed) {
                    while (true) {
                        const double2 iv = ld_relaxed2(&a.incl[k]);
                        if (!is_unset(iv.x) && !is_unset(iv.y)) {  // final value of that tile's first element
                            qr = iv.x, qa = iv.y;
                            break;
                        }
                        const double2 av = ld_relaxed2(&a.agg[k]);
                        if (!is_unset(av.x) && !is_unset(av.y)) {  // an open tile: value for a zero carry-in, slope tp
                            qr = av.x, qa = av.y, final_value = false;
                            break;
                        }
                    }
                }
                const unsigned fin = __ballot_sync(0xffffffffu, final_value);
                const int f = fin ? __ffs(fin) - 1 : 32;  // the window is used up to and including lane f
                double wr = lane <= f ? qr * lt_r : 0.0, wa = lane <= f ? qa * lt_a : 0.0;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) {
                    wr += __shfl_xor_sync(0xffffffffu, wr, d);
                    wa += __shfl_xor_sync(0xffffffffu, wa, d);
                }
                cr = fma(sc_r, wr, cr), ca = fma(sc_a, wa, ca);
                if (f < 32) break;
                sc_r *= S.tpow[31].x * a.tp_r, sc_a *= S.tpow[31].y * a.tp_a;
                base += 32;
            }
        }
        if (lane == 0 && !closed) st_relaxed2(&a.incl[tile], fma(a.tp_r, cr, xr), fma(a.tp_a, ca, xa));
        // carry into warp w: the warps to its right for a zero tile carry, plus the tile carry across them when open
        double nr = __shfl_down_sync(0xffffffffu, xr, 1), na = __shfl_down_sync(0xffffffffu, xa, 1);
        if (lane < SCAN_WARPS) {
            if (lane == SCAN_WARPS - 1) nr = 0.0, na = 0.0;
            if ((wabove >> 1) == 0u) {  // no path ends in the warps to the right of this one
                double qr = 1.0, qa = 1.0;
                for (int i = lane; i < SCAN_WARPS - 1; ++i) qr *= a.wp_r, qa *= a.wp_a;
                nr = fma(qr, cr, nr), na = fma(qa, ca, na);
            }
            S.wcarry[lane] = make_double2(nr, na);
        }
    }
    if (!CHUNK) consumer_sync();
    SCT(3, sct_me);  // warp 0's combine + look-back (as seen by a waiting warp)
    if (CHUNK && warm) return;

    // ---- pass 2: add the carry, round to f32, statistics, stores straight from registers -------------------------
    double s1 = 0.0, s2 = 0.0;
    {
        double cr = er, ca = ea;
        if (lane_open) {
            if (!CHUNK) {
                const double2 wc = S.wcarry[warp];
                wc_r = wc.x, wc_a = wc.y;
            }
            cr = fma(lane_r, wc_r, cr), ca = fma(lane_a, wc_a, ca);
        }
        float ao[SCAN_VPT], ro[SCAN_VPT];
        if (em == 0u && full) {
#pragma unroll
            for (int k = 0; k < SCAN_VPT; ++k) {
                ao[k] = narrow(fma(a.pw_a[SCAN_VPT - k], ca, la[k]));
                ro[k] = narrow(fma(a.pw_r[SCAN_VPT - k], cr, lr[k]));
                const double x = widen(ao[k]);  // the f32-rounded advantage, as the reference's buffer holds it
                s1 += x;
                s2 = fma(x, x, s2);
            }
        } else {
            const int kmin = eb ? 32 - __clz(eb) : 0;  // elements at or left of the run's last path end take no carry
#pragma unroll
            for (int k = 0; k < SCAN_VPT; ++k) {
                double ya = la[k], yr = lr[k];
                if (k >= kmin) ya = fma(a.pw_a[SCAN_VPT - k], ca, ya), yr = fma(a.pw_r[SCAN_VPT - k], cr, yr);
                ao[k] = narrow(ya), ro[k] = narrow(yr);
                if (j0 + k < cnt) {
                    const double x = widen(ao[k]);
                    s1 += x;
                    s2 = fma(x, x, s2);
                }
            }
        }
        SCT(6, sct_me);  // pass 2: arithmetic
        float *ap = a.adv + t0 + j0, *rp = a.ret + t0 + j0;
        if (full || j0 + SCAN_VPT <= cnt) {
            st_stream4(ap, ao[0], ao[1], ao[2], ao[3]);
            st_stream4(ap + 4, ao[4], ao[5], ao[6], ao[7]);
            st_stream4(rp, ro[0], ro[1], ro[2], ro[3]);
            st_stream4(rp + 4, ro[4], ro[5], ro[6], ro[7]);
        } else {
#pragma unroll
            for (int k = 0; k < SCAN_VPT; ++k)
                if (j0 + k < cnt) ap[k] = ao[k], rp[k] = ro[k];
        }
    }
    SCT(7, sct_me);  // pass 2: stores
    if (CHUNK) {  // the CTA's tiles are its own: the sums ride in registers until its run ends
        acc1 += s1, acc2 += s2;
        return;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        s1 += __shfl_xor_sync(0xffffffffu, s1, d);
        s2 += __shfl_xor_sync(0xffffffffu, s2, d);
    }
    if (lane == 0) S.wstat[parity][warp] = make_double2(s1, s2);
    SCT(4, sct_me);  // pass 2 + stores + statistics
}

template <bool WB>
__global__ void __launch_bounds__(SCAN_CTA_THREADS, SCAN_MIN_CTAS) scan_kernel(ScanArgs a)
{
    extern __shared__ __align__(16) unsigned char scan_dyn[];
    TileBuf *buf = reinterpret_cast<TileBuf *>(scan_dyn);
    __shared__ ScanCtl ctl;

    const int tid = threadIdx.x, lane = tid & 31;
    const bool producer = tid >= SCAN_THREADS;
    const int64_t n_scan = a.offsets[a.n_ep];  // end of the last closed path
    if (n_scan > a.n_total || n_scan < 0) return;  // inconsistent offsets: stats_reduce_kernel reports it (NaN statistics)
    const int num_tiles = (int)((n_scan + SCAN_TILE - 1) / SCAN_TILE);

    double lane_r = 1.0, lane_a = 1.0;  // c^(8 * (31 - lane))
    for (int i = lane; i < 31; ++i) lane_r *= a.st_r[0], lane_a *= a.st_a[0];

    if (tid < 32) {
        double tr = 1.0, ta = 1.0;
        for (int i = 0; i < lane; ++i) tr *= a.tp_r, ta *= a.tp_a;
        ctl.tpow[lane] = make_double2(tr, ta);
    }
    // CTA c owns the logical tiles c, c + G, c + 2G, ...: the launch is cooperative (all G CTAs co-resident), the CTAs
    // move through their tiles in step, and with the look-back cut-off a tile only ever waits for the few tiles next to
    // it, which the neighbouring CTAs publish at the same moment.  (Tiles handed out by a ticket counter were tried:
    // the producer has to claim a tile one step before its CTA starts it, and that skew made the look-back wait
    // erratically -- 0.6 to 3.6 ms for the same input.)
    const int G = gridDim.x;
    int j = blockIdx.x;
    if (producer) scan_prepare<WB>(a, buf[0], n_scan, num_tiles, lane, j);
    __syncthreads();
    int b = 0, prev_tile = -1;
    double unused1 = 0.0, unused2 = 0.0;
    while (true) {
        const int tile = buf[b].tile;
        // the previous tile's statistics: its eight warp sums in a fixed order (the result does not depend on which
        // CTA took the tile)
        if (tid == 0 && prev_tile >= 0) {
            double t1 = 0.0, t2 = 0.0;
            for (int w = 0; w < SCAN_WARPS; ++w) t1 += ctl.wstat[b ^ 1][w].x, t2 += ctl.wstat[b ^ 1][w].y;
            a.partials[prev_tile] = make_double2(t1, t2);
        }
        if (tile < 0) break;
        j += G;
        if (producer)
            scan_prepare<WB>(a, buf[b ^ 1], n_scan, num_tiles, lane, j);
        else
            scan_consume<WB, false>(a, buf[b], tile, n_scan, num_tiles, ctl, tid, lane_r, lane_a, b, false, unused1, unused2);
        SCT_DECL;
        __syncthreads();
        SCT(5, tid == 32);                // consumers waiting for the producer
        SCT(11, tid == SCAN_THREADS);     // producer waiting for the consumers
        SCT(12, tid == 32);               // (tile count: every span adds ~0, used as a divisor via slot 13)
#ifdef SCAN_TIMING
        if (blockIdx.x == gridDim.x / 2 && tid == 32) atomicAdd(&g_scan_timing[13], 1ull);
#endif
        prev_tile = tile;
        b ^= 1;
    }
}

// Chunk mode: CTA c owns the contiguous logical tiles [c K, (c + 1) K) and walks them right-to-left, so the carry
// from tile to tile never leaves the CTA.  What comes in from the CTA's right-hand neighbour is recomputed instead of
// waited for: the CTA first runs `warm` extra tiles to the right of its run through pass 1 only.  After W open tiles the
// influence of anything further right is slope^W < SCAN_CUT (below the float64 recurrence's own rounding, see SCAN_CUT),
// and a path end inside the warm-up makes it exact.  For gamma = 0.98
// that is 2 tiles of reads for every ~74 tiles of work, no descriptors, no polling, no memset, no co-residency.  The
// look-back kernel above stays for discounts so close to 1 that the warm-up would outgrow the run (scan_launch picks).
template <bool WB>
__global__ void __launch_bounds__(SCAN_CTA_THREADS, SCAN_MIN_CTAS) scan_chunk_kernel(ScanArgs a, int tiles_per_cta, int warm)
{
    extern __shared__ __align__(16) unsigned char scan_dyn[];
    TileBuf *buf = reinterpret_cast<TileBuf *>(scan_dyn);
    __shared__ ScanCtl ctl;

    const int tid = threadIdx.x, lane = tid & 31;
    const bool producer = tid >= SCAN_THREADS;
    const int64_t n_scan = a.offsets[a.n_ep];
    if (n_scan > a.n_total || n_scan < 0) return;
    const int num_tiles = (int)((n_scan + SCAN_TILE - 1) / SCAN_TILE);
    const int j_begin = blockIdx.x * tiles_per_cta, j_end = min(j_begin + tiles_per_cta, num_tiles);
    if (j_begin >= num_tiles) return;
    int j = max(j_begin - warm, 0);  // logical tile (0 = right-most); tiles before j_begin are warm-up

    double lane_r = 1.0, lane_a = 1.0;  // c^(8 * (31 - lane))
    for (int i = lane; i < 31; ++i) lane_r *= a.st_r[0], lane_a *= a.st_a[0];
    if (tid == 0) ctl.chain[0] = ctl.chain[1] = make_double2(0.0, 0.0);
    __shared__ OffCache win;
    ProdState ps;
    ps.win = &win;

    if (producer) scan_prepare<WB>(a, buf[0], n_scan, j < j_end ? num_tiles : 0, lane, j, &ps);
    __syncthreads();
    int b = 0;
    double acc1 = 0.0, acc2 = 0.0;  // this thread's (sum, sum of squares) of the f32-rounded advantages over the run
    while (true) {
        const int tile = buf[b].tile;
        if (tile < 0) break;
        const bool is_warm = j < j_begin;
        ++j;
        if (producer)
            scan_prepare<WB>(a, buf[b ^ 1], n_scan, j < j_end ? num_tiles : 0, lane, j, &ps);
        else
            scan_consume<WB, true>(a, buf[b], tile, n_scan, num_tiles, ctl, tid, lane_r, lane_a, b, is_warm, acc1, acc2);
        __syncthreads();
        b ^= 1;
    }
    // the run's statistics, reduced once in a fixed order (the tile -> CTA map is static, so the result is reproducible):
    // the CTA's first tile carries the total, its other tiles zero (stats_reduce_kernel sums per-tile entries)
    if (!producer) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            acc1 += __shfl_xor_sync(0xffffffffu, acc1, d);
            acc2 += __shfl_xor_sync(0xffffffffu, acc2, d);
        }
        if (lane == 0) ctl.wstat[0][tid >> 5] = make_double2(acc1, acc2);
    }
    __syncthreads();
    for (int jj = j_begin + tid; jj < j_end; jj += SCAN_CTA_THREADS) {
        double2 t = make_double2(0.0, 0.0);
        if (jj == j_begin)
            for (int w = 0; w < SCAN_WARPS; ++w) t.x += ctl.wstat[0][w].x, t.y += ctl.wstat[0][w].y;
        a.partials[num_tiles - 1 - jj] = t;
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
    const int tid = threadIdx.x;
    if (n_scan > n_total || n_scan < 0) {  // offsets that do not fit the buffers: the scan kernel refused to run
        if (tid == 0) {
            const double nan = __longlong_as_double(0x7ff8000000000000ll);
            stats[0] = nan, stats[1] = nan, stats[2] = (double)n_total, stats[3] = nan, stats[4] = nan;
        }
        return;
    }
    const int num_tiles = (int)((n_scan + SCAN_TILE - 1) / SCAN_TILE);
    __shared__ double s_a[1024], s_b[1024];
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
    size_t ticket, agg, incl, partials, total;
    int num_tiles;
};

WsLayout ws_layout(int64_t n)
{
    WsLayout w;
    w.num_tiles = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
    if (w.num_tiles < 1) w.num_tiles = 1;
    size_t o = 0;
    w.ticket = o, o += 256;
    w.agg = o, o += align_up(sizeof(double2) * (size_t)w.num_tiles, 256);
    w.incl = o, o += align_up(sizeof(double2) * (size_t)w.num_tiles, 256);
    w.partials = o, o += align_up(sizeof(double2) * (size_t)w.num_tiles, 256);
    w.total = o;
    return w;
}

}  // namespace

#ifdef SCAN_TIMING
extern "C" int32_t rrl_debug_scan_timing(unsigned long long *out16, int32_t reset)
{
    if (out16 && cudaMemcpyFromSymbol(out16, g_scan_timing, sizeof(g_scan_timing)) != cudaSuccess) return RRL_ECUDA;
    if (reset) {
        unsigned long long z[16] = {};
        if (cudaMemcpyToSymbol(g_scan_timing, z, sizeof(z)) != cudaSuccess) return RRL_ECUDA;
    }
    return RRL_OK;
}
#endif

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
    a.n_ep = n_ep, a.n_total = n_total;
    a.gamma = gamma, a.gl = gamma * lam;  // the reference forms gamma * lam in double (replay_buffer.py:70)
    // powers by repeated multiplication / squaring (the products a serial evaluation would form)
    auto powers = [](double c, double *pw, double *st, double &wp, double &tp) {
        pw[0] = 1.0;
        for (int i = 1; i <= SCAN_VPT; ++i) pw[i] = pw[i - 1] * c;
        st[0] = pw[SCAN_VPT];
        for (int i = 1; i < 5; ++i) st[i] = st[i - 1] * st[i - 1];
        wp = st[4] * st[4];                 // 32 threads
        tp = wp;
        for (int w = 1; w < SCAN_WARPS; w <<= 1) tp *= tp;  // 8 warps
    };
    static_assert(SCAN_WARPS == 8, "tp = wp^8 is formed by three squarings");
    powers(a.gamma, a.pw_r, a.st_r, a.wp_r, a.tp_r);
    powers(a.gl, a.pw_a, a.st_a, a.wp_a, a.tp_a);
    a.adv = adv, a.ret = ret;
    a.ticket = reinterpret_cast<int *>(base + w.ticket);
    a.agg = reinterpret_cast<double2 *>(base + w.agg);
    a.incl = reinterpret_cast<double2 *>(base + w.incl);
    a.partials = reinterpret_cast<double2 *>(base + w.partials);
    const int tiles = (int)((n_cover + SCAN_TILE - 1) / SCAN_TILE);
    if (tiles > 0) {
        const size_t smem = 2 * sizeof(TileBuf);
        static int per_sm_cached[4] = {0, 0, 0, 0}, sms_cached = 0;  // [chunk? 2 : 0] + with_baseline
        auto occupancy = [&](const void *fn, int slot) -> int32_t {
            if (per_sm_cached[slot]) return RRL_OK;
            RRL_CUDA_OK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), err);
            int dev = 0, sms = 0, per_sm = 0;
            RRL_CUDA_OK(cudaGetDevice(&dev), err);
            RRL_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev), err);
            RRL_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, SCAN_CTA_THREADS, smem), err);
            if (per_sm < 1) {
                err.set("scan: kernel does not fit on an SM (%zu B shared memory)", smem);
                return RRL_ECUDA;
            }
            sms_cached = sms, per_sm_cached[slot] = per_sm;
            return RRL_OK;
        };
        // Warm-up length of the chunk kernel: open tiles after which a carry has decayed below SCAN_CUT.
        int warm = 1 << 30;
        {
            const double tp = std::max(a.tp_r, a.tp_a);
            if (tp <= 0.0) {
                warm = 1;
            } else if (tp < 1.0) {
                const double w_ = std::ceil(std::log(SCAN_CUT_HOST) / std::log(tp));
                if (w_ < 1e6) warm = std::max(1, (int)w_);
            }
        }
        const int wb = with_baseline ? 1 : 0;
        const void *fn_chunk = with_baseline ? reinterpret_cast<const void *>(scan_chunk_kernel<true>)
                                             : reinterpret_cast<const void *>(scan_chunk_kernel<false>);
        RRL_TRY(occupancy(fn_chunk, 2 + wb));
        const int max_ctas = per_sm_cached[2 + wb] * sms_cached;
        // a run is at least four warm-ups long (the warm-up then reads at most a quarter on top), even if that leaves
        // CTA slots empty at small sizes; discounts whose warm-up exceeds 64 tiles (gamma > ~0.998) take the look-back
        const int per_cta = std::max((tiles + max_ctas - 1) / max_ctas, std::min(4 * warm, tiles));
#ifdef SCAN_FORCE_LOOKBACK
        const bool chunk = false;
#else
        const bool chunk = warm <= 64;
#endif
        if (chunk) {
            const int grid = (tiles + per_cta - 1) / per_cta;
            if (with_baseline)
                scan_chunk_kernel<true><<<grid, SCAN_CTA_THREADS, smem, stream>>>(a, per_cta, warm);
            else
                scan_chunk_kernel<false><<<grid, SCAN_CTA_THREADS, smem, stream>>>(a, per_cta, warm);
            RRL_CUDA_OK(cudaGetLastError(), err);
        } else {
            // every descriptor word starts as the "unset" sentinel
            RRL_CUDA_OK(cudaMemsetAsync(base, 0xFF, w.partials, stream), err);
            const void *fn = with_baseline ? reinterpret_cast<const void *>(scan_kernel<true>)
                                           : reinterpret_cast<const void *>(scan_kernel<false>);
            RRL_TRY(occupancy(fn, wb));
            // co-resident by construction: the look-back spins on tiles owned by other CTAs
            const int grid = std::min(tiles, per_sm_cached[wb] * sms_cached);
            void *args[] = {&a};
            RRL_CUDA_OK(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(SCAN_CTA_THREADS), args, smem, stream), err);
        }
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
// the library image: every GPU that loads librelayrl_b200.so carries these 393 472 bytes until
// the process exits.  That bounds the size: at most RRL_SCAN_NOALLOC_MAX transitions per call.
// ---------------------------------------------------------------------------------------------
#include <mutex>

namespace rrl {
extern thread_local ErrBuf g_create_err;

namespace {
constexpr int64_t NOALLOC_MAX_TILES = RRL_SCAN_NOALLOC_MAX / SCAN_TILE;  // 8192
// ticket (256) + aggregates + inclusive values + statistic partials, each 256-aligned
constexpr size_t NOALLOC_WS_BYTES = 256 + NOALLOC_MAX_TILES * 3 * sizeof(double2);
static_assert(RRL_SCAN_NOALLOC_MAX % SCAN_TILE == 0, "limit must be a whole number of tiles");
static_assert(NOALLOC_WS_BYTES == 393472, "documented size of the static scratch");

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
