// 5/3 reversible and 9/7 irreversible DWT for sm_100a: one decomposition level per launch, vertical
// and horizontal lifting fused, no shared memory.
//
// Work decomposition: one warp owns a strip of 120 columns (lanes 1..30, four adjacent columns per
// lane, one 16-byte access per lane and row; lanes 0 and 31 carry the 4-column halos) and a segment
// of SEG row pairs.  It streams down the rows once: the vertical lifting runs as a register
// recurrence per column (each input row is loaded exactly once per warp, plus a short warm-up), the
// horizontal lifting of each finished row runs across the lanes with one warp shuffle per lifting
// step, and the four sub-band pieces go out as 8-byte stores.  Boundaries use whole-sample symmetric
// extension by index reflection, which is what the reference's index clamping computes
// (ext/grok/src/lib/openjp2/dwt.cpp:182-188, 313-399, 769-921).
//
// Order of the two directions follows the reference: forward = vertical then horizontal
// (dwt.cpp:420-506), inverse = horizontal then vertical (dwt.cpp:615-704, 927-1062).
#include "device.h"

namespace j2k {
namespace {

constexpr int STRIP_COLS = 120;
constexpr int WARPS_PER_CTA = 4;

__device__ __forceinline__ int mirror(int i, int lo, int hi) {
    if (hi - lo == 1) return lo;
    while (i < lo || i >= hi) {
        if (i < lo) i = 2 * lo - i;
        if (i >= hi) i = 2 * (hi - 1) - i;
    }
    return i;
}

__device__ __forceinline__ int fm(int a, int b) {          // opj_intmath.h:255-266
    return (int)(((long long)a * (long long)b + 4096) >> 13);
}

template <typename T> struct Q { T v[4]; };

// ---- lifting kernels (forward) -----------------------------------------------------------------
// Each provides: vertical streaming step (new odd row xo = x(n-1), new even row xe = x(n); emits
// the finished pair lo = row n-LAG, hi = row n-LAG+1) and the horizontal lifting across lanes.
struct Fwd53 {
    static constexpr int LAG = 2, WARM = 1;
    Q<int> xprev, dprev;      // x(n-2), d(n-3)
    __device__ void init(const Q<int>& x) {
        xprev = x;
#pragma unroll
        for (int k = 0; k < 4; ++k) dprev.v[k] = 0;
    }
    __device__ void step(const Q<int>& xo, const Q<int>& xe, Q<int>& lo, Q<int>& hi) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int d = xo.v[k] - ((xprev.v[k] + xe.v[k]) >> 1);
            lo.v[k] = xprev.v[k] + ((dprev.v[k] + d + 2) >> 2);
            hi.v[k] = d;
            dprev.v[k] = d;
            xprev.v[k] = xe.v[k];
        }
    }
    static __device__ void horizontal(Q<int>& r) {
        int nx = __shfl_down_sync(0xffffffffu, r.v[0], 1);
        r.v[1] -= (r.v[0] + r.v[2]) >> 1;
        r.v[3] -= (r.v[2] + nx) >> 1;
        int pv = __shfl_up_sync(0xffffffffu, r.v[3], 1);
        r.v[0] += (pv + r.v[1] + 2) >> 2;
        r.v[2] += (r.v[1] + r.v[3] + 2) >> 2;
    }
    static __device__ int single(int x, int odd) { return odd ? x * 2 : x; }   // dwt.cpp:318,323
};

struct Fwd97 {                 // Q13 fixed point, dwt.cpp:365-399
    static constexpr int LAG = 4, WARM = 3;
    Q<int> x2, d1, s1, d2;     // x(n-2), d1(n-3), s1(n-4), d2(n-5)
    __device__ void init(const Q<int>& x) {
        x2 = x;
#pragma unroll
        for (int k = 0; k < 4; ++k) d1.v[k] = s1.v[k] = d2.v[k] = 0;
    }
    __device__ void step(const Q<int>& xo, const Q<int>& xe, Q<int>& lo, Q<int>& hi) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int nd1 = xo.v[k] - fm(x2.v[k] + xe.v[k], 12993);
            int ns1 = x2.v[k] - fm(d1.v[k] + nd1, 434);
            int nd2 = d1.v[k] + fm(s1.v[k] + ns1, 7233);
            int ns2 = s1.v[k] + fm(d2.v[k] + nd2, 3633);
            lo.v[k] = fm(ns2, 6659);
            hi.v[k] = fm(nd2, 5038);
            x2.v[k] = xe.v[k]; d1.v[k] = nd1; s1.v[k] = ns1; d2.v[k] = nd2;
        }
    }
    static __device__ void horizontal(Q<int>& r) {
        const unsigned m = 0xffffffffu;
        int nx = __shfl_down_sync(m, r.v[0], 1);
        r.v[1] -= fm(r.v[0] + r.v[2], 12993);
        r.v[3] -= fm(r.v[2] + nx, 12993);
        int pv = __shfl_up_sync(m, r.v[3], 1);
        r.v[0] -= fm(pv + r.v[1], 434);
        r.v[2] -= fm(r.v[1] + r.v[3], 434);
        nx = __shfl_down_sync(m, r.v[0], 1);
        r.v[1] += fm(r.v[0] + r.v[2], 7233);
        r.v[3] += fm(r.v[2] + nx, 7233);
        pv = __shfl_up_sync(m, r.v[3], 1);
        r.v[0] += fm(pv + r.v[1], 3633);
        r.v[2] += fm(r.v[1] + r.v[3], 3633);
        r.v[0] = fm(r.v[0], 6659); r.v[2] = fm(r.v[2], 6659);
        r.v[1] = fm(r.v[1], 5038); r.v[3] = fm(r.v[3], 5038);
    }
    static __device__ int single(int x, int) { return x; }                     // dwt.cpp:368
};

// ---- loads / stores ----------------------------------------------------------------------------
struct Region {
    int x0, y0, x1, y1, pitch;
    int snx, sny;             // number of low-pass columns / rows
};

// four samples at absolute columns a0..a0+3 of memory row `row` (already reflected) of an
// interleaved region
template <typename T>
__device__ __forceinline__ Q<T> load_quad(const T* __restrict__ base, const Region& g, int row, int a0, bool vec_ok) {
    Q<T> q;
    const T* p = base + (size_t)row * g.pitch;
    if (vec_ok && a0 >= g.x0 && a0 + 3 < g.x1) {
        int4 t = __ldg(reinterpret_cast<const int4*>(p + (a0 - g.x0)));
        q.v[0] = *reinterpret_cast<T*>(&t.x); q.v[1] = *reinterpret_cast<T*>(&t.y);
        q.v[2] = *reinterpret_cast<T*>(&t.z); q.v[3] = *reinterpret_cast<T*>(&t.w);
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) q.v[k] = __ldg(p + (mirror(a0 + k, g.x0, g.x1) - g.x0));
    }
    return q;
}

// store the low pair (columns a0, a0+2) and high pair (a0+1, a0+3) of a finished row into the
// Mallat row `mrow` of dst
template <typename T>
__device__ __forceinline__ void store_split(T* __restrict__ dst, const Region& g, int mrow, int a0, const Q<T>& r) {
    T* p = dst + (size_t)mrow * g.pitch;
    int li = (a0 >> 1) - ((g.x0 + 1) >> 1);
    int hi = (a0 >> 1) - (g.x0 >> 1);
    T* pl = p + li;
    T* ph = p + g.snx + hi;
    bool in0 = a0 >= g.x0 && a0 < g.x1, in1 = a0 + 1 >= g.x0 && a0 + 1 < g.x1;
    bool in2 = a0 + 2 >= g.x0 && a0 + 2 < g.x1, in3 = a0 + 3 >= g.x0 && a0 + 3 < g.x1;
    if (in0 && in2 && ((reinterpret_cast<uintptr_t>(pl) & 7) == 0)) {
        *reinterpret_cast<int2*>(pl) = make_int2(*reinterpret_cast<const int*>(&r.v[0]), *reinterpret_cast<const int*>(&r.v[2]));
    } else {
        if (in0) pl[0] = r.v[0];
        if (in2) pl[1] = r.v[2];
    }
    if (in1 && in3 && ((reinterpret_cast<uintptr_t>(ph) & 7) == 0)) {
        *reinterpret_cast<int2*>(ph) = make_int2(*reinterpret_cast<const int*>(&r.v[1]), *reinterpret_cast<const int*>(&r.v[3]));
    } else {
        if (in1) ph[0] = r.v[1];
        if (in3) ph[1] = r.v[3];
    }
}

template <class L>
__global__ void __launch_bounds__(32 * WARPS_PER_CTA)
k_dwt_forward(const DwtJob* __restrict__ jobs, int seg_pairs) {
    const DwtJob job = jobs[blockIdx.y];
    const int lane = threadIdx.x & 31;
    const int item = blockIdx.x * WARPS_PER_CTA + (threadIdx.x >> 5);
    if (item >= job.nstrips * job.nsegs) return;
    const int strip = item % job.nstrips, seg = item / job.nstrips;
    Region g;
    g.x0 = job.x0; g.y0 = job.y0; g.x1 = job.x1; g.y1 = job.y1; g.pitch = job.pitch;
    g.snx = ((g.x1 + 1) >> 1) - ((g.x0 + 1) >> 1);
    g.sny = ((g.y1 + 1) >> 1) - ((g.y0 + 1) >> 1);
    const int* __restrict__ src = job.src;
    int* __restrict__ dst = job.dst;
    const bool vec_ok = ((g.x0 & 1) == 0) && ((g.pitch & 3) == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
    const int xe = g.x0 & ~1, ye = g.y0 & ~1;
    const int a0 = xe + strip * STRIP_COLS - 4 + lane * 4;       // even absolute column
    const bool owner = lane >= 1 && lane <= 30;
    const bool wide = (g.x1 - g.x0) > 1;
    const int e_first = ye + 2 * seg * seg_pairs;
    int e_last = e_first + 2 * (seg_pairs - 1);
    if (e_last > g.y1 - 1) e_last = (g.y1 - 1) & ~1;

    if (g.y1 - g.y0 == 1) {                    // single row: no vertical filtering (dwt.cpp:318,323)
        if (seg != 0) return;
        Q<int> r = load_quad<int>(src, g, 0, a0, vec_ok);
        int odd = g.y0 & 1;
#pragma unroll
        for (int k = 0; k < 4; ++k) r.v[k] = L::single(r.v[k], odd);
        if (wide) L::horizontal(r);
        else { int oddx = g.x0 & 1; r.v[oddx] = L::single(r.v[oddx], oddx); }
        if (owner) store_split<int>(dst, g, 0, a0, r);
        return;
    }

    L lift;
    const int n0 = e_first + L::LAG - 2 * L::WARM;
    lift.init(load_quad<int>(src, g, mirror(n0 - 2, g.y0, g.y1) - g.y0, a0, vec_ok));
#pragma unroll 2
    for (int n = n0; n - L::LAG <= e_last; n += 2) {
        Q<int> xo = load_quad<int>(src, g, mirror(n - 1, g.y0, g.y1) - g.y0, a0, vec_ok);
        Q<int> xv = load_quad<int>(src, g, mirror(n, g.y0, g.y1) - g.y0, a0, vec_ok);
        Q<int> lo, hi;
        lift.step(xo, xv, lo, hi);
        const int e = n - L::LAG;
        if (e < e_first) continue;                        // warm-up (warp-uniform)
        if (wide) { L::horizontal(lo); L::horizontal(hi); }
        else {
            int oddx = g.x0 & 1;                          // one column: lives at a0 + oddx of the lane owning x0
            lo.v[oddx] = L::single(lo.v[oddx], oddx); hi.v[oddx] = L::single(hi.v[oddx], oddx);
        }
        if (owner) {
            if (e >= g.y0 && e < g.y1) store_split<int>(dst, g, (e >> 1) - ((g.y0 + 1) >> 1), a0, lo);
            if (e + 1 >= g.y0 && e + 1 < g.y1) store_split<int>(dst, g, g.sny + ((e + 1) >> 1) - (g.y0 >> 1), a0, hi);
        }
    }
}

// ---- lifting kernels (inverse) -----------------------------------------------------------------
// Vertical streaming step: new low row a = L(n), new high row b = H(n+1); emits the finished pair
// (row n-LAG, row n-LAG+1).
struct Inv53 {                 // dwt.cpp:335-352
    typedef int T;
    static constexpr int LAG = 2, WARM = 1;
    Q<int> hprev, sprev;       // H(n-1), s(n-2)
    __device__ void init(const Q<int>& h) {
        hprev = h;
#pragma unroll
        for (int k = 0; k < 4; ++k) sprev.v[k] = 0;
    }
    __device__ void step(const Q<int>& a, const Q<int>& b, Q<int>& even, Q<int>& odd) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int s = a.v[k] - ((hprev.v[k] + b.v[k] + 2) >> 2);
            even.v[k] = sprev.v[k];
            odd.v[k] = hprev.v[k] + ((sprev.v[k] + s) >> 1);
            sprev.v[k] = s;
            hprev.v[k] = b.v[k];
        }
    }
    static __device__ void horizontal(Q<int>& r) {
        const unsigned m = 0xffffffffu;
        int pv = __shfl_up_sync(m, r.v[3], 1);
        r.v[0] -= (pv + r.v[1] + 2) >> 2;
        r.v[2] -= (r.v[1] + r.v[3] + 2) >> 2;
        int nx = __shfl_down_sync(m, r.v[0], 1);
        r.v[1] += (r.v[0] + r.v[2]) >> 1;
        r.v[3] += (r.v[2] + nx) >> 1;
    }
    static __device__ int single(int x, int odd) { return odd ? x / 2 : x; }   // dwt.cpp:340,345
};

struct Inv97 {                 // float, dwt.cpp:769-921; constants :94-100
    typedef float T;
    static constexpr int LAG = 4, WARM = 3;
    static constexpr float K = 1.230174105f, C = 1.625732422f;
    static constexpr float ALPHA = 1.586134342f, BETA = 0.052980118f, GAMMA = -0.882911075f, DELTA = -0.443506852f;
    Q<float> b1, e1, o1, e2;   // b(n-1), e1(n-2), o1(n-3), e2(n-4)
    static __device__ __forceinline__ float lift(float x, float l, float r, float c) {
        return __fadd_rn(x, __fmul_rn(__fadd_rn(l, r), c));     // no FMA contraction
    }
    __device__ void init(const Q<float>& h) {
#pragma unroll
        for (int k = 0; k < 4; ++k) { b1.v[k] = __fmul_rn(h.v[k], C); e1.v[k] = o1.v[k] = e2.v[k] = 0.f; }
    }
    __device__ void step(const Q<float>& a, const Q<float>& b, Q<float>& even, Q<float>& odd) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float an = __fmul_rn(a.v[k], K), bn = __fmul_rn(b.v[k], C);
            float ne1 = lift(an, b1.v[k], bn, DELTA);
            float no1 = lift(b1.v[k], e1.v[k], ne1, GAMMA);
            float ne2 = lift(e1.v[k], o1.v[k], no1, BETA);
            float no2 = lift(o1.v[k], e2.v[k], ne2, ALPHA);
            even.v[k] = e2.v[k];
            odd.v[k] = no2;
            b1.v[k] = bn; e1.v[k] = ne1; o1.v[k] = no1; e2.v[k] = ne2;
        }
    }
    static __device__ void horizontal(Q<float>& r) {
        const unsigned m = 0xffffffffu;
        r.v[0] = __fmul_rn(r.v[0], K); r.v[2] = __fmul_rn(r.v[2], K);
        r.v[1] = __fmul_rn(r.v[1], C); r.v[3] = __fmul_rn(r.v[3], C);
        float pv = __shfl_up_sync(m, r.v[3], 1);
        r.v[0] = lift(r.v[0], pv, r.v[1], DELTA);
        r.v[2] = lift(r.v[2], r.v[1], r.v[3], DELTA);
        float nx = __shfl_down_sync(m, r.v[0], 1);
        r.v[1] = lift(r.v[1], r.v[0], r.v[2], GAMMA);
        r.v[3] = lift(r.v[3], r.v[2], nx, GAMMA);
        pv = __shfl_up_sync(m, r.v[3], 1);
        r.v[0] = lift(r.v[0], pv, r.v[1], BETA);
        r.v[2] = lift(r.v[2], r.v[1], r.v[3], BETA);
        nx = __shfl_down_sync(m, r.v[0], 1);
        r.v[1] = lift(r.v[1], r.v[0], r.v[2], ALPHA);
        r.v[3] = lift(r.v[3], r.v[2], nx, ALPHA);
    }
    static __device__ float single(float x, int) { return x; }
};

// four interleaved samples at absolute columns a0..a0+3 gathered from the low and high halves of
// Mallat row `mrow`
template <typename T>
__device__ __forceinline__ Q<T> load_split(const T* __restrict__ base, const Region& g, int mrow, int a0) {
    Q<T> q;
    const T* p = base + (size_t)mrow * g.pitch;
    const int cl = (g.x0 + 1) >> 1, ch = g.x0 >> 1;
    if (a0 >= g.x0 && a0 + 3 < g.x1) {
        const T* pl = p + ((a0 >> 1) - cl);
        const T* ph = p + g.snx + ((a0 >> 1) - ch);
        if (((reinterpret_cast<uintptr_t>(pl) | reinterpret_cast<uintptr_t>(ph)) & 7) == 0) {
            int2 l = __ldg(reinterpret_cast<const int2*>(pl)), h = __ldg(reinterpret_cast<const int2*>(ph));
            q.v[0] = *reinterpret_cast<T*>(&l.x); q.v[2] = *reinterpret_cast<T*>(&l.y);
            q.v[1] = *reinterpret_cast<T*>(&h.x); q.v[3] = *reinterpret_cast<T*>(&h.y);
        } else {
            q.v[0] = __ldg(pl); q.v[2] = __ldg(pl + 1); q.v[1] = __ldg(ph); q.v[3] = __ldg(ph + 1);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int a = mirror(a0 + k, g.x0, g.x1);
            q.v[k] = (a & 1) ? __ldg(p + g.snx + ((a >> 1) - ch)) : __ldg(p + ((a >> 1) - cl));
        }
    }
    return q;
}

template <typename T>
__device__ __forceinline__ void store_quad(T* __restrict__ dst, const Region& g, int row, int a0, const Q<T>& r, bool vec_ok) {
    T* p = dst + (size_t)row * g.pitch + (a0 - g.x0);
    if (vec_ok && a0 >= g.x0 && a0 + 3 < g.x1) {
        *reinterpret_cast<int4*>(p) = make_int4(*reinterpret_cast<const int*>(&r.v[0]), *reinterpret_cast<const int*>(&r.v[1]),
                                                *reinterpret_cast<const int*>(&r.v[2]), *reinterpret_cast<const int*>(&r.v[3]));
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) if (a0 + k >= g.x0 && a0 + k < g.x1) p[k] = r.v[k];
    }
}

// Mallat row index of absolute row n (reflected into the region first)
__device__ __forceinline__ int mallat_row(const Region& g, int n) {
    int r = mirror(n, g.y0, g.y1);
    return (r & 1) ? g.sny + ((r >> 1) - (g.y0 >> 1)) : (r >> 1) - ((g.y0 + 1) >> 1);
}

template <class L>
__global__ void __launch_bounds__(32 * WARPS_PER_CTA)
k_dwt_inverse(const DwtJob* __restrict__ jobs, int seg_pairs) {
    typedef typename L::T T;
    const DwtJob job = jobs[blockIdx.y];
    const int lane = threadIdx.x & 31;
    const int item = blockIdx.x * WARPS_PER_CTA + (threadIdx.x >> 5);
    if (item >= job.nstrips * job.nsegs) return;
    const int strip = item % job.nstrips, seg = item / job.nstrips;
    Region g;
    g.x0 = job.x0; g.y0 = job.y0; g.x1 = job.x1; g.y1 = job.y1; g.pitch = job.pitch;
    g.snx = ((g.x1 + 1) >> 1) - ((g.x0 + 1) >> 1);
    g.sny = ((g.y1 + 1) >> 1) - ((g.y0 + 1) >> 1);
    const T* __restrict__ src = reinterpret_cast<const T*>(job.src);
    T* __restrict__ dst = reinterpret_cast<T*>(job.dst);
    const bool vec_ok = ((g.x0 & 1) == 0) && ((g.pitch & 3) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0);
    const int xe = g.x0 & ~1, ye = g.y0 & ~1;
    const int a0 = xe + strip * STRIP_COLS - 4 + lane * 4;
    const bool owner = lane >= 1 && lane <= 30;
    const bool wide = (g.x1 - g.x0) > 1;
    const int oddx = g.x0 & 1;
    const int e_first = ye + 2 * seg * seg_pairs;
    int e_last = e_first + 2 * (seg_pairs - 1);
    if (e_last > g.y1 - 1) e_last = (g.y1 - 1) & ~1;

    if (g.y1 - g.y0 == 1) {
        if (seg != 0) return;
        Q<T> r = load_split<T>(src, g, 0, a0);
        if (wide) L::horizontal(r); else r.v[oddx] = L::single(r.v[oddx], oddx);
        int odd = g.y0 & 1;
#pragma unroll
        for (int k = 0; k < 4; ++k) r.v[k] = L::single(r.v[k], odd);
        if (owner) store_quad<T>(dst, g, 0, a0, r, vec_ok);
        return;
    }

    L lift;
    const int n0 = e_first + L::LAG - 2 * L::WARM;
    {
        Q<T> h = load_split<T>(src, g, mallat_row(g, n0 - 1), a0);
        if (wide) L::horizontal(h); else h.v[oddx] = L::single(h.v[oddx], oddx);
        lift.init(h);
    }
#pragma unroll 2
    for (int n = n0; n - L::LAG <= e_last; n += 2) {
        Q<T> a = load_split<T>(src, g, mallat_row(g, n), a0);
        Q<T> b = load_split<T>(src, g, mallat_row(g, n + 1), a0);
        if (wide) { L::horizontal(a); L::horizontal(b); }
        else { a.v[oddx] = L::single(a.v[oddx], oddx); b.v[oddx] = L::single(b.v[oddx], oddx); }
        Q<T> ev, od;
        lift.step(a, b, ev, od);
        const int e = n - L::LAG;
        if (e < e_first || !owner) continue;
        if (e >= g.y0 && e < g.y1) store_quad<T>(dst, g, e - g.y0, a0, ev, vec_ok);
        if (e + 1 >= g.y0 && e + 1 < g.y1) store_quad<T>(dst, g, e + 1 - g.y0, a0, od, vec_ok);
    }
}

}  // namespace

// Segment length: long enough that the warm-up rows are a few percent, short enough that one
// level of one 4K frame still gives every SM a dozen or more warps.
static int pick_seg_pairs(const DwtJob* jobs, int njobs) {
    long long best = 64;
    for (int sp = 64; sp >= 8; sp >>= 1) {
        long long items = 0;
        for (int i = 0; i < njobs; ++i) {
            int xe = jobs[i].x0 & ~1, ye = jobs[i].y0 & ~1;
            long long strips = (jobs[i].x1 - xe + STRIP_COLS - 1) / STRIP_COLS;
            long long pairs = (((jobs[i].y1 + 1) & ~1) - ye) / 2;
            items += strips * ((pairs + sp - 1) / sp);
        }
        best = sp;
        if (items >= 148LL * 24) break;
    }
    return (int)best;
}

int dwt_plan_jobs(DwtJob* jobs, int njobs) {
    int sp = pick_seg_pairs(jobs, njobs);
    int max_items = 0;
    for (int i = 0; i < njobs; ++i) {
        int xe = jobs[i].x0 & ~1, ye = jobs[i].y0 & ~1;
        jobs[i].nstrips = (jobs[i].x1 - xe + STRIP_COLS - 1) / STRIP_COLS;
        int pairs = (((jobs[i].y1 + 1) & ~1) - ye) / 2;
        jobs[i].nsegs = (pairs + sp - 1) / sp;
        jobs[i].strip0 = sp;
        int items = jobs[i].nstrips * jobs[i].nsegs;
        if (items > max_items) max_items = items;
    }
    return max_items;
}

void launch_dwt_forward(const DwtJob* d_jobs, int njobs, int max_items, int seg_pairs, int kind, cudaStream_t st) {
    if (njobs <= 0 || max_items <= 0) return;
    dim3 grid((max_items + WARPS_PER_CTA - 1) / WARPS_PER_CTA, njobs), block(32 * WARPS_PER_CTA);
    if (kind == DWT_53) k_dwt_forward<Fwd53><<<grid, block, 0, st>>>(d_jobs, seg_pairs);
    else k_dwt_forward<Fwd97><<<grid, block, 0, st>>>(d_jobs, seg_pairs);
}

void launch_dwt_inverse(const DwtJob* d_jobs, int njobs, int max_items, int seg_pairs, int kind, cudaStream_t st) {
    if (njobs <= 0 || max_items <= 0) return;
    dim3 grid((max_items + WARPS_PER_CTA - 1) / WARPS_PER_CTA, njobs), block(32 * WARPS_PER_CTA);
    if (kind == DWT_53) k_dwt_inverse<Inv53><<<grid, block, 0, st>>>(d_jobs, seg_pairs);
    else k_dwt_inverse<Inv97><<<grid, block, 0, st>>>(d_jobs, seg_pairs);
}

}  // namespace j2k
