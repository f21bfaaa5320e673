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
#include <cstdlib>

namespace j2k {
namespace {

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

// N adjacent samples of one row held by one lane (N = 4 or 8; v[0] sits on an even column)
template <typename T, int N> struct Q { T v[N]; };

// ---- lifting kernels (forward) -----------------------------------------------------------------
// Each provides: vertical streaming step (new odd row xo = x(n-1), new even row xe = x(n); emits
// the finished pair lo = row n-LAG, hi = row n-LAG+1) and the horizontal lifting across lanes.
template <int N> struct Fwd53 {
    typedef int T;
    static constexpr int LAG = 2, WARM = 1;
    Q<int, N> xprev, dprev;      // x(n-2), d(n-3)
    __device__ void init(const Q<int, N>& x) {
        xprev = x;
#pragma unroll
        for (int k = 0; k < N; ++k) dprev.v[k] = 0;
    }
    __device__ void step(const Q<int, N>& xo, const Q<int, N>& xe, Q<int, N>& lo, Q<int, N>& hi) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
            int d = xo.v[k] - ((xprev.v[k] + xe.v[k]) >> 1);
            lo.v[k] = xprev.v[k] + ((dprev.v[k] + d + 2) >> 2);
            hi.v[k] = d;
            dprev.v[k] = d;
            xprev.v[k] = xe.v[k];
        }
    }
    static __device__ void horizontal(Q<int, N>& r) {
        const int nx = __shfl_down_sync(0xffffffffu, r.v[0], 1);
#pragma unroll
        for (int k = 1; k < N; k += 2) r.v[k] -= (r.v[k - 1] + (k + 1 < N ? r.v[k + 1 < N ? k + 1 : 0] : nx)) >> 1;
        const int pv = __shfl_up_sync(0xffffffffu, r.v[N - 1], 1);
#pragma unroll
        for (int k = 0; k < N; k += 2) r.v[k] += ((k ? r.v[k ? k - 1 : 0] : pv) + r.v[k + 1] + 2) >> 2;
    }
    static __device__ int single(int x, int odd) { return odd ? x * 2 : x; }   // dwt.cpp:318,323
};

template <int N> struct Fwd97 {                 // Q13 fixed point, dwt.cpp:365-399
    typedef int T;
    static constexpr int LAG = 4, WARM = 3;
    Q<int, N> x2, d1, s1, d2;     // x(n-2), d1(n-3), s1(n-4), d2(n-5)
    __device__ void init(const Q<int, N>& x) {
        x2 = x;
#pragma unroll
        for (int k = 0; k < N; ++k) d1.v[k] = s1.v[k] = d2.v[k] = 0;
    }
    __device__ void step(const Q<int, N>& xo, const Q<int, N>& xe, Q<int, N>& lo, Q<int, N>& hi) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
            int nd1 = xo.v[k] - fm(x2.v[k] + xe.v[k], 12993);
            int ns1 = x2.v[k] - fm(d1.v[k] + nd1, 434);
            int nd2 = d1.v[k] + fm(s1.v[k] + ns1, 7233);
            int ns2 = s1.v[k] + fm(d2.v[k] + nd2, 3633);
            lo.v[k] = fm(ns2, 6659);
            hi.v[k] = fm(nd2, 5038);
            x2.v[k] = xe.v[k]; d1.v[k] = nd1; s1.v[k] = ns1; d2.v[k] = nd2;
        }
    }
    // one lifting step on the odd (ODD = 1) or even samples: r[k] += sign * fm(r[k-1] + r[k+1], c)
    template <int ODD> static __device__ __forceinline__ void lift(Q<int, N>& r, int c, int sign) {
        const unsigned m = 0xffffffffu;
        if (ODD) {
            const int nx = __shfl_down_sync(m, r.v[0], 1);
#pragma unroll
            for (int k = 1; k < N; k += 2) r.v[k] += sign * fm(r.v[k - 1] + (k + 1 < N ? r.v[k + 1 < N ? k + 1 : 0] : nx), c);
        } else {
            const int pv = __shfl_up_sync(m, r.v[N - 1], 1);
#pragma unroll
            for (int k = 0; k < N; k += 2) r.v[k] += sign * fm((k ? r.v[k ? k - 1 : 0] : pv) + r.v[k + 1], c);
        }
    }
    static __device__ void horizontal(Q<int, N>& r) {
        lift<1>(r, 12993, -1);
        lift<0>(r, 434, -1);
        lift<1>(r, 7233, 1);
        lift<0>(r, 3633, 1);
#pragma unroll
        for (int k = 0; k < N; k += 2) { r.v[k] = fm(r.v[k], 6659); r.v[k + 1] = fm(r.v[k + 1], 5038); }
    }
    static __device__ int single(int x, int) { return x; }                     // dwt.cpp:368
};

// ---- loads / stores ----------------------------------------------------------------------------
struct Region {
    int x0, y0, x1, y1, pitch;
    int snx, sny;             // number of low-pass columns / rows
};

// Which 4-column groups of its N columns a lane has to load: owner lanes all of those that come
// within 4 columns of the region, the halo lane on the left only its last group, the one on the
// right only its first.
template <int N>
__device__ __forceinline__ unsigned group_mask(const Region& g, int a0, int lane, int owners) {
    unsigned m = 0;
    if (lane > owners + 1) return 0;
#pragma unroll
    for (int q = 0; q < N / 4; ++q) {
        const int c = a0 + 4 * q;
        if (c + 4 > g.x0 - 4 && c < g.x1 + 4) m |= 1u << q;
    }
    if (lane == 0) m &= 1u << (N / 4 - 1);
    if (lane == owners + 1) m &= 1u;
    return m;
}

// N samples at absolute columns a0..a0+N-1 of memory row `row` (already reflected) of an
// interleaved region
template <typename T, int N>
__device__ __forceinline__ Q<T, N> load_row(const T* __restrict__ base, const Region& g, int row, int a0, bool vec_ok, unsigned gm) {
    Q<T, N> q;
    const T* p = base + (size_t)row * g.pitch;
#pragma unroll
    for (int s = 0; s < N / 4; ++s) {
        const int c = a0 + 4 * s;
        if (!((gm >> s) & 1u)) {
#pragma unroll
            for (int k = 0; k < 4; ++k) q.v[4 * s + k] = T(0);
        } else if (vec_ok && c >= g.x0 && c + 3 < g.x1) {
            int4 t = __ldg(reinterpret_cast<const int4*>(p + (c - g.x0)));
            q.v[4 * s + 0] = *reinterpret_cast<T*>(&t.x); q.v[4 * s + 1] = *reinterpret_cast<T*>(&t.y);
            q.v[4 * s + 2] = *reinterpret_cast<T*>(&t.z); q.v[4 * s + 3] = *reinterpret_cast<T*>(&t.w);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) q.v[4 * s + k] = __ldg(p + (mirror(c + k, g.x0, g.x1) - g.x0));
        }
    }
    return q;
}

// store the low samples (even columns) and high samples (odd columns) of a finished row into the
// Mallat row `mrow` of dst
template <typename T, int N>
__device__ __forceinline__ void store_split(T* __restrict__ dst, const Region& g, int mrow, int a0, const Q<T, N>& r) {
    T* p = dst + (size_t)mrow * g.pitch;
    T* pl = p + ((a0 >> 1) - ((g.x0 + 1) >> 1));
    T* ph = p + g.snx + ((a0 >> 1) - (g.x0 >> 1));
    const bool inside = a0 >= g.x0 && a0 + N - 1 < g.x1;
    if (N == 8 && inside && (((reinterpret_cast<uintptr_t>(pl) | reinterpret_cast<uintptr_t>(ph)) & 15) == 0)) {
        *reinterpret_cast<int4*>(pl) = make_int4(*reinterpret_cast<const int*>(&r.v[0]), *reinterpret_cast<const int*>(&r.v[2]),
                                                 *reinterpret_cast<const int*>(&r.v[N > 4 ? 4 : 0]), *reinterpret_cast<const int*>(&r.v[N > 4 ? 6 : 0]));
        *reinterpret_cast<int4*>(ph) = make_int4(*reinterpret_cast<const int*>(&r.v[1]), *reinterpret_cast<const int*>(&r.v[3]),
                                                 *reinterpret_cast<const int*>(&r.v[N > 4 ? 5 : 0]), *reinterpret_cast<const int*>(&r.v[N > 4 ? 7 : 0]));
        return;
    }
#pragma unroll
    for (int s = 0; s < N / 4; ++s) {
        const int c = a0 + 4 * s;
        T* ql = pl + 2 * s;
        T* qh = ph + 2 * s;
        const bool in0 = c >= g.x0 && c < g.x1, in1 = c + 1 >= g.x0 && c + 1 < g.x1;
        const bool in2 = c + 2 >= g.x0 && c + 2 < g.x1, in3 = c + 3 >= g.x0 && c + 3 < g.x1;
        if (in0 && in2 && ((reinterpret_cast<uintptr_t>(ql) & 7) == 0)) {
            *reinterpret_cast<int2*>(ql) = make_int2(*reinterpret_cast<const int*>(&r.v[4 * s]), *reinterpret_cast<const int*>(&r.v[4 * s + 2]));
        } else {
            if (in0) ql[0] = r.v[4 * s];
            if (in2) ql[1] = r.v[4 * s + 2];
        }
        if (in1 && in3 && ((reinterpret_cast<uintptr_t>(qh) & 7) == 0)) {
            *reinterpret_cast<int2*>(qh) = make_int2(*reinterpret_cast<const int*>(&r.v[4 * s + 1]), *reinterpret_cast<const int*>(&r.v[4 * s + 3]));
        } else {
            if (in1) qh[0] = r.v[4 * s + 1];
            if (in3) qh[1] = r.v[4 * s + 3];
        }
    }
}

// Column-interior strips (every column a lane needs lies inside the region and all accesses are
// naturally aligned) take a branch-free path: predicated 16-byte loads, 8/16-byte stores.
template <typename T, int N>
__device__ __forceinline__ Q<T, N> load_row_fast(const T* __restrict__ p, unsigned gm) {
    Q<T, N> q;
#pragma unroll
    for (int s = 0; s < N / 4; ++s) {
        int4 t = make_int4(0, 0, 0, 0);
        if ((gm >> s) & 1u) t = __ldg(reinterpret_cast<const int4*>(p + 4 * s));
        q.v[4 * s + 0] = *reinterpret_cast<T*>(&t.x); q.v[4 * s + 1] = *reinterpret_cast<T*>(&t.y);
        q.v[4 * s + 2] = *reinterpret_cast<T*>(&t.z); q.v[4 * s + 3] = *reinterpret_cast<T*>(&t.w);
    }
    return q;
}
template <typename T, int N>
__device__ __forceinline__ void store_split_fast(T* __restrict__ pl, int snx, const Q<T, N>& r) {
    T* ph = pl + snx;
    if (N == 8) {
        *reinterpret_cast<int4*>(pl) = make_int4(*reinterpret_cast<const int*>(&r.v[0]), *reinterpret_cast<const int*>(&r.v[2]),
                                                 *reinterpret_cast<const int*>(&r.v[N > 4 ? 4 : 0]), *reinterpret_cast<const int*>(&r.v[N > 4 ? 6 : 0]));
        *reinterpret_cast<int4*>(ph) = make_int4(*reinterpret_cast<const int*>(&r.v[1]), *reinterpret_cast<const int*>(&r.v[3]),
                                                 *reinterpret_cast<const int*>(&r.v[N > 4 ? 5 : 0]), *reinterpret_cast<const int*>(&r.v[N > 4 ? 7 : 0]));
    } else {
        *reinterpret_cast<int2*>(pl) = make_int2(*reinterpret_cast<const int*>(&r.v[0]), *reinterpret_cast<const int*>(&r.v[2]));
        *reinterpret_cast<int2*>(ph) = make_int2(*reinterpret_cast<const int*>(&r.v[1]), *reinterpret_cast<const int*>(&r.v[3]));
    }
}

struct StripCtx {
    Region g;
    int a0, owners, e_first, e_last;
    unsigned gm;
    bool owner, vec_ok, wide;
};

template <class L, int N, int PF, bool FAST>
__device__ __forceinline__ void forward_rows(const int* __restrict__ src, int* __restrict__ dst, const StripCtx& c) {
    const Region& g = c.g;
    const int a0 = c.a0;
    const int* __restrict__ scol = src + (a0 - g.x0);                       // FAST: this lane's column in row 0
    int* __restrict__ dcol = dst + ((a0 >> 1) - ((g.x0 + 1) >> 1));         // FAST: low-band position (x0 is even there)
    auto ld = [&](int n) -> Q<int, N> {
        const int row = mirror(n, g.y0, g.y1) - g.y0;
        if (FAST) return load_row_fast<int, N>(scol + (size_t)row * g.pitch, c.gm);
        return load_row<int, N>(src, g, row, a0, c.vec_ok, c.gm);
    };
    L lift;
    const int n0 = c.e_first + L::LAG - 2 * L::WARM;
    lift.init(ld(n0 - 2));
    const int n_end = c.e_last + L::LAG;         // last n of the loop
    constexpr int D = PF > 0 ? PF : 1;           // row pairs fetched ahead (ring of D pairs)
    Q<int, N> po[D], pe[D];
    if (PF) {
#pragma unroll
        for (int j = 0; j < D; ++j) if (n0 + 2 * j <= n_end) { po[j] = ld(n0 + 2 * j - 1); pe[j] = ld(n0 + 2 * j); }
    }
#pragma unroll 1
    for (int nb = n0; nb <= n_end; nb += 2 * D * (PF ? 1 : 2)) {
#pragma unroll
        for (int j = 0; j < (PF ? D : 2); ++j) {
            const int n = nb + 2 * j;
            if (n > n_end) break;
            Q<int, N> xo, xv;
            if (PF) {
                xo = po[j % D]; xv = pe[j % D];
                if (n + 2 * D <= n_end) { po[j % D] = ld(n + 2 * D - 1); pe[j % D] = ld(n + 2 * D); }
            } else {
                xo = ld(n - 1); xv = ld(n);
            }
            Q<int, N> lo, hi;
            lift.step(xo, xv, lo, hi);
            const int e = n - L::LAG;
            if (e < c.e_first) continue;                      // warm-up (warp-uniform)
            if (FAST || c.wide) { L::horizontal(lo); L::horizontal(hi); }
            else {
                int oddx = g.x0 & 1;                          // one column: lives at a0 + oddx of the lane owning x0
                lo.v[oddx] = L::single(lo.v[oddx], oddx); hi.v[oddx] = L::single(hi.v[oddx], oddx);
            }
            if (c.owner) {
                const int mlo = (e >> 1) - ((g.y0 + 1) >> 1), mhi = g.sny + ((e + 1) >> 1) - (g.y0 >> 1);
                if (FAST) {
                    if (e >= g.y0 && e < g.y1) store_split_fast<int, N>(dcol + (size_t)mlo * g.pitch, g.snx, lo);
                    if (e + 1 >= g.y0 && e + 1 < g.y1) store_split_fast<int, N>(dcol + (size_t)mhi * g.pitch, g.snx, hi);
                } else {
                    if (e >= g.y0 && e < g.y1) store_split<int, N>(dst, g, mlo, a0, lo);
                    if (e + 1 >= g.y0 && e + 1 < g.y1) store_split<int, N>(dst, g, mhi, a0, hi);
                }
            }
        }
    }
}

template <class L, int N, int PF>
__global__ void __launch_bounds__(32 * WARPS_PER_CTA)
k_dwt_forward(const DwtJob* __restrict__ jobs) {
    const DwtJob job = jobs[blockIdx.y];
    const int lane = threadIdx.x & 31;
    const int item = blockIdx.x * WARPS_PER_CTA + (threadIdx.x >> 5);
    if (item >= job.nstrips * job.nsegs) return;
    const int strip = item % job.nstrips, seg = item / job.nstrips;
    const int seg_pairs = job.strip0;
    StripCtx c;
    Region& g = c.g;
    g.x0 = job.x0; g.y0 = job.y0; g.x1 = job.x1; g.y1 = job.y1; g.pitch = job.pitch;
    g.snx = ((g.x1 + 1) >> 1) - ((g.x0 + 1) >> 1);
    g.sny = ((g.y1 + 1) >> 1) - ((g.y0 + 1) >> 1);
    const int* __restrict__ src = job.src;
    int* __restrict__ dst = job.dst;
    c.vec_ok = ((g.x0 & 1) == 0) && ((g.pitch & 3) == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
    const int xe = g.x0 & ~1, ye = g.y0 & ~1;
    c.owners = job.strip_cols / N;
    const int s0 = xe + strip * job.strip_cols;                   // first owned column of the strip
    c.a0 = s0 - N + lane * N;                                     // even absolute column
    c.owner = lane >= 1 && lane <= c.owners;
    c.gm = group_mask<N>(g, c.a0, lane, c.owners);
    c.wide = (g.x1 - g.x0) > 1;
    c.e_first = ye + 2 * seg * seg_pairs;
    c.e_last = c.e_first + 2 * (seg_pairs - 1);
    if (c.e_last > g.y1 - 1) c.e_last = (g.y1 - 1) & ~1;

    if (g.y1 - g.y0 == 1) {                    // single row: no vertical filtering (dwt.cpp:318,323)
        if (seg != 0) return;
        Q<int, N> r = load_row<int, N>(src, g, 0, c.a0, c.vec_ok, c.gm);
        int odd = g.y0 & 1;
#pragma unroll
        for (int k = 0; k < N; ++k) r.v[k] = L::single(r.v[k], odd);
        if (c.wide) L::horizontal(r);
        else { int oddx = g.x0 & 1; r.v[oddx] = L::single(r.v[oddx], oddx); }
        if (c.owner) store_split<int, N>(dst, g, 0, c.a0, r);
        return;
    }
    // warp-uniform: the strip with its 4-column halos lies inside the region, everything aligned
    const int align = N == 8 ? 7 : 3;
    const bool fast = c.vec_ok && s0 - 4 >= g.x0 && s0 + c.owners * N + 4 <= g.x1 && ((s0 - g.x0) & align) == 0 &&
                      ((g.x0 >> 1) & (align >> 1)) == 0 && (g.snx & (align >> 1)) == 0 &&
                      ((reinterpret_cast<uintptr_t>(dst) & 15) == 0);
    if (fast) forward_rows<L, N, PF, true>(src, dst, c);
    else forward_rows<L, N, PF, false>(src, dst, c);
}

// ---- level 1 fused with sample unpacking, DC level shift and RCT / ICT ---------------------------------------------------
// The first decomposition level reads the caller's interleaved pixels directly (2 or 1 bytes per sample instead of the 4-byte
// planes the separate MCT kernel would have written and this kernel read back): per row a lane takes its four pixels with
// 16/8/4-byte loads, applies the level shift and the component transform in registers (tcd.cpp:1659-1696, mct.cpp:94-157,
// 241-401) and feeds NC vertical lifting recurrences at once.  Everything after the load is the code of k_dwt_forward.

__device__ __forceinline__ int fm13(int a, int b) { return (int)(((long long)a * (long long)b + 4096) >> 13); }

template <int S> __device__ __forceinline__ int px_sample(const uint32_t* w, int idx) {
    if (S == 1) return (int)((w[idx >> 2] >> ((idx & 3) * 8)) & 0xffu);
    return (int)((w[idx >> 1] >> ((idx & 1) * 16)) & 0xffffu);
}

// four pixels at columns a0..a0+3 of image row `row` -> NC component quads (shifted, transformed)
template <int NC, int NCH, int S, bool FAST>
__device__ __forceinline__ void load_px(const DwtPxJob& j, const DwtPxSrc& ps, int row, int a0, unsigned gm, Q<int, 4>* out) {
    constexpr int NW = NCH * S;                         // 32-bit words of four pixels
    uint32_t w[NW];
#pragma unroll
    for (int k = 0; k < NW; ++k) w[k] = 0;
    const uint8_t* rp = ps.px + (long long)row * ps.rowbytes;
    if (gm) {
        if (FAST || (a0 >= j.x0 && a0 + 3 < j.x1 && ((reinterpret_cast<uintptr_t>(rp + (long long)a0 * NCH * S)) & (NW % 4 == 0 ? 15 : (NW % 2 == 0 ? 7 : 3))) == 0)) {
            const uint8_t* p = rp + (long long)a0 * NCH * S;
            if (NW % 4 == 0) {
#pragma unroll
                for (int k = 0; k < NW / 4; ++k) { const uint4 t = __ldg(reinterpret_cast<const uint4*>(p) + k); w[4 * k] = t.x; w[4 * k + 1] = t.y; w[4 * k + 2] = t.z; w[4 * k + 3] = t.w; }
            } else if (NW % 2 == 0) {
#pragma unroll
                for (int k = 0; k < NW / 2; ++k) { const uint2 t = __ldg(reinterpret_cast<const uint2*>(p) + k); w[2 * k] = t.x; w[2 * k + 1] = t.y; }
            } else {
#pragma unroll
                for (int k = 0; k < NW; ++k) w[k] = __ldg(reinterpret_cast<const uint32_t*>(p) + k);
            }
        } else {
            // region edge: reflected columns, sample by sample
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int col = mirror(a0 + k, j.x0, j.x1);
                const uint8_t* p = rp + (long long)col * NCH * S;
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    const uint32_t v = S == 1 ? (uint32_t)__ldg(p + c) : (uint32_t)__ldg(reinterpret_cast<const uint16_t*>(p) + c);
                    const int idx = k * NCH + c;
                    if (S == 1) w[idx >> 2] |= v << ((idx & 3) * 8); else w[idx >> 1] |= v << ((idx & 1) * 16);
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int v[4];
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            int s = px_sample<S>(w, k * NCH + 0);
            if (NCH > 1 && ps.chan[c] == 1) s = px_sample<S>(w, k * NCH + 1);
            if (NCH > 2 && ps.chan[c] == 2) s = px_sample<S>(w, k * NCH + 2);
            if (NCH > 3 && ps.chan[c] == 3) s = px_sample<S>(w, k * NCH + 3);
            s -= j.dc[c];
            v[c] = j.irreversible ? s << 11 : s;
        }
        if (NC >= 3 && j.mct) {
            const int r = v[0], g = v[1], b = v[2];
            if (!j.irreversible) { v[0] = (r + 2 * g + b) >> 2; v[1] = b - g; v[2] = r - g; }
            else {
                v[0] = fm13(r, 2449) + fm13(g, 4809) + fm13(b, 934);
                v[1] = -fm13(r, 1382) - fm13(g, 2714) + fm13(b, 4096);
                v[2] = fm13(r, 4096) - fm13(g, 3430) - fm13(b, 666);
            }
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) out[c].v[k] = gm ? v[c] : 0;
    }
}

template <class L, int NC, int NCH, int S, bool FAST>
__device__ __forceinline__ void forward_rows_px(const DwtPxJob& j, const DwtPxSrc& ps, const StripCtx& c) {
    constexpr int N = 4;
    const Region& g = c.g;
    const int a0 = c.a0;
    auto ld = [&](int n, Q<int, N>* q) { load_px<NC, NCH, S, FAST>(j, ps, mirror(n, g.y0, g.y1), a0, c.gm, q); };
    L lift[NC];
    const int n0 = c.e_first + L::LAG - 2 * L::WARM;
    {
        Q<int, N> q[NC];
        ld(n0 - 2, q);
#pragma unroll
        for (int k = 0; k < NC; ++k) lift[k].init(q[k]);
    }
    const int n_end = c.e_last + L::LAG;
    Q<int, N> po[NC], pe[NC];                       // the next row pair, fetched one iteration ahead
    ld(n0 - 1, po); ld(n0, pe);
#pragma unroll 1
    for (int n = n0; n <= n_end; n += 2) {
        Q<int, N> xo[NC], xv[NC];
#pragma unroll
        for (int k = 0; k < NC; ++k) { xo[k] = po[k]; xv[k] = pe[k]; }
        if (n + 2 <= n_end) { ld(n + 1, po); ld(n + 2, pe); }
        const int e = n - L::LAG;
        const int mlo = (e >> 1) - ((g.y0 + 1) >> 1), mhi = g.sny + ((e + 1) >> 1) - (g.y0 >> 1);
#pragma unroll
        for (int k = 0; k < NC; ++k) {
            Q<int, N> lo, hi;
            lift[k].step(xo[k], xv[k], lo, hi);
            if (e < c.e_first) continue;                      // warm-up (warp-uniform)
            if (FAST || c.wide) { L::horizontal(lo); L::horizontal(hi); }
            else {
                const int oddx = g.x0 & 1;
                lo.v[oddx] = L::single(lo.v[oddx], oddx); hi.v[oddx] = L::single(hi.v[oddx], oddx);
            }
            if (!c.owner) continue;
            int* dst = j.dst[k];
            if (FAST) {
                int* dcol = dst + ((a0 >> 1) - ((g.x0 + 1) >> 1));
                if (e >= g.y0 && e < g.y1) store_split_fast<int, N>(dcol + (size_t)mlo * g.pitch, g.snx, lo);
                if (e + 1 >= g.y0 && e + 1 < g.y1) store_split_fast<int, N>(dcol + (size_t)mhi * g.pitch, g.snx, hi);
            } else {
                if (e >= g.y0 && e < g.y1) store_split<int, N>(dst, g, mlo, a0, lo);
                if (e + 1 >= g.y0 && e + 1 < g.y1) store_split<int, N>(dst, g, mhi, a0, hi);
            }
        }
    }
}

template <class L, int NC, int NCH, int S>
__global__ void __launch_bounds__(32 * WARPS_PER_CTA)
k_dwt_forward_px(const DwtPxJob* __restrict__ jobs, const DwtPxSrc ps) {
    constexpr int N = 4;
    const DwtPxJob job = jobs[blockIdx.y];
    const int lane = threadIdx.x & 31;
    const int item = blockIdx.x * WARPS_PER_CTA + (threadIdx.x >> 5);
    if (item >= job.nstrips * job.nsegs) return;
    const int strip = item % job.nstrips, seg = item / job.nstrips;
    StripCtx c;
    Region& g = c.g;
    g.x0 = job.x0; g.y0 = job.y0; g.x1 = job.x1; g.y1 = job.y1; g.pitch = job.pitch;
    g.snx = ((g.x1 + 1) >> 1) - ((g.x0 + 1) >> 1);
    g.sny = ((g.y1 + 1) >> 1) - ((g.y0 + 1) >> 1);
    const int xe = g.x0 & ~1, ye = g.y0 & ~1;
    c.owners = job.strip_cols / N;
    const int s0 = xe + strip * job.strip_cols;
    c.a0 = s0 - N + lane * N;
    c.owner = lane >= 1 && lane <= c.owners;
    c.gm = group_mask<N>(g, c.a0, lane, c.owners);
    c.wide = (g.x1 - g.x0) > 1;
    c.vec_ok = false;
    c.e_first = ye + 2 * seg * job.strip0;
    c.e_last = c.e_first + 2 * (job.strip0 - 1);
    if (c.e_last > g.y1 - 1) c.e_last = (g.y1 - 1) & ~1;
    // interior strip, everything aligned: branch-free loads and stores
    constexpr int NW = NCH * S;
    constexpr int LA = NW % 4 == 0 ? 15 : (NW % 2 == 0 ? 7 : 3);
    bool fast = ((g.x0 & 1) == 0) && ((g.pitch & 3) == 0) && s0 - 4 >= g.x0 && s0 + c.owners * N + 4 <= g.x1 && ((s0 - g.x0) & 3) == 0 &&
                ((g.x0 >> 1) & 1) == 0 && (g.snx & 1) == 0 && ((reinterpret_cast<uintptr_t>(ps.px) | (uintptr_t)ps.rowbytes) & LA) == 0 &&
                (((long long)s0 * NCH * S) & LA) == 0;
#pragma unroll
    for (int k = 0; k < NC; ++k) fast = fast && ((reinterpret_cast<uintptr_t>(job.dst[k]) & 15) == 0);
    if (fast) forward_rows_px<L, NC, NCH, S, true>(job, ps, c);
    else forward_rows_px<L, NC, NCH, S, false>(job, ps, c);
}

// ---- lifting kernels (inverse) -----------------------------------------------------------------
// Vertical streaming step: new low row a = L(n), new high row b = H(n+1); emits the finished pair
// (row n-LAG, row n-LAG+1).
template <int N> struct Inv53 {                 // dwt.cpp:335-352
    typedef int T;
    static constexpr int LAG = 2, WARM = 1;
    Q<int, N> hprev, sprev;       // H(n-1), s(n-2)
    __device__ void init(const Q<int, N>& h) {
        hprev = h;
#pragma unroll
        for (int k = 0; k < N; ++k) sprev.v[k] = 0;
    }
    __device__ void step(const Q<int, N>& a, const Q<int, N>& b, Q<int, N>& even, Q<int, N>& odd) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
            int s = a.v[k] - ((hprev.v[k] + b.v[k] + 2) >> 2);
            even.v[k] = sprev.v[k];
            odd.v[k] = hprev.v[k] + ((sprev.v[k] + s) >> 1);
            sprev.v[k] = s;
            hprev.v[k] = b.v[k];
        }
    }
    static __device__ void horizontal(Q<int, N>& r) {
        const unsigned m = 0xffffffffu;
        const int pv = __shfl_up_sync(m, r.v[N - 1], 1);
#pragma unroll
        for (int k = 0; k < N; k += 2) r.v[k] -= ((k ? r.v[k ? k - 1 : 0] : pv) + r.v[k + 1] + 2) >> 2;
        const int nx = __shfl_down_sync(m, r.v[0], 1);
#pragma unroll
        for (int k = 1; k < N; k += 2) r.v[k] += (r.v[k - 1] + (k + 1 < N ? r.v[k + 1 < N ? k + 1 : 0] : nx)) >> 1;
    }
    static __device__ int single(int x, int odd) { return odd ? x / 2 : x; }   // dwt.cpp:340,345
};

template <int N> struct Inv97 {                 // float, dwt.cpp:769-921; constants :94-100
    typedef float T;
    static constexpr int LAG = 4, WARM = 3;
    static constexpr float K = 1.230174105f, C = 1.625732422f;
    static constexpr float ALPHA = 1.586134342f, BETA = 0.052980118f, GAMMA = -0.882911075f, DELTA = -0.443506852f;
    Q<float, N> b1, e1, o1, e2;   // b(n-1), e1(n-2), o1(n-3), e2(n-4)
    static __device__ __forceinline__ float lift(float x, float l, float r, float c) {
        return __fadd_rn(x, __fmul_rn(__fadd_rn(l, r), c));     // no FMA contraction
    }
    __device__ void init(const Q<float, N>& h) {
#pragma unroll
        for (int k = 0; k < N; ++k) { b1.v[k] = __fmul_rn(h.v[k], C); e1.v[k] = o1.v[k] = e2.v[k] = 0.f; }
    }
    __device__ void step(const Q<float, N>& a, const Q<float, N>& b, Q<float, N>& even, Q<float, N>& odd) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
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
    template <int ODD> static __device__ __forceinline__ void hlift(Q<float, N>& r, float c) {
        const unsigned m = 0xffffffffu;
        if (ODD) {
            const float nx = __shfl_down_sync(m, r.v[0], 1);
#pragma unroll
            for (int k = 1; k < N; k += 2) r.v[k] = lift(r.v[k], r.v[k - 1], (k + 1 < N ? r.v[k + 1 < N ? k + 1 : 0] : nx), c);
        } else {
            const float pv = __shfl_up_sync(m, r.v[N - 1], 1);
#pragma unroll
            for (int k = 0; k < N; k += 2) r.v[k] = lift(r.v[k], (k ? r.v[k ? k - 1 : 0] : pv), r.v[k + 1], c);
        }
    }
    static __device__ void horizontal(Q<float, N>& r) {
#pragma unroll
        for (int k = 0; k < N; k += 2) { r.v[k] = __fmul_rn(r.v[k], K); r.v[k + 1] = __fmul_rn(r.v[k + 1], C); }
        hlift<0>(r, DELTA);
        hlift<1>(r, GAMMA);
        hlift<0>(r, BETA);
        hlift<1>(r, ALPHA);
    }
    static __device__ float single(float x, int) { return x; }
};

// N interleaved samples at absolute columns a0..a0+N-1 gathered from the low and high halves of
// Mallat row `mrow`
template <typename T, int N>
__device__ __forceinline__ Q<T, N> load_split(const T* __restrict__ base, const Region& g, int mrow, int a0, unsigned gm) {
    Q<T, N> q;
    const T* p = base + (size_t)mrow * g.pitch;
    const int cl = (g.x0 + 1) >> 1, ch = g.x0 >> 1;
    if (N == 8 && gm == 3u && a0 >= g.x0 && a0 + N - 1 < g.x1) {
        const T* pl = p + ((a0 >> 1) - cl);
        const T* ph = p + g.snx + ((a0 >> 1) - ch);
        if (((reinterpret_cast<uintptr_t>(pl) | reinterpret_cast<uintptr_t>(ph)) & 15) == 0) {
            int4 l = __ldg(reinterpret_cast<const int4*>(pl)), h = __ldg(reinterpret_cast<const int4*>(ph));
            q.v[0] = *reinterpret_cast<T*>(&l.x); q.v[2] = *reinterpret_cast<T*>(&l.y);
            q.v[N > 4 ? 4 : 0] = *reinterpret_cast<T*>(&l.z); q.v[N > 4 ? 6 : 0] = *reinterpret_cast<T*>(&l.w);
            q.v[1] = *reinterpret_cast<T*>(&h.x); q.v[3] = *reinterpret_cast<T*>(&h.y);
            q.v[N > 4 ? 5 : 0] = *reinterpret_cast<T*>(&h.z); q.v[N > 4 ? 7 : 0] = *reinterpret_cast<T*>(&h.w);
            return q;
        }
    }
#pragma unroll
    for (int s = 0; s < N / 4; ++s) {
        const int c = a0 + 4 * s;
        if (!((gm >> s) & 1u)) {
#pragma unroll
            for (int k = 0; k < 4; ++k) q.v[4 * s + k] = T(0);
        } else if (c >= g.x0 && c + 3 < g.x1) {
            const T* pl = p + ((c >> 1) - cl);
            const T* ph = p + g.snx + ((c >> 1) - ch);
            if (((reinterpret_cast<uintptr_t>(pl) | reinterpret_cast<uintptr_t>(ph)) & 7) == 0) {
                int2 l = __ldg(reinterpret_cast<const int2*>(pl)), h = __ldg(reinterpret_cast<const int2*>(ph));
                q.v[4 * s + 0] = *reinterpret_cast<T*>(&l.x); q.v[4 * s + 2] = *reinterpret_cast<T*>(&l.y);
                q.v[4 * s + 1] = *reinterpret_cast<T*>(&h.x); q.v[4 * s + 3] = *reinterpret_cast<T*>(&h.y);
            } else {
                q.v[4 * s + 0] = __ldg(pl); q.v[4 * s + 2] = __ldg(pl + 1); q.v[4 * s + 1] = __ldg(ph); q.v[4 * s + 3] = __ldg(ph + 1);
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int a = mirror(c + k, g.x0, g.x1);
                q.v[4 * s + k] = (a & 1) ? __ldg(p + g.snx + ((a >> 1) - ch)) : __ldg(p + ((a >> 1) - cl));
            }
        }
    }
    return q;
}

template <typename T, int N>
__device__ __forceinline__ void store_row(T* __restrict__ dst, const Region& g, int row, int a0, const Q<T, N>& r, bool vec_ok) {
    T* p = dst + (size_t)row * g.pitch + (a0 - g.x0);
#pragma unroll
    for (int s = 0; s < N / 4; ++s) {
        const int c = a0 + 4 * s;
        if (vec_ok && c >= g.x0 && c + 3 < g.x1) {
            *reinterpret_cast<int4*>(p + 4 * s) = make_int4(*reinterpret_cast<const int*>(&r.v[4 * s]), *reinterpret_cast<const int*>(&r.v[4 * s + 1]),
                                                            *reinterpret_cast<const int*>(&r.v[4 * s + 2]), *reinterpret_cast<const int*>(&r.v[4 * s + 3]));
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) if (c + k >= g.x0 && c + k < g.x1) p[4 * s + k] = r.v[4 * s + k];
        }
    }
}

// Mallat row index of absolute row n (reflected into the region first)
__device__ __forceinline__ int mallat_row(const Region& g, int n) {
    int r = mirror(n, g.y0, g.y1);
    return (r & 1) ? g.sny + ((r >> 1) - (g.y0 >> 1)) : (r >> 1) - ((g.y0 + 1) >> 1);
}

template <typename T, int N>
__device__ __forceinline__ Q<T, N> load_split_fast(const T* __restrict__ pl, int snx, unsigned gm) {
    Q<T, N> q;
    const T* ph = pl + snx;
    if (N == 8) {
        // owner lanes take both groups with 16-byte loads, halo lanes one group with 8-byte loads
        int4 l = make_int4(0, 0, 0, 0), h = make_int4(0, 0, 0, 0);
        if (gm == 3u) { l = __ldg(reinterpret_cast<const int4*>(pl)); h = __ldg(reinterpret_cast<const int4*>(ph)); }
        else if (gm == 1u) {
            int2 a = __ldg(reinterpret_cast<const int2*>(pl)), b = __ldg(reinterpret_cast<const int2*>(ph));
            l.x = a.x; l.y = a.y; h.x = b.x; h.y = b.y;
        } else if (gm == 2u) {
            int2 a = __ldg(reinterpret_cast<const int2*>(pl + 2)), b = __ldg(reinterpret_cast<const int2*>(ph + 2));
            l.z = a.x; l.w = a.y; h.z = b.x; h.w = b.y;
        }
        q.v[0] = *reinterpret_cast<T*>(&l.x); q.v[2] = *reinterpret_cast<T*>(&l.y);
        q.v[N > 4 ? 4 : 0] = *reinterpret_cast<T*>(&l.z); q.v[N > 4 ? 6 : 0] = *reinterpret_cast<T*>(&l.w);
        q.v[1] = *reinterpret_cast<T*>(&h.x); q.v[3] = *reinterpret_cast<T*>(&h.y);
        q.v[N > 4 ? 5 : 0] = *reinterpret_cast<T*>(&h.z); q.v[N > 4 ? 7 : 0] = *reinterpret_cast<T*>(&h.w);
    } else {
        int2 l = make_int2(0, 0), h = make_int2(0, 0);
        if (gm & 1u) { l = __ldg(reinterpret_cast<const int2*>(pl)); h = __ldg(reinterpret_cast<const int2*>(ph)); }
        q.v[0] = *reinterpret_cast<T*>(&l.x); q.v[2] = *reinterpret_cast<T*>(&l.y);
        q.v[1] = *reinterpret_cast<T*>(&h.x); q.v[3] = *reinterpret_cast<T*>(&h.y);
    }
    return q;
}
template <typename T, int N>
__device__ __forceinline__ void store_row_fast(T* __restrict__ p, const Q<T, N>& r) {
#pragma unroll
    for (int s = 0; s < N / 4; ++s)
        *reinterpret_cast<int4*>(p + 4 * s) = make_int4(*reinterpret_cast<const int*>(&r.v[4 * s]), *reinterpret_cast<const int*>(&r.v[4 * s + 1]),
                                                        *reinterpret_cast<const int*>(&r.v[4 * s + 2]), *reinterpret_cast<const int*>(&r.v[4 * s + 3]));
}

template <class L, int N, int PF, bool FAST>
__device__ __forceinline__ void inverse_rows(const typename L::T* __restrict__ src, typename L::T* __restrict__ dst, const StripCtx& c) {
    typedef typename L::T T;
    const Region& g = c.g;
    const int a0 = c.a0;
    const int oddx = g.x0 & 1;
    const T* __restrict__ scol = src + ((a0 >> 1) - ((g.x0 + 1) >> 1));     // FAST: low-band position
    T* __restrict__ dcol = dst + (a0 - g.x0);
    auto ld = [&](int n) -> Q<T, N> {
        const int mrow = mallat_row(g, n);
        Q<T, N> q = FAST ? load_split_fast<T, N>(scol + (size_t)mrow * g.pitch, g.snx, c.gm)
                         : load_split<T, N>(src, g, mrow, a0, c.gm);
        return q;
    };
    auto hz = [&](Q<T, N>& q) {
        if (FAST || c.wide) L::horizontal(q); else q.v[oddx] = L::single(q.v[oddx], oddx);
    };
    L lift;
    const int n0 = c.e_first + L::LAG - 2 * L::WARM;
    {
        Q<T, N> h = ld(n0 - 1);
        hz(h);
        lift.init(h);
    }
    const int n_end = c.e_last + L::LAG;
    constexpr int D = PF > 0 ? PF : 1;
    Q<T, N> pa[D], pb[D];
    if (PF) {
#pragma unroll
        for (int j = 0; j < D; ++j) if (n0 + 2 * j <= n_end) { pa[j] = ld(n0 + 2 * j); pb[j] = ld(n0 + 2 * j + 1); }
    }
#pragma unroll 1
    for (int nb = n0; nb <= n_end; nb += 2 * D * (PF ? 1 : 2)) {
#pragma unroll
        for (int j = 0; j < (PF ? D : 2); ++j) {
            const int n = nb + 2 * j;
            if (n > n_end) break;
            Q<T, N> a, b;
            if (PF) {
                a = pa[j % D]; b = pb[j % D];
                if (n + 2 * D <= n_end) { pa[j % D] = ld(n + 2 * D); pb[j % D] = ld(n + 2 * D + 1); }
            } else {
                a = ld(n); b = ld(n + 1);
            }
            hz(a); hz(b);
            Q<T, N> ev, od;
            lift.step(a, b, ev, od);
            const int e = n - L::LAG;
            if (e < c.e_first || !c.owner) continue;
            if (FAST) {
                if (e >= g.y0 && e < g.y1) store_row_fast<T, N>(dcol + (size_t)(e - g.y0) * g.pitch, ev);
                if (e + 1 >= g.y0 && e + 1 < g.y1) store_row_fast<T, N>(dcol + (size_t)(e + 1 - g.y0) * g.pitch, od);
            } else {
                if (e >= g.y0 && e < g.y1) store_row<T, N>(dst, g, e - g.y0, a0, ev, c.vec_ok);
                if (e + 1 >= g.y0 && e + 1 < g.y1) store_row<T, N>(dst, g, e + 1 - g.y0, a0, od, c.vec_ok);
            }
        }
    }
}

template <class L, int N, int PF>
__global__ void __launch_bounds__(32 * WARPS_PER_CTA)
k_dwt_inverse(const DwtJob* __restrict__ jobs) {
    typedef typename L::T T;
    const DwtJob job = jobs[blockIdx.y];
    const int lane = threadIdx.x & 31;
    const int item = blockIdx.x * WARPS_PER_CTA + (threadIdx.x >> 5);
    if (item >= job.nstrips * job.nsegs) return;
    const int strip = item % job.nstrips, seg = item / job.nstrips;
    const int seg_pairs = job.strip0;
    StripCtx c;
    Region& g = c.g;
    g.x0 = job.x0; g.y0 = job.y0; g.x1 = job.x1; g.y1 = job.y1; g.pitch = job.pitch;
    g.snx = ((g.x1 + 1) >> 1) - ((g.x0 + 1) >> 1);
    g.sny = ((g.y1 + 1) >> 1) - ((g.y0 + 1) >> 1);
    const T* __restrict__ src = reinterpret_cast<const T*>(job.src);
    T* __restrict__ dst = reinterpret_cast<T*>(job.dst);
    c.vec_ok = ((g.x0 & 1) == 0) && ((g.pitch & 3) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0);
    const int xe = g.x0 & ~1, ye = g.y0 & ~1;
    c.owners = job.strip_cols / N;
    const int s0 = xe + strip * job.strip_cols;
    c.a0 = s0 - N + lane * N;
    c.owner = lane >= 1 && lane <= c.owners;
    c.gm = group_mask<N>(g, c.a0, lane, c.owners);
    c.wide = (g.x1 - g.x0) > 1;
    const int oddx = g.x0 & 1;
    c.e_first = ye + 2 * seg * seg_pairs;
    c.e_last = c.e_first + 2 * (seg_pairs - 1);
    if (c.e_last > g.y1 - 1) c.e_last = (g.y1 - 1) & ~1;

    if (g.y1 - g.y0 == 1) {
        if (seg != 0) return;
        Q<T, N> r = load_split<T, N>(src, g, 0, c.a0, c.gm);
        if (c.wide) L::horizontal(r); else r.v[oddx] = L::single(r.v[oddx], oddx);
        int odd = g.y0 & 1;
#pragma unroll
        for (int k = 0; k < N; ++k) r.v[k] = L::single(r.v[k], odd);
        if (c.owner) store_row<T, N>(dst, g, 0, c.a0, r, c.vec_ok);
        return;
    }
    const int align = N == 8 ? 7 : 3;
    const bool fast = c.vec_ok && s0 - 4 >= g.x0 && s0 + c.owners * N + 4 <= g.x1 && ((s0 - g.x0) & align) == 0 &&
                      ((g.x0 >> 1) & (align >> 1)) == 0 && (g.snx & (align >> 1)) == 0 &&
                      ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
    if (fast) inverse_rows<L, N, PF, true>(src, dst, c);
    else inverse_rows<L, N, PF, false>(src, dst, c);
}

}  // namespace

// Tunables (overridable through the environment for sweeps): columns per lane, software prefetch,
// and the number of warp work items a launch should at least offer the 148 SMs.
struct DwtTune { int ncol, pf_fwd, pf_inv, min_items, owners, gran, min_sp, floor_items; };
static const DwtTune& tune() {
    static DwtTune t = [] {
        DwtTune v{4, 1, 2, 148 * 96, 30, 1, 8, 148 * 16};
        auto geti = [](const char* name, int dflt) { const char* s = getenv(name); return s ? atoi(s) : dflt; };
        v.ncol = geti("J2KCUDA_DWT_NCOL", v.ncol) == 8 ? 8 : 4;
        v.pf_fwd = geti("J2KCUDA_DWT_PF", v.pf_fwd);
        v.pf_inv = geti("J2KCUDA_DWT_PF_INV", getenv("J2KCUDA_DWT_PF") ? v.pf_fwd : v.pf_inv);
        if (v.pf_fwd < 0) v.pf_fwd = 0; if (v.pf_fwd > 3) v.pf_fwd = 3;
        if (v.pf_inv < 0) v.pf_inv = 0; if (v.pf_inv > 3) v.pf_inv = 3;
        v.min_items = geti("J2KCUDA_DWT_ITEMS", v.min_items);
        if (v.min_items < 1) v.min_items = 1;
        v.owners = geti("J2KCUDA_DWT_OWNERS", v.owners);
        if (v.owners < 1 || v.owners > 30) v.owners = 30;
        v.min_sp = geti("J2KCUDA_DWT_MINSP", v.min_sp);
        v.floor_items = geti("J2KCUDA_DWT_FLOOR", v.floor_items);
        v.gran = geti("J2KCUDA_DWT_GRAN", v.gran);
        if (v.gran < 1 || v.gran > v.owners) v.gran = 1;
        return v;
    }();
    return t;
}

// Strips are balanced: the row is cut into chunks of `ncol` columns, at most `owners` chunks (owner
// lanes, <= 30) per warp, and every strip of a job gets the same number of chunks (a multiple of
// `gran`, which keeps strip starts on sector / line boundaries).
static void strip_geometry(const DwtJob& j, int ncol, int& nstrips, int& strip_cols) {
    const DwtTune& t = tune();
    const int xe = j.x0 & ~1;
    const int chunks = (j.x1 - xe + ncol - 1) / ncol;
    nstrips = (chunks + t.owners - 1) / t.owners;
    if (nstrips < 1) nstrips = 1;
    int per = (chunks + nstrips - 1) / nstrips;
    per = (per + t.gran - 1) / t.gran * t.gran;
    if (per > 30) per = 30;
    if (per < 1) per = 1;
    nstrips = (chunks + per - 1) / per;
    if (nstrips < 1) nstrips = 1;
    strip_cols = per * ncol;
}

// Segment length: long enough that the warm-up rows are a few percent, short enough that one
// level of one 4K frame still gives every SM a dozen or more warps.
static int pick_seg_pairs(const DwtJob* jobs, int njobs) {
    const DwtTune& t = tune();
    long long best = 64;
    for (int sp = 64; sp >= 2; sp >>= 1) {
        long long items = 0;
        for (int i = 0; i < njobs; ++i) {
            int ns, sc;
            strip_geometry(jobs[i], t.ncol, ns, sc);
            const int ye = jobs[i].y0 & ~1;
            long long pairs = (((jobs[i].y1 + 1) & ~1) - ye) / 2;
            items += (long long)ns * ((pairs + sp - 1) / sp);
        }
        best = sp;
        // preferred: enough items to hide the row-to-row dependency of every warp; below 8 row pairs per
        // segment the warm-up rows (1 pair for 5/3, 3 for 9/7) cost more than the extra warps bring, so the
        // small levels only go shorter while the launch would not even give every SM `floor_items` warps
        if (items >= t.min_items) break;
        if (sp <= t.min_sp && items >= t.floor_items) break;
    }
    return (int)best;
}

int dwt_plan_jobs(DwtJob* jobs, int njobs) {
    const DwtTune& t = tune();
    int sp = pick_seg_pairs(jobs, njobs);
    int max_items = 0;
    for (int i = 0; i < njobs; ++i) {
        const int ye = jobs[i].y0 & ~1;
        strip_geometry(jobs[i], t.ncol, jobs[i].nstrips, jobs[i].strip_cols);
        int pairs = (((jobs[i].y1 + 1) & ~1) - ye) / 2;
        jobs[i].nsegs = (pairs + sp - 1) / sp;
        jobs[i].strip0 = sp;
        int items = jobs[i].nstrips * jobs[i].nsegs;
        if (items > max_items) max_items = items;
    }
    return max_items;
}

template <template <int> class L, int N>
static void launch_fwd_n(dim3 grid, dim3 block, cudaStream_t st, const DwtJob* d_jobs, int pf) {
    switch (pf) {
    case 0: k_dwt_forward<L<N>, N, 0><<<grid, block, 0, st>>>(d_jobs); break;
    case 1: k_dwt_forward<L<N>, N, 1><<<grid, block, 0, st>>>(d_jobs); break;
    case 2: k_dwt_forward<L<N>, N, 2><<<grid, block, 0, st>>>(d_jobs); break;
    default: k_dwt_forward<L<N>, N, 3><<<grid, block, 0, st>>>(d_jobs); break;
    }
}
template <template <int> class L, int N>
static void launch_inv_n(dim3 grid, dim3 block, cudaStream_t st, const DwtJob* d_jobs, int pf) {
    switch (pf) {
    case 0: k_dwt_inverse<L<N>, N, 0><<<grid, block, 0, st>>>(d_jobs); break;
    case 1: k_dwt_inverse<L<N>, N, 1><<<grid, block, 0, st>>>(d_jobs); break;
    case 2: k_dwt_inverse<L<N>, N, 2><<<grid, block, 0, st>>>(d_jobs); break;
    default: k_dwt_inverse<L<N>, N, 3><<<grid, block, 0, st>>>(d_jobs); break;
    }
}
template <template <int> class L>
static void launch_fwd(dim3 grid, dim3 block, cudaStream_t st, const DwtJob* d_jobs) {
    const DwtTune& t = tune();
    if (t.ncol == 8) launch_fwd_n<L, 8>(grid, block, st, d_jobs, t.pf_fwd); else launch_fwd_n<L, 4>(grid, block, st, d_jobs, t.pf_fwd);
}
template <template <int> class L>
static void launch_inv(dim3 grid, dim3 block, cudaStream_t st, const DwtJob* d_jobs) {
    const DwtTune& t = tune();
    if (t.ncol == 8) launch_inv_n<L, 8>(grid, block, st, d_jobs, t.pf_inv); else launch_inv_n<L, 4>(grid, block, st, d_jobs, t.pf_inv);
}

void launch_dwt_forward(const DwtJob* d_jobs, int njobs, int max_items, int /*seg_pairs*/, int kind, cudaStream_t st) {
    if (njobs <= 0 || max_items <= 0) return;
    dim3 grid((max_items + WARPS_PER_CTA - 1) / WARPS_PER_CTA, njobs), block(32 * WARPS_PER_CTA);
    if (kind == DWT_53) launch_fwd<Fwd53>(grid, block, st, d_jobs);
    else launch_fwd<Fwd97>(grid, block, st, d_jobs);
}

// Measured on B200 (4K RGBA16, profiles/README.md): the fused kernel needs 164 registers (12 warps per SM) and takes 122 us where
// k_mct_forward + level 1 take 46 + 60 us, so it is opt-in (J2KCUDA_FUSE_L1=1) until it beats them.
bool dwt_px_fusion_available() { const char* e = getenv("J2KCUDA_FUSE_L1"); return tune().ncol == 4 && e && e[0] == '1'; }

template <template <int> class L>
static bool launch_fwd_px(dim3 grid, dim3 block, cudaStream_t st, const DwtPxJob* j, const DwtPxSrc& ps, int nc, int nch, int S) {
#define J2K_PX(NC_, NCH_, S_) if (nc == NC_ && nch == NCH_ && S == S_) { k_dwt_forward_px<L<4>, NC_, NCH_, S_><<<grid, block, 0, st>>>(j, ps); return true; }
    J2K_PX(3, 3, 1) J2K_PX(3, 3, 2) J2K_PX(4, 4, 1) J2K_PX(4, 4, 2) J2K_PX(3, 4, 1) J2K_PX(3, 4, 2)
#undef J2K_PX
    return false;
}
bool launch_dwt_forward_px(const DwtPxJob* d_jobs, const DwtPxSrc& src, int njobs, int max_items, int kind, int ncomp, int nch, int sample_bytes,
                           cudaStream_t st) {
    if (njobs <= 0 || max_items <= 0) return false;
    dim3 grid((max_items + WARPS_PER_CTA - 1) / WARPS_PER_CTA, njobs), block(32 * WARPS_PER_CTA);
    return kind == DWT_53 ? launch_fwd_px<Fwd53>(grid, block, st, d_jobs, src, ncomp, nch, sample_bytes)
                          : launch_fwd_px<Fwd97>(grid, block, st, d_jobs, src, ncomp, nch, sample_bytes);
}

void launch_dwt_inverse(const DwtJob* d_jobs, int njobs, int max_items, int /*seg_pairs*/, int kind, cudaStream_t st) {
    if (njobs <= 0 || max_items <= 0) return;
    dim3 grid((max_items + WARPS_PER_CTA - 1) / WARPS_PER_CTA, njobs), block(32 * WARPS_PER_CTA);
    if (kind == DWT_53) launch_inv<Inv53>(grid, block, st, d_jobs);
    else launch_inv<Inv97>(grid, block, st, d_jobs);
}

}  // namespace j2k

// ---- j2kcuda_dwt_forward_noalloc: the forward transform for callers that cannot allocate -----------------------------
// No counterpart in the reference (include/j2kcuda.h).  The multi-level
// forward transform needs a second plane to ping-pong with and a table of per-level job descriptors
// in device memory.  The existing paths cudaMalloc both; this one keeps them in a file-scope device array of fixed size,
// which every GPU that loads libj2kcuda.so holds until the process ends.
//
//   level 1:            caller's input  -> caller's output (all four bands of level 1 at their Mallat position)
//   level 2, 4, ...:    output's LL     -> scratch plane   (same pitch as the output, only the rows of LL1 exist)
//   level 3, 5, ...:    scratch's LL    -> output
//   afterwards the detail bands of the even levels (and the last LL when the level count is even) are copied from the
//   scratch plane to their place in the output (device-to-device 2D copies on the same stream).
//
// Nothing in the scratch has to start at zero: every sample a level reads was written by the level before it in the same
// call.  Every call on a device uses the same array: two calls whose work overlaps on the GPU (different streams without
// an event between them, or different host threads) race on it.
#include "../../include/j2kcuda.h"
#include "common.h"
#include "geometry.h"
#include <mutex>

namespace {

constexpr size_t kNoAllocHeaderBytes = 4096;                              // job table: 32 levels x sizeof(DwtJob)
constexpr size_t kNoAllocPlaneSamples = (size_t)J2KCUDA_NOALLOC_MAX_SCRATCH_SAMPLES;
static_assert(32 * sizeof(j2k::DwtJob) <= kNoAllocHeaderBytes, "job table does not fit the scratch header");

static __device__ __align__(256) unsigned char g_dwt_noalloc_scratch[kNoAllocHeaderBytes + kNoAllocPlaneSamples * 4];

struct NoAllocJobs { j2k::DwtJob job[32]; int n; };

// Device code may name the array directly: the job table goes in by value (a kernel parameter, no host-to-device copy)
__global__ void k_dwt_noalloc_jobs(NoAllocJobs t) {
    j2k::DwtJob* dst = reinterpret_cast<j2k::DwtJob*>(g_dwt_noalloc_scratch);
    for (int i = threadIdx.x; i < t.n; i += blockDim.x) dst[i] = t.job[i];
}

// Host code needs the array's DEVICE address (the symbol's name is only a host-side stand-in): asked for once per device,
// with that device current.
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[64] = {};

int noalloc_scratch(unsigned char** out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        cudaGetLastError();
        return j2k::set_error(J2KCUDA_ERR_NO_DEVICE, "no CUDA device: libj2kcuda has no CPU path");
    }
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_dwt_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return j2k::set_error(J2KCUDA_ERR_CUDA, std::string("cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(e));
        }
        g_noalloc_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_noalloc_ptr[dev];
    return J2KCUDA_OK;
}

}  // namespace

extern "C" int j2kcuda_dwt_forward_noalloc(const int32_t* d_in, int32_t* d_out, int pitch, int x0, int y0, int x1, int y1,
                                           int numres, int irreversible, void* stream) { return j2k::guarded([&]() -> int {
    using namespace j2k;
    if (!d_in || !d_out || d_in == d_out) return set_error(J2KCUDA_ERR_PARAM, "j2kcuda_dwt_forward_noalloc: null or aliased plane");
    if (x0 < 0 || y0 < 0 || x1 <= x0 || y1 <= y0 || numres < 1 || numres > 33) return set_error(J2KCUDA_ERR_PARAM, "j2kcuda_dwt_forward_noalloc: bad region");
    const int w = x1 - x0, h = y1 - y0, levels = numres - 1;
    if (pitch < w || (pitch & 3) || (((uintptr_t)d_in | (uintptr_t)d_out) & 15))
        return set_error(J2KCUDA_ERR_PARAM, "j2kcuda_dwt_forward_noalloc: pitch must be >= width and a multiple of 4 samples, planes 16-byte aligned");
    // rows of the level-1 LL band = rows the scratch plane needs (it has the output's pitch)
    const int ll_rows = (int)ceildivpow2((uint32_t)y1, 1) - (int)ceildivpow2((uint32_t)y0, 1);
    if (levels >= 2 && (size_t)pitch * (size_t)ll_rows > kNoAllocPlaneSamples)
        return set_error(J2KCUDA_ERR_UNSUPPORTED, "j2kcuda_dwt_forward_noalloc: pitch * ceil(height / 2) = " + std::to_string((size_t)pitch * (size_t)ll_rows) +
                         " samples exceeds the fixed scratch of J2KCUDA_NOALLOC_MAX_SCRATCH_SAMPLES = " + std::to_string(kNoAllocPlaneSamples) +
                         "; use j2kcuda_stage_dwt or an encoder context for larger planes");
    unsigned char* base = nullptr;
    if (int rc = noalloc_scratch(&base)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    DwtJob* d_jobs = reinterpret_cast<DwtJob*>(base);
    int32_t* scratch = reinterpret_cast<int32_t*>(base + kNoAllocHeaderBytes);
    const int kind = irreversible ? DWT_97_FIX : DWT_53;

    if (levels == 0) {                                   // nothing to transform: the output is the input
        if (cudaMemcpy2DAsync(d_out, (size_t)pitch * 4, d_in, (size_t)pitch * 4, (size_t)w * 4, (size_t)h, cudaMemcpyDeviceToDevice, st) != cudaSuccess)
            return set_error(J2KCUDA_ERR_CUDA, cudaGetErrorString(cudaGetLastError()));
        return J2KCUDA_OK;
    }
    // plane(0) = input, odd levels land in the output, even levels in the scratch
    auto plane = [&](int l) -> int32_t* { return l == 0 ? const_cast<int32_t*>(d_in) : (l & 1) ? d_out : scratch; };
    NoAllocJobs tab{};
    int items[32], done = 0;
    for (int l = 1; l <= levels && l <= 32; ++l) {
        DwtJob& j = tab.job[tab.n];
        j.x0 = (int)ceildivpow2((uint32_t)x0, l - 1); j.x1 = (int)ceildivpow2((uint32_t)x1, l - 1);
        j.y0 = (int)ceildivpow2((uint32_t)y0, l - 1); j.y1 = (int)ceildivpow2((uint32_t)y1, l - 1);
        j.pitch = pitch; j.strip0 = j.nstrips = j.nsegs = 0; j.strip_cols = 0; j.pad = 0;
        if (j.x1 <= j.x0 || j.y1 <= j.y0) break;         // every deeper level is empty as well
        j.src = plane(l - 1); j.dst = plane(l);
        items[tab.n] = dwt_plan_jobs(&j, 1);
        ++tab.n; done = l;
    }
    k_dwt_noalloc_jobs<<<1, 32, 0, st>>>(tab);
    for (int i = 0; i < tab.n; ++i) launch_dwt_forward(d_jobs + i, 1, items[i], tab.job[i].strip0, kind, st);
    if (cudaError_t e = cudaGetLastError()) return set_error(J2KCUDA_ERR_CUDA, std::string("j2kcuda_dwt_forward_noalloc: ") + cudaGetErrorString(e));
    // even levels: their detail bands (and the final LL when the deepest level is even) move from the scratch to the output
    for (int l = 2; l <= done; l += 2) {
        const DwtJob& j = tab.job[l - 1];
        const int rw = j.x1 - j.x0, rh = j.y1 - j.y0;
        const int snx = (int)ceildivpow2((uint32_t)j.x1, 1) - (int)ceildivpow2((uint32_t)j.x0, 1);
        const int sny = (int)ceildivpow2((uint32_t)j.y1, 1) - (int)ceildivpow2((uint32_t)j.y0, 1);
        cudaError_t e = cudaSuccess;
        if (l == done) {
            e = cudaMemcpy2DAsync(d_out, (size_t)pitch * 4, scratch, (size_t)pitch * 4, (size_t)rw * 4, (size_t)rh, cudaMemcpyDeviceToDevice, st);
        } else {
            if (rw > snx) e = cudaMemcpy2DAsync(d_out + snx, (size_t)pitch * 4, scratch + snx, (size_t)pitch * 4, (size_t)(rw - snx) * 4, (size_t)rh, cudaMemcpyDeviceToDevice, st);
            if (e == cudaSuccess && rh > sny && snx > 0)
                e = cudaMemcpy2DAsync(d_out + (size_t)sny * pitch, (size_t)pitch * 4, scratch + (size_t)sny * pitch, (size_t)pitch * 4, (size_t)snx * 4, (size_t)(rh - sny), cudaMemcpyDeviceToDevice, st);
        }
        if (e != cudaSuccess) { cudaGetLastError(); return set_error(J2KCUDA_ERR_CUDA, std::string("j2kcuda_dwt_forward_noalloc: ") + cudaGetErrorString(e)); }
    }
    return J2KCUDA_OK;
}); }
