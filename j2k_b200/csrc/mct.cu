// Sample unpacking + DC level shift + RCT/ICT fused into one pass (and the inverse: RCT/ICT inverse +
// level shift + clamp + packing).  HBM-bound: every thread moves four horizontally adjacent pixels,
// reading the caller's interleaved samples with 16/8/4-byte vector loads and writing one int4 per
// component plane.
// Reference arithmetic: ext/grok/src/lib/openjp2/tcd.cpp:1659-1696 (shift, 11 fractional bits for
// 9/7), mct.cpp:94-157 / 163-227 (RCT), :241-401 (ICT, Q13), :406-471 (inverse ICT, float),
// tcd.cpp:1495-1556 (inverse shift, lrintf, clamp); sample conversion src/common/j2k_codec.cpp:222-427.
#include "device.h"

namespace j2k {
namespace {

__device__ __forceinline__ int fm(int a, int b) { return (int)(((long long)a * (long long)b + 4096) >> 13); }

__device__ __forceinline__ void mct_fwd_px(int& a, int& b, int& c, int irreversible) {
    int r = a, g = b, bl = c;
    if (!irreversible) { a = (r + 2 * g + bl) >> 2; b = bl - g; c = r - g; }
    else {
        a = fm(r, 2449) + fm(g, 4809) + fm(bl, 934);
        b = -fm(r, 1382) - fm(g, 2714) + fm(bl, 4096);
        c = fm(r, 4096) - fm(g, 3430) - fm(bl, 666);
    }
}
__device__ __forceinline__ void rct_inv_px(int& a, int& b, int& c) {
    int y = a, u = b, v = c;
    int g = y - ((u + v) >> 2);
    a = v + g; b = g; c = u + g;
}
__device__ __forceinline__ void ict_inv_px(float& a, float& b, float& c) {
    float y = a, u = b, v = c;
    a = __fadd_rn(y, __fmul_rn(v, 1.402f));
    b = __fsub_rn(__fsub_rn(y, __fmul_rn(u, 0.34413f)), __fmul_rn(v, 0.71414f));
    c = __fadd_rn(y, __fmul_rn(u, 1.772f));
}

template <int S> __device__ __forceinline__ int load_sample(const uint8_t* p, int sgnd) {
    if (S == 1) return sgnd ? (int)*reinterpret_cast<const int8_t*>(p) : (int)*p;
    if (S == 2) return sgnd ? (int)*reinterpret_cast<const int16_t*>(p) : (int)*reinterpret_cast<const uint16_t*>(p);
    return *reinterpret_cast<const int*>(p);
}
template <int S> __device__ __forceinline__ void store_sample(uint8_t* p, int v) {
    if (S == 1) *p = (uint8_t)v;
    else if (S == 2) *reinterpret_cast<uint16_t*>(p) = (uint16_t)v;
    else *reinterpret_cast<int*>(p) = v;
}
template <int S> __device__ __forceinline__ int word_sample(const uint32_t* w, int idx, int sgnd) {   // idx = sample index
    if (S == 1) { uint32_t v = (w[idx >> 2] >> ((idx & 3) * 8)) & 0xff; return sgnd ? (int)(int8_t)v : (int)v; }
    if (S == 2) { uint32_t v = (w[idx >> 1] >> ((idx & 1) * 16)) & 0xffff; return sgnd ? (int)(int16_t)v : (int)v; }
    return (int)w[idx];
}

struct FwdArgs {
    PixelSrc src;
    MctParams mp;
    int32_t* planes[4];
    int nch;            // channels per pixel when interleaved, else 0
    int chan[4];        // channel index of component c
};

// S = sample bytes, NCH = interleaved channel count (0: generic strided access)
template <int S, int NCH>
__global__ void __launch_bounds__(256) k_mct_forward(const FwdArgs a) {
    const int quads = (a.mp.width + 3) >> 2;
    const long long total = (long long)quads * a.mp.height;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int y = (int)(i / quads), x = (int)(i % quads) * 4;
        const int nv = min(4, a.mp.width - x);
        int v[4][4];
        if (NCH > 0 && nv == 4) {
            constexpr int NW = (NCH > 0 ? NCH : 1) * S;          // 32-bit words for four pixels
            uint32_t w[NW];
            const uint8_t* p = a.src.base[0] + (long long)y * a.src.rowbytes[0] + (long long)x * NCH * S;
            if (NW % 4 == 0) {
#pragma unroll
                for (int k = 0; k < NW / 4; ++k) {
                    uint4 t = __ldg(reinterpret_cast<const uint4*>(p) + k);
                    w[4 * k] = t.x; w[4 * k + 1] = t.y; w[4 * k + 2] = t.z; w[4 * k + 3] = t.w;
                }
            } else if (NW % 2 == 0) {
#pragma unroll
                for (int k = 0; k < NW / 2; ++k) {
                    uint2 t = __ldg(reinterpret_cast<const uint2*>(p) + k);
                    w[2 * k] = t.x; w[2 * k + 1] = t.y;
                }
            } else {
#pragma unroll
                for (int k = 0; k < NW; ++k) w[k] = __ldg(reinterpret_cast<const uint32_t*>(p) + k);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int ch[4];
#pragma unroll
                for (int j = 0; j < (NCH > 0 ? NCH : 1); ++j) ch[j] = word_sample<S>(w, k * NCH + j, 0);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    int sel = a.chan[c];
                    int s = ch[0];
                    if (NCH > 1 && sel == 1) s = ch[1];
                    if (NCH > 2 && sel == 2) s = ch[2];
                    if (NCH > 3 && sel == 3) s = ch[3];
                    v[c][k] = s;
                }
            }
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (c >= a.mp.ncomp) continue;
                const uint8_t* p = a.src.base[c] + (long long)y * a.src.rowbytes[c] + (long long)x * a.src.colbytes[c];
#pragma unroll
                for (int k = 0; k < 4; ++k) v[c][k] = k < nv ? load_sample<S>(p + (long long)k * a.src.colbytes[c], 0) : 0;
            }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (c >= a.mp.ncomp) continue;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int s = v[c][k];
                if (!a.mp.conv[c].plain) s = (int)copy_convert(a.mp.conv[c].src_type == 2 ? (long long)(uint32_t)s : (long long)s, a.mp.conv[c]);
                s -= a.mp.dc_shift[c];
                v[c][k] = a.mp.irreversible ? s << 11 : s;
            }
        }
        if (a.mp.mct && a.mp.ncomp >= 3) {
#pragma unroll
            for (int k = 0; k < 4; ++k) mct_fwd_px(v[0][k], v[1][k], v[2][k], a.mp.irreversible);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (c >= a.mp.ncomp) continue;
            int32_t* q = a.planes[c] + (long long)y * a.mp.pitch + x;
            if (nv == 4) *reinterpret_cast<int4*>(q) = make_int4(v[c][0], v[c][1], v[c][2], v[c][3]);
            else for (int k = 0; k < nv; ++k) q[k] = v[c][k];
        }
    }
}

struct InvArgs {
    PixelDst dst;
    MctParams mp;
    const int32_t* planes[4];
    int nch;
    int chan[4];
};

template <int S, int NCH>
__global__ void __launch_bounds__(256) k_mct_inverse(const InvArgs a) {
    const int quads = (a.mp.width + 3) >> 2;
    const long long total = (long long)quads * a.mp.height;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int y = (int)(i / quads), x = (int)(i % quads) * 4;
        const int nv = min(4, a.mp.width - x);
        int v[4][4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (c >= a.mp.ncomp) { v[c][0] = v[c][1] = v[c][2] = v[c][3] = 0; continue; }
            const int32_t* q = a.planes[c] + (long long)y * a.mp.pitch + x;
            if (nv == 4) { int4 t = __ldg(reinterpret_cast<const int4*>(q)); v[c][0] = t.x; v[c][1] = t.y; v[c][2] = t.z; v[c][3] = t.w; }
            else for (int k = 0; k < 4; ++k) v[c][k] = k < nv ? q[k] : 0;
        }
        if (a.mp.irreversible) {
            // planes hold float bit patterns (dwt.cpp:927-1062 leaves floats; tcd.cpp:1540-1548)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float f0 = __int_as_float(v[0][k]), f1 = __int_as_float(v[1][k]), f2 = __int_as_float(v[2][k]);
                if (a.mp.mct && a.mp.ncomp >= 3) ict_inv_px(f0, f1, f2);
                v[0][k] = (int)lrintf(f0); v[1][k] = (int)lrintf(f1); v[2][k] = (int)lrintf(f2);
                v[3][k] = (int)lrintf(__int_as_float(v[3][k]));
            }
        } else if (a.mp.mct && a.mp.ncomp >= 3) {
#pragma unroll
            for (int k = 0; k < 4; ++k) rct_inv_px(v[0][k], v[1][k], v[2][k]);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (c >= a.mp.ncomp) continue;
            const int lo = a.dst.sgnd ? -(1 << (a.mp.prec[c] - 1)) : 0;
            const int hi = a.dst.sgnd ? (1 << (a.mp.prec[c] - 1)) - 1 : (1 << a.mp.prec[c]) - 1;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                v[c][k] = min(max(v[c][k] + a.mp.dc_shift[c], lo), hi);
            }
        }
        if (NCH > 0 && nv == 4) {
            constexpr int NC = NCH > 0 ? NCH : 1;
            constexpr int NW = NC * S;
            uint32_t w[NW];
#pragma unroll
            for (int k = 0; k < NW; ++k) w[k] = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int j = 0; j < NC; ++j) {
                    int s = 0;
#pragma unroll
                    for (int c = 0; c < 4; ++c) if (c < NC && a.chan[c] == j) s = v[c][k];
                    const int idx = k * NC + j;
                    if (S == 1) w[idx >> 2] |= (uint32_t)(s & 0xff) << ((idx & 3) * 8);
                    else if (S == 2) w[idx >> 1] |= (uint32_t)(s & 0xffff) << ((idx & 1) * 16);
                    else w[idx] = (uint32_t)s;
                }
            uint8_t* p = a.dst.base[0] + (long long)y * a.dst.rowbytes[0] + (long long)x * NC * S;
            if (NW % 4 == 0) {
#pragma unroll
                for (int k = 0; k < NW / 4; ++k) reinterpret_cast<uint4*>(p)[k] = make_uint4(w[4 * k], w[4 * k + 1], w[4 * k + 2], w[4 * k + 3]);
            } else if (NW % 2 == 0) {
#pragma unroll
                for (int k = 0; k < NW / 2; ++k) reinterpret_cast<uint2*>(p)[k] = make_uint2(w[2 * k], w[2 * k + 1]);
            } else {
#pragma unroll
                for (int k = 0; k < NW; ++k) reinterpret_cast<uint32_t*>(p)[k] = w[k];
            }
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (c >= a.mp.ncomp) continue;
                uint8_t* p = a.dst.base[c] + (long long)y * a.dst.rowbytes[c] + (long long)x * a.dst.colbytes[c];
                for (int k = 0; k < nv; ++k) store_sample<S>(p + (long long)k * a.dst.colbytes[c], v[c][k]);
            }
        }
    }
}

__global__ void __launch_bounds__(256) k_copy_channel(const uint8_t* __restrict__ img, int S, int ncomp, int c, int sgnd, int img_w, int w, int h,
                                                      int sx, int sy, uint8_t* __restrict__ dst, const CopyConv cc) {
    const long long total = (long long)w * h;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int y = (int)(i / w), x = (int)(i % w);
        const uint8_t* p = img + (((long long)y * sy * img_w + (long long)x * sx) * ncomp + c) * S;
        long long s = S == 1 ? (sgnd ? (long long)*reinterpret_cast<const int8_t*>(p) : (long long)*p)
                             : (sgnd ? (long long)*reinterpret_cast<const int16_t*>(p) : (long long)*reinterpret_cast<const uint16_t*>(p));
        const long long r = cc.plain ? s : copy_convert(s, cc);
        if (cc.dst_type == 0) dst[i] = (uint8_t)r;
        else if (cc.dst_type == 1) reinterpret_cast<uint16_t*>(dst)[i] = (uint16_t)r;
        else reinterpret_cast<uint32_t*>(dst)[i] = (uint32_t)r;
    }
}

__global__ void __launch_bounds__(256) k_output_subsampled(const SubOut a, uint8_t* __restrict__ out) {
    const long long total = (long long)a.width * a.height;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int y = (int)(i / a.width), x = (int)(i % a.width);
        for (int c = 0; c < a.ncomp; ++c) {
            const int xs = min(x / a.dx[c], a.cw[c] - 1), ys = min(y / a.dy[c], a.ch[c] - 1);
            int v = a.plane[c][(long long)ys * a.pitch + xs];
            if (a.irreversible) v = (int)lrintf(__int_as_float(v));
            const int lo = a.sgnd ? -(1 << (a.prec[c] - 1)) : 0, hi = a.sgnd ? (1 << (a.prec[c] - 1)) - 1 : (1 << a.prec[c]) - 1;
            v = min(max(v + a.dc_shift[c], lo), hi);
            uint8_t* p = out + (i * a.ncomp + c) * a.sample_bytes;
            if (a.sample_bytes == 1) *p = (uint8_t)v; else *reinterpret_cast<uint16_t*>(p) = (uint16_t)v;
        }
    }
}

__global__ void __launch_bounds__(256) k_mct_planes(int32_t* c0, int32_t* c1, int32_t* c2, size_t n, int irreversible, int inverse) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        int a = c0[i], b = c1[i], c = c2[i];
        if (!inverse) mct_fwd_px(a, b, c, irreversible);
        else if (!irreversible) rct_inv_px(a, b, c);
        else {
            float fa = __int_as_float(a), fb = __int_as_float(b), fc = __int_as_float(c);
            ict_inv_px(fa, fb, fc);
            a = __float_as_int(fa); b = __float_as_int(fb); c = __float_as_int(fc);
        }
        c0[i] = a; c1[i] = b; c2[i] = c;
    }
}

// Are the components channels of one interleaved pixel array?  Returns channels per pixel or 0.
template <class P>
int detect_interleave(const P& s, int ncomp, int chan[4]) {
    const int S = s.sample_bytes;
    const uint8_t* lo = s.base[0];
    for (int c = 1; c < ncomp; ++c) if (s.base[c] < lo) lo = s.base[c];
    int64_t cb = s.colbytes[0];
    if (cb % S || cb / S < ncomp || cb / S > 4 || cb / S < 3) return 0;
    int nch = (int)(cb / S);
    int used = 0;
    for (int c = 0; c < 4; ++c) chan[c] = 0;
    for (int c = 0; c < ncomp; ++c) {
        if (s.colbytes[c] != cb || s.rowbytes[c] != s.rowbytes[0]) return 0;
        int64_t d = s.base[c] - lo;
        if (d % S || d / S >= nch) return 0;
        chan[c] = (int)(d / S);
        if (used & (1 << chan[c])) return 0;
        used |= 1 << chan[c];
    }
    // vector width used by the kernels: 4*nch*S bytes per thread, aligned to its largest
    // power-of-two divisor up to 16
    int bytes = 4 * nch * S;
    int align = bytes % 16 == 0 ? 16 : (bytes % 8 == 0 ? 8 : 4);
    if ((reinterpret_cast<uintptr_t>(lo) % align) || (s.rowbytes[0] % align)) return 0;
    return nch;
}

int grid_for(long long threads) {
    long long blocks = (threads + 255) / 256;
    const long long cap = 148LL * 8;          // 8 resident CTAs of 256 threads per SM, grid-stride beyond
    return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace

#define MCT_DISPATCH(KERNEL, ARGS, S, NCH, GRID, ST)                                              \
    do {                                                                                          \
        if (S == 1) { if (NCH == 3) KERNEL<1, 3><<<GRID, 256, 0, ST>>>(ARGS); else if (NCH == 4) KERNEL<1, 4><<<GRID, 256, 0, ST>>>(ARGS); else KERNEL<1, 0><<<GRID, 256, 0, ST>>>(ARGS); } \
        else if (S == 2) { if (NCH == 3) KERNEL<2, 3><<<GRID, 256, 0, ST>>>(ARGS); else if (NCH == 4) KERNEL<2, 4><<<GRID, 256, 0, ST>>>(ARGS); else KERNEL<2, 0><<<GRID, 256, 0, ST>>>(ARGS); } \
        else KERNEL<4, 0><<<GRID, 256, 0, ST>>>(ARGS);                                            \
    } while (0)

int detect_pixel_interleave(const PixelSrc& s, int ncomp, int chan[4]) { return detect_interleave(s, ncomp, chan); }

void launch_mct_forward(const PixelSrc& src, const MctParams& mp, int32_t* const planes[4], cudaStream_t st) {
    FwdArgs a;
    a.src = src; a.mp = mp;
    for (int c = 0; c < 4; ++c) a.planes[c] = planes[c];
    a.nch = detect_interleave(src, mp.ncomp, a.chan);
    if (a.nch) { const uint8_t* lo = src.base[0]; for (int c = 1; c < mp.ncomp; ++c) if (src.base[c] < lo) lo = src.base[c]; a.src.base[0] = lo; }
    int grid = grid_for((long long)((mp.width + 3) / 4) * mp.height);
    MCT_DISPATCH(k_mct_forward, a, src.sample_bytes, a.nch, grid, st);
}

void launch_mct_inverse(const PixelDst& dst, const MctParams& mp, const int32_t* const planes[4], cudaStream_t st) {
    InvArgs a;
    a.dst = dst; a.mp = mp;
    for (int c = 0; c < 4; ++c) a.planes[c] = planes[c];
    a.nch = detect_interleave(dst, mp.ncomp, a.chan);
    if (a.nch != mp.ncomp) a.nch = 0;          // never write a channel the codestream does not carry
    if (a.nch) { uint8_t* lo = dst.base[0]; for (int c = 1; c < mp.ncomp; ++c) if (dst.base[c] < lo) lo = dst.base[c]; a.dst.base[0] = lo; }
    int grid = grid_for((long long)((mp.width + 3) / 4) * mp.height);
    MCT_DISPATCH(k_mct_inverse, a, dst.sample_bytes, a.nch, grid, st);
}

void launch_copy_channel(const uint8_t* img, int S, int ncomp, int c, int sgnd, int img_w, int w, int h, int sx, int sy, uint8_t* dst,
                         const CopyConv& conv, cudaStream_t st) {
    if (w <= 0 || h <= 0) return;
    k_copy_channel<<<grid_for((long long)w * h), 256, 0, st>>>(img, S, ncomp, c, sgnd, img_w, w, h, sx, sy, dst, conv);
}

void launch_output_subsampled(const SubOut& so, uint8_t* d_out, cudaStream_t st) {
    if (so.width <= 0 || so.height <= 0) return;
    k_output_subsampled<<<grid_for((long long)so.width * so.height), 256, 0, st>>>(so, d_out);
}

void launch_mct_planes(int32_t* c0, int32_t* c1, int32_t* c2, size_t n, int irreversible, int inverse, cudaStream_t st) {
    if (!n) return;
    k_mct_planes<<<grid_for((long long)n), 256, 0, st>>>(c0, c1, c2, n, irreversible, inverse);
}

}  // namespace j2k
