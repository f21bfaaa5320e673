/*
 * j2k_oracle.c -- TEST INFRASTRUCTURE ONLY.  See j2k_oracle.h.
 *
 * CPU restatement of the reference hot path (ext/grok/src/lib/openjp2).  Each function cites the
 * reference lines it follows.  Parity: pinned against oracle/_ref/libgrokref.so (the reference's
 * own sources compiled here) by tests/test_oracle_vs_ref.py and against tests/golden/ fixtures.
 * The product (libj2kcuda) never links or calls this file.
 */
#include "j2k_oracle.h"
#include <stdlib.h>
#include <string.h>
#include <math.h>

/* ------------------------------------------------------------------------------------------ */
/* fixed point helpers: opj_intmath.h:255-279                                                  */
static inline int32_t fix_mul(int32_t a, int32_t b) {
    int64_t t = (int64_t)a * (int64_t)b + 4096;
    return (int32_t)(t >> 13);
}
static inline int32_t fix_mul_t1(int32_t a, int32_t b) {
    int64_t t = (int64_t)a * (int64_t)b + 4096;
    return (int32_t)(t >> (13 + 11 - 6));
}
static inline int ceildivpow2(int a, int b) { return (int)(((int64_t)a + ((int64_t)1 << b) - 1) >> b); }

/* ------------------------------------------------------------------------------------------ */
/* MCT                                                                                         */
void orc_rct_fwd(int32_t* c0, int32_t* c1, int32_t* c2, size_t n) {      /* mct.cpp:94-157 */
    for (size_t i = 0; i < n; ++i) {
        int32_t r = c0[i], g = c1[i], b = c2[i];
        c0[i] = (r + 2 * g + b) >> 2; c1[i] = b - g; c2[i] = r - g;
    }
}
void orc_rct_inv(int32_t* c0, int32_t* c1, int32_t* c2, size_t n) {      /* mct.cpp:163-227 */
    for (size_t i = 0; i < n; ++i) {
        int32_t y = c0[i], u = c1[i], v = c2[i];
        int32_t g = y - ((u + v) >> 2);
        c0[i] = v + g; c1[i] = g; c2[i] = u + g;
    }
}
void orc_ict_fwd(int32_t* c0, int32_t* c1, int32_t* c2, size_t n) {      /* mct.cpp:241-401 */
    for (size_t i = 0; i < n; ++i) {
        int32_t r = c0[i], g = c1[i], b = c2[i];
        c0[i] = fix_mul(r, 2449) + fix_mul(g, 4809) + fix_mul(b, 934);
        c1[i] = -fix_mul(r, 1382) - fix_mul(g, 2714) + fix_mul(b, 4096);
        c2[i] = fix_mul(r, 4096) - fix_mul(g, 3430) - fix_mul(b, 666);
    }
}
void orc_ict_inv(float* c0, float* c1, float* c2, size_t n) {            /* mct.cpp:406-471 */
    for (size_t i = 0; i < n; ++i) {
        float y = c0[i], u = c1[i], v = c2[i];
        float r = y + (v * 1.402f);
        float g = y - (u * 0.34413f) - (v * 0.71414f);
        float b = y + (u * 1.772f);
        c0[i] = r; c1[i] = g; c2[i] = b;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* DWT.  One 1-D line is handled in absolute canvas coordinates: sample at even coordinate is  */
/* low-pass, odd is high-pass; whole-sample symmetric extension at both ends equals the index   */
/* clamping of dwt.cpp:182-188.                                                                */
static inline int mirror(int i, int lo, int hi) { /* reflect i into [lo, hi) */
    if (hi - lo == 1) return lo;
    while (i < lo || i >= hi) {
        if (i < lo) i = 2 * lo - i;
        if (i >= hi) i = 2 * (hi - 1) - i;
    }
    return i;
}
#define X(i) x[mirror((i), u0, u1) - u0]

/* dwt.cpp:313-330 */
static void lift53_fwd(int32_t* x, int u0, int u1) {
    if (u1 - u0 == 1) { if (u0 & 1) x[0] *= 2; return; }
    for (int i = u0; i < u1; ++i) if (i & 1) x[i - u0] -= (X(i - 1) + X(i + 1)) >> 1;
    for (int i = u0; i < u1; ++i) if (!(i & 1)) x[i - u0] += (X(i - 1) + X(i + 1) + 2) >> 2;
}
/* dwt.cpp:335-352 */
static void lift53_inv(int32_t* x, int u0, int u1) {
    if (u1 - u0 == 1) { if (u0 & 1) x[0] /= 2; return; }
    for (int i = u0; i < u1; ++i) if (!(i & 1)) x[i - u0] -= (X(i - 1) + X(i + 1) + 2) >> 2;
    for (int i = u0; i < u1; ++i) if (i & 1) x[i - u0] += (X(i - 1) + X(i + 1)) >> 1;
}
/* dwt.cpp:365-399 (Q13 constants) */
static void lift97_fwd(int32_t* x, int u0, int u1) {
    if (u1 - u0 == 1) return;
    for (int i = u0; i < u1; ++i) if (i & 1) x[i - u0] -= fix_mul(X(i - 1) + X(i + 1), 12993);
    for (int i = u0; i < u1; ++i) if (!(i & 1)) x[i - u0] -= fix_mul(X(i - 1) + X(i + 1), 434);
    for (int i = u0; i < u1; ++i) if (i & 1) x[i - u0] += fix_mul(X(i - 1) + X(i + 1), 7233);
    for (int i = u0; i < u1; ++i) if (!(i & 1)) x[i - u0] += fix_mul(X(i - 1) + X(i + 1), 3633);
    for (int i = u0; i < u1; ++i) x[i - u0] = fix_mul(x[i - u0], (i & 1) ? 5038 : 6659);
}
#undef X
/* dwt.cpp:769-921; constants :94-100.  The reference adds 2*c*neighbour at a boundary where one
 * neighbour is missing; with symmetric extension (l + r) * c has l == r, i.e. (2l)*c.  The
 * reference computes (c+c)*l there; both are exact doublings so the products agree. */
static void lift97_inv(float* x, int u0, int u1) {
    const float K = 1.230174105f, c13318 = 1.625732422f;
    const float alpha = 1.586134342f, beta = 0.052980118f, gamma = -0.882911075f, delta = -0.443506852f;
    int n = u1 - u0;
    if (n == 1) return;
#define XF(i) x[mirror((i), u0, u1) - u0]
    for (int i = u0; i < u1; ++i) x[i - u0] *= (i & 1) ? c13318 : K;
    for (int i = u0; i < u1; ++i) if (!(i & 1)) x[i - u0] += (XF(i - 1) + XF(i + 1)) * delta;
    for (int i = u0; i < u1; ++i) if (i & 1) x[i - u0] += (XF(i - 1) + XF(i + 1)) * gamma;
    for (int i = u0; i < u1; ++i) if (!(i & 1)) x[i - u0] += (XF(i - 1) + XF(i + 1)) * beta;
    for (int i = u0; i < u1; ++i) if (i & 1) x[i - u0] += (XF(i - 1) + XF(i + 1)) * alpha;
#undef XF
}

/* split interleaved line (canvas coords u0..u1) into [lows | highs]; dwt.cpp:219-263 */
static void deinterleave(const int32_t* x, int32_t* dst, int stride, int u0, int u1) {
    int sn = ceildivpow2(u1, 1) - ceildivpow2(u0, 1);
    int il = 0, ih = 0;
    for (int i = u0; i < u1; ++i) {
        if (i & 1) dst[(size_t)(sn + ih++) * stride] = x[i - u0];
        else dst[(size_t)(il++) * stride] = x[i - u0];
    }
}
static void interleave(const int32_t* src, int stride, int32_t* x, int u0, int u1) { /* :268-307 */
    int sn = ceildivpow2(u1, 1) - ceildivpow2(u0, 1);
    int il = 0, ih = 0;
    for (int i = u0; i < u1; ++i) {
        if (i & 1) x[i - u0] = src[(size_t)(sn + ih++) * stride];
        else x[i - u0] = src[(size_t)(il++) * stride];
    }
}

typedef void (*lift_fn)(int32_t*, int, int);

/* dwt.cpp:420-506: finest level first; per level vertical (columns) then horizontal (rows). */
static void dwt_fwd(int32_t* a, int x0, int y0, int x1, int y1, int numres, lift_fn lift) {
    int w = x1 - x0;
    int maxlen = (x1 - x0) > (y1 - y0) ? (x1 - x0) : (y1 - y0);
    int32_t* line = (int32_t*)malloc(sizeof(int32_t) * (size_t)(maxlen + 1));
    for (int lvl = 0; lvl < numres - 1; ++lvl) {
        int rx0 = ceildivpow2(x0, lvl), rx1 = ceildivpow2(x1, lvl);
        int ry0 = ceildivpow2(y0, lvl), ry1 = ceildivpow2(y1, lvl);
        int rw = rx1 - rx0, rh = ry1 - ry0;
        for (int j = 0; j < rw; ++j) {
            for (int k = 0; k < rh; ++k) line[k] = a[(size_t)k * w + j];
            lift(line, ry0, ry1);
            deinterleave(line, a + j, w, ry0, ry1);
        }
        for (int j = 0; j < rh; ++j) {
            memcpy(line, a + (size_t)j * w, sizeof(int32_t) * (size_t)rw);
            lift(line, rx0, rx1);
            deinterleave(line, a + (size_t)j * w, 1, rx0, rx1);
        }
    }
    free(line);
}
/* dwt.cpp:615-704: coarsest first; per level horizontal then vertical. */
static void dwt_inv(int32_t* a, int x0, int y0, int x1, int y1, int numres, lift_fn lift) {
    int w = x1 - x0;
    int maxlen = (x1 - x0) > (y1 - y0) ? (x1 - x0) : (y1 - y0);
    int32_t* line = (int32_t*)malloc(sizeof(int32_t) * (size_t)(maxlen + 1));
    for (int lvl = numres - 2; lvl >= 0; --lvl) {
        int rx0 = ceildivpow2(x0, lvl), rx1 = ceildivpow2(x1, lvl);
        int ry0 = ceildivpow2(y0, lvl), ry1 = ceildivpow2(y1, lvl);
        int rw = rx1 - rx0, rh = ry1 - ry0;
        for (int j = 0; j < rh; ++j) {
            interleave(a + (size_t)j * w, 1, line, rx0, rx1);
            lift(line, rx0, rx1);
            memcpy(a + (size_t)j * w, line, sizeof(int32_t) * (size_t)rw);
        }
        for (int j = 0; j < rw; ++j) {
            interleave(a + j, w, line, ry0, ry1);
            lift(line, ry0, ry1);
            for (int k = 0; k < rh; ++k) a[(size_t)k * w + j] = line[k];
        }
    }
    free(line);
}
void orc_dwt53_fwd(int32_t* a, int x0, int y0, int x1, int y1, int numres) { dwt_fwd(a, x0, y0, x1, y1, numres, lift53_fwd); }
void orc_dwt53_inv(int32_t* a, int x0, int y0, int x1, int y1, int numres) { dwt_inv(a, x0, y0, x1, y1, numres, lift53_inv); }
void orc_dwt97_fwd(int32_t* a, int x0, int y0, int x1, int y1, int numres) { dwt_fwd(a, x0, y0, x1, y1, numres, lift97_fwd); }
void orc_dwt97_inv(float* a, int x0, int y0, int x1, int y1, int numres) {
    /* floats are moved as 32-bit words by the shared driver */
    dwt_inv((int32_t*)a, x0, y0, x1, y1, numres, (lift_fn)(void*)lift97_inv);
}

/* ------------------------------------------------------------------------------------------ */
/* Quantiser: T1Encoder.cpp:103-174                                                            */
uint32_t orc_quantize_block(const int32_t* coeff, int stride, int w, int h, int qmfbid,
                            int32_t bandconst, uint32_t* out) {
    uint32_t max = 0;
    for (int j = 0; j < h; ++j)
        for (int i = 0; i < w; ++i) {
            int32_t v = coeff[(size_t)j * stride + i];
            v = qmfbid == 1 ? v * 64 : fix_mul_t1(v, bandconst);
            uint32_t mag = (uint32_t)(v < 0 ? -v : v);
            if (mag > max) max = mag;
            out[j * w + i] = mag | ((uint32_t)(v < 0) << 31);
        }
    return max;
}

/* ------------------------------------------------------------------------------------------ */
/* MQ coder: mqc.cpp:120-215 (state table = ISO 15444-1 Table C.2)                             */
static const uint16_t MQ_QE[47] = {
    0x5601, 0x3401, 0x1801, 0x0AC1, 0x0521, 0x0221, 0x5601, 0x5401, 0x4801, 0x3801, 0x3001, 0x2401,
    0x1C01, 0x1601, 0x5601, 0x5401, 0x5101, 0x4801, 0x3801, 0x3401, 0x3001, 0x2801, 0x2401, 0x2201,
    0x1C01, 0x1801, 0x1601, 0x1401, 0x1201, 0x1101, 0x0AC1, 0x09C1, 0x08A1, 0x0521, 0x0441, 0x02A1,
    0x0221, 0x0141, 0x0111, 0x0085, 0x0049, 0x0025, 0x0015, 0x0009, 0x0005, 0x0001, 0x5601 };
static const uint8_t MQ_NMPS[47] = {
    1, 2, 3, 4, 5, 38, 7, 8, 9, 10, 11, 12, 13, 29, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27,
    28, 29, 30, 31, 32, 33, 34, 35, 36, 37, 38, 39, 40, 41, 42, 43, 44, 45, 45, 46 };
static const uint8_t MQ_NLPS[47] = {
    1, 6, 9, 12, 29, 33, 6, 14, 14, 14, 17, 18, 20, 21, 14, 14, 15, 16, 17, 18, 19, 19, 20, 21, 22, 23,
    24, 25, 26, 27, 28, 29, 30, 31, 32, 33, 34, 35, 36, 37, 38, 39, 40, 41, 42, 43, 46 };
static const uint8_t MQ_SW[47] = {
    1, 0, 0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0,
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0 };
void orc_mq_table(uint16_t* qe, uint8_t* nmps, uint8_t* nlps, uint8_t* sw) {
    memcpy(qe, MQ_QE, sizeof(MQ_QE)); memcpy(nmps, MQ_NMPS, 47); memcpy(nlps, MQ_NLPS, 47); memcpy(sw, MQ_SW, 47);
}

enum { CTX_ZC = 0, CTX_SC = 9, CTX_MAG = 14, CTX_AGG = 17, CTX_UNI = 18, NCTX = 19 };

typedef struct {
    uint32_t a, c, ct;
    uint8_t *bp, *start;
    uint8_t st[NCTX];   /* state index */
    uint8_t mps[NCTX];
} mqenc_t;

static void mq_reset(uint8_t* st, uint8_t* mps) {          /* t1_opt.cpp:686-690 */
    memset(st, 0, NCTX); memset(mps, 0, NCTX);
    st[CTX_UNI] = 46; st[CTX_AGG] = 3; st[CTX_ZC] = 4;
}
static void mqe_init(mqenc_t* q, uint8_t* bp) {             /* mqc.cpp:406-420 */
    mq_reset(q->st, q->mps);
    q->a = 0x8000; q->c = 0; q->bp = bp - 1; q->ct = 12; q->start = bp;
}
static void mqe_byteout(mqenc_t* q) {                       /* mqc.cpp:223-257 */
    if (q->bp < q->start) {
        q->bp++; *q->bp = (uint8_t)(q->c >> 19); q->c &= 0x7ffff; q->ct = 8;
    } else if (*q->bp == 0xff) {
        q->bp++; *q->bp = (uint8_t)(q->c >> 20); q->c &= 0xfffff; q->ct = 7;
    } else if ((q->c & 0x8000000) == 0) {
        q->bp++; *q->bp = (uint8_t)(q->c >> 19); q->c &= 0x7ffff; q->ct = 8;
    } else {
        (*q->bp)++;
        if (*q->bp == 0xff) {
            q->c &= 0x7ffffff;
            q->bp++; *q->bp = (uint8_t)(q->c >> 20); q->c &= 0xfffff; q->ct = 7;
        } else {
            q->bp++; *q->bp = (uint8_t)(q->c >> 19); q->c &= 0x7ffff; q->ct = 8;
        }
    }
}
static void mqe_renorm(mqenc_t* q) {                        /* mqc.cpp:259-269 */
    do {
        q->a <<= 1; q->c <<= 1; q->ct--;
        if (q->ct == 0) mqe_byteout(q);
    } while ((q->a & 0x8000) == 0);
}
/* optional trace of the (context, decision) sequence: what the CUDA context kernel hands to its MQ kernels */
static uint8_t* g_trace = NULL; static int g_trace_cap = 0, g_trace_n = 0;
static void mqe_encode(mqenc_t* q, int ctx, uint32_t d) {   /* mqc.cpp:271-297, 428-439 */
    if (g_trace) { if (g_trace_n < g_trace_cap) g_trace[g_trace_n] = (uint8_t)((ctx << 1) | d); ++g_trace_n; }
    uint32_t s = q->st[ctx], qe = MQ_QE[s];
    q->a -= qe;
    if (q->mps[ctx] == d) {
        if ((q->a & 0x8000) == 0) {
            if (q->a < qe) q->a = qe; else q->c += qe;
            q->st[ctx] = MQ_NMPS[s];
            mqe_renorm(q);
        } else q->c += qe;
    } else {
        if (q->a < qe) q->c += qe; else q->a = qe;
        if (MQ_SW[s]) q->mps[ctx] ^= 1;
        q->st[ctx] = MQ_NLPS[s];
        mqe_renorm(q);
    }
}
static int mqe_numbytes(const mqenc_t* q) { long d = q->bp - q->start; return d < 0 ? 0 : (int)d; }
static void mqe_flush(mqenc_t* q) {                         /* mqc.cpp:299-306, 441-452 */
    uint32_t tempc = q->c + q->a;
    q->c |= 0xffff;
    if (q->c >= tempc) q->c -= 0x8000;
    q->c <<= q->ct; mqe_byteout(q);
    q->c <<= q->ct; mqe_byteout(q);
    if (*q->bp != 0xff) q->bp++;
}

/* ------------------------------------------------------------------------------------------ */
/* Context rules, ISO 15444-1 Tables D.1-D.3 (t1_luts.h tables are these rules tabulated)      */
int orc_zc_context(int orient, int h, int v, int d) {
    if (orient == 1) { int t = h; h = v; v = t; }
    if (orient == 3) {
        int hv = h + v;
        if (d >= 3) return 8;
        if (d == 2) return hv >= 1 ? 7 : 6;
        if (d == 1) return hv >= 2 ? 5 : (hv == 1 ? 4 : 3);
        return hv >= 2 ? 2 : (hv == 1 ? 1 : 0);
    }
    if (h == 2) return 8;
    if (h == 1) return v >= 1 ? 7 : (d >= 1 ? 6 : 5);
    if (v == 2) return 4;
    if (v == 1) return 3;
    return d >= 2 ? 2 : (d == 1 ? 1 : 0);
}
static void sc_context(int hc, int vc, int* ctx, int* xorbit) {
    if (hc > 1) hc = 1;
    if (hc < -1) hc = -1;
    if (vc > 1) vc = 1;
    if (vc < -1) vc = -1;
    *xorbit = 0;
    if (hc < 0 || (hc == 0 && vc < 0)) { *xorbit = 1; hc = -hc; vc = -vc; }
    if (hc == 1) *ctx = CTX_SC + (vc == 1 ? 4 : (vc == 0 ? 3 : 2));
    else *ctx = CTX_SC + (vc == 1 ? 1 : 0);
}

/* nmsedec tables: t1_generate_luts.cpp:434-456 ; lookups t1.cpp:298-314 */
static int16_t nm_sig[128], nm_sig0[128], nm_ref[128], nm_ref0[128];
static int nm_ready = 0;
static int imax(int a, int b) { return a > b ? a : b; }
static void nm_init(void) {
    if (nm_ready) return;
    for (int i = 0; i < 128; ++i) {
        double t = i / 64.0, u = t, v = t - 1.5;
        nm_sig[i] = (int16_t)imax(0, (int)(floor((u * u - v * v) * 64.0 + 0.5) / 64.0 * 8192.0));
        nm_sig0[i] = (int16_t)imax(0, (int)(floor((u * u) * 64.0 + 0.5) / 64.0 * 8192.0));
        u = t - 1.0;
        v = (i & 64) ? t - 1.5 : t - 0.5;
        nm_ref[i] = (int16_t)imax(0, (int)(floor((u * u - v * v) * 64.0 + 0.5) / 64.0 * 8192.0));
        nm_ref0[i] = (int16_t)imax(0, (int)(floor((u * u) * 64.0 + 0.5) / 64.0 * 8192.0));
    }
    nm_ready = 1;
}
static int nmsedec_sig(uint32_t x, int bpno) { return bpno > 0 ? nm_sig[(x >> bpno) & 127] : nm_sig0[x & 127]; }
static int nmsedec_ref(uint32_t x, int bpno) { return bpno > 0 ? nm_ref[(x >> bpno) & 127] : nm_ref0[x & 127]; }

static const double DWT_NORMS[4][10] = {     /* dwt.cpp:193-198 */
    {1.000, 1.500, 2.750, 5.375, 10.68, 21.34, 42.67, 85.33, 170.7, 341.3},
    {1.038, 1.592, 2.919, 5.703, 11.33, 22.64, 45.25, 90.48, 180.9},
    {1.038, 1.592, 2.919, 5.703, 11.33, 22.64, 45.25, 90.48, 180.9},
    {.7186, .9218, 1.586, 3.043, 6.019, 12.01, 24.00, 47.97, 95.93}};
static const double DWT_NORMS_REAL[4][10] = { /* dwt.cpp:203-208 */
    {1.000, 1.965, 4.177, 8.403, 16.90, 33.84, 67.69, 135.3, 270.6, 540.9},
    {2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0},
    {2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0},
    {2.080, 3.865, 8.307, 17.18, 34.71, 69.59, 139.3, 278.6, 557.2}};
double orc_dwt_norm(int level, int orient, int real) { return real ? DWT_NORMS_REAL[orient][level] : DWT_NORMS[orient][level]; }
double orc_mct_norm(int compno, int real) {   /* mct.cpp:73,78 */
    static const double n0[3] = {1.732, .8292, .8292}, n1[3] = {1.732, 1.805, 1.573};
    return real ? n1[compno] : n0[compno];
}
double orc_wmsedec(int nmsedec, double mct_norm, double dwt_norm, double stepsize, int bpno) { /* t1.cpp:1114-1143 */
    double w = mct_norm * dwt_norm * stepsize * (1 << bpno);
    w *= w * nmsedec / 8192.0;
    return w;
}

/* ------------------------------------------------------------------------------------------ */
/* Tier-1 state, one byte-plane per flag with a one-sample border                              */
typedef struct {
    int w, h, sw;            /* sw = w + 2 */
    uint8_t *sig, *neg, *vis, *ref;
} t1state_t;
static int t1s_init(t1state_t* s, int w, int h) {
    s->w = w; s->h = h; s->sw = w + 2;
    size_t n = (size_t)(w + 2) * (h + 2);
    s->sig = (uint8_t*)calloc(4, n);
    if (!s->sig) return -1;
    s->neg = s->sig + n; s->vis = s->neg + n; s->ref = s->vis + n;
    return 0;
}
static void t1s_free(t1state_t* s) { free(s->sig); }
#define IDX(s, x, y) ((size_t)((y) + 1) * (s)->sw + (x) + 1)

static int zc_ctx_at(const t1state_t* s, int x, int y, int orient) {
    size_t p = IDX(s, x, y); int sw = s->sw;
    int h = s->sig[p - 1] + s->sig[p + 1];
    int v = s->sig[p - sw] + s->sig[p + sw];
    int d = s->sig[p - sw - 1] + s->sig[p - sw + 1] + s->sig[p + sw - 1] + s->sig[p + sw + 1];
    return CTX_ZC + orc_zc_context(orient, h, v, d);
}
static int any_sig_neighbour(const t1state_t* s, int x, int y) {
    size_t p = IDX(s, x, y); int sw = s->sw;
    return s->sig[p - 1] | s->sig[p + 1] | s->sig[p - sw] | s->sig[p + sw] | s->sig[p - sw - 1] |
           s->sig[p - sw + 1] | s->sig[p + sw - 1] | s->sig[p + sw + 1];
}
static void sc_ctx_at(const t1state_t* s, int x, int y, int* ctx, int* xorbit) {
    size_t p = IDX(s, x, y); int sw = s->sw;
#define CONTRIB(q) (s->sig[q] ? (s->neg[q] ? -1 : 1) : 0)
    int hc = CONTRIB(p - 1) + CONTRIB(p + 1);
    int vc = CONTRIB(p - sw) + CONTRIB(p + sw);
#undef CONTRIB
    sc_context(hc, vc, ctx, xorbit);
}

/* t1_opt.cpp:330-401 */
static void enc_sigpass(t1state_t* s, mqenc_t* q, const uint32_t* d, int bpno, int orient, int* nmsedec) {
    int one = bpno + 6;
    *nmsedec = 0;
    for (int k = 0; k < s->h; k += 4)
        for (int x = 0; x < s->w; ++x)
            for (int y = k; y < k + 4 && y < s->h; ++y) {
                size_t p = IDX(s, x, y);
                if (s->sig[p] || s->vis[p] || !any_sig_neighbour(s, x, y)) continue;
                uint32_t v = d[y * s->w + x];
                uint32_t bit = (v >> one) & 1;
                mqe_encode(q, zc_ctx_at(s, x, y, orient), bit);
                if (bit) {
                    int ctx, xb;
                    *nmsedec += nmsedec_sig(v, bpno);
                    sc_ctx_at(s, x, y, &ctx, &xb);
                    mqe_encode(q, ctx, (v >> 31) ^ (uint32_t)xb);
                    s->sig[p] = 1; s->neg[p] = (uint8_t)(v >> 31);
                }
                s->vis[p] = 1;
            }
}
/* t1_opt.cpp:403-467 */
static void enc_refpass(t1state_t* s, mqenc_t* q, const uint32_t* d, int bpno, int* nmsedec) {
    int one = bpno + 6;
    *nmsedec = 0;
    for (int k = 0; k < s->h; k += 4)
        for (int x = 0; x < s->w; ++x)
            for (int y = k; y < k + 4 && y < s->h; ++y) {
                size_t p = IDX(s, x, y);
                if (!s->sig[p] || s->vis[p]) continue;
                uint32_t v = d[y * s->w + x];
                *nmsedec += nmsedec_ref(v, bpno);
                int ctx = s->ref[p] ? CTX_MAG + 2 : (any_sig_neighbour(s, x, y) ? CTX_MAG + 1 : CTX_MAG);
                mqe_encode(q, ctx, (v >> one) & 1);
                s->ref[p] = 1;
            }
}
/* t1_opt.cpp:469-576 */
static void enc_clnpass(t1state_t* s, mqenc_t* q, const uint32_t* d, int bpno, int orient, int* nmsedec) {
    int one = bpno + 6;
    *nmsedec = 0;
    for (int k = 0; k < s->h; k += 4)
        for (int x = 0; x < s->w; ++x) {
            int y = k, partial = 0;
            int agg = (k + 4 <= s->h);
            for (int j = k; agg && j < k + 4; ++j) {
                size_t p = IDX(s, x, j);
                if (s->sig[p] || s->vis[p] || any_sig_neighbour(s, x, j)) agg = 0;
            }
            if (agg) {
                int runlen = 0;
                for (; runlen < 4; ++runlen)
                    if ((d[(k + runlen) * s->w + x] >> one) & 1) break;
                mqe_encode(q, CTX_AGG, runlen != 4);
                if (runlen == 4) continue;
                mqe_encode(q, CTX_UNI, (uint32_t)runlen >> 1);
                mqe_encode(q, CTX_UNI, (uint32_t)runlen & 1);
                y = k + runlen; partial = 1;
            }
            for (; y < k + 4 && y < s->h; ++y) {
                size_t p = IDX(s, x, y);
                uint32_t v = d[y * s->w + x];
                if (partial || (!s->sig[p] && !s->vis[p])) {
                    uint32_t bit = 1;
                    if (!partial) {
                        bit = (v >> one) & 1;
                        mqe_encode(q, zc_ctx_at(s, x, y, orient), bit);
                    }
                    partial = 0;
                    if (bit) {
                        int ctx, xb;
                        *nmsedec += nmsedec_sig(v, bpno);
                        sc_ctx_at(s, x, y, &ctx, &xb);
                        mqe_encode(q, ctx, (v >> 31) ^ (uint32_t)xb);
                        s->sig[p] = 1; s->neg[p] = (uint8_t)(v >> 31);
                    }
                }
                s->vis[p] = 0;
            }
            /* samples skipped by the run still get their visit flag cleared (t1_opt.cpp:504-515) */
            for (int j = k; j < k + 4 && j < s->h; ++j) s->vis[IDX(s, x, j)] = 0;
        }
}

static int floorlog2(uint32_t v) { int l = 0; while (v > 1) { v >>= 1; ++l; } return l; }

int orc_t1_encode_cblk(const uint32_t* smag, int w, int h, int orient, uint32_t maxmag,
                       uint8_t* out, int cap, int* numbps, int* npasses, int* rates, int* nmsedec) {
    nm_init();
    t1state_t s;
    if (t1s_init(&s, w, h)) return -1;
    int worst = w * h * 4 + 16;
    uint8_t* buf = (uint8_t*)calloc(1, (size_t)worst + 2);
    mqenc_t q;
    mqe_init(&q, buf + 1);
    int nb = maxmag ? floorlog2(maxmag) + 1 - 6 : 0;     /* t1_opt.cpp:680 */
    if (nb < 0) nb = 0;
    *numbps = nb;
    int bpno = nb - 1, passtype = 2, passno = 0;
    for (; bpno >= 0; ++passno) {
        int nm = 0;
        if (passtype == 0) enc_sigpass(&s, &q, smag, bpno, orient, &nm);
        else if (passtype == 1) enc_refpass(&s, &q, smag, bpno, &nm);
        else enc_clnpass(&s, &q, smag, bpno, orient, &nm);
        nmsedec[passno] = nm;
        rates[passno] = mqe_numbytes(&q) + 3;
        if (++passtype == 3) { passtype = 0; bpno--; }
    }
    mqe_flush(&q);
    *npasses = passno;
    int total = mqe_numbytes(&q);
    const uint8_t* data = buf + 1;
    for (int i = 0; i < passno; ++i) {                   /* t1_opt.cpp:727-736 */
        if (rates[i] > total) rates[i] = total;
        if (rates[i] > 1 && data[rates[i] - 1] == 0xFF) rates[i]--;
    }
    int used = passno ? rates[passno - 1] : 0;
    int rc = used;
    if (used > cap) rc = -1; else memcpy(out, data, (size_t)used);
    free(buf); t1s_free(&s);
    return rc;
}

/* The GENERIC encoder loop (t1.cpp:1537-1673) for the code-block styles that only touch the arithmetic coder: RESET (0x02,
 * contexts back to their initial states after every pass) and TERMALL (0x04, the coder is flushed and restarted after every
 * pass: every pass is a codeword segment of its own).  Context modelling is that of cblksty == 0.  Differences from the
 * optimised loop above: the rate of a pass is numbytes + 3, or + 1 when the pass was terminated, and only when numbytes > 0;
 * the restart steps back over the last byte (mqc.cpp:521-532) and the final flush comes on top of the last pass's own.
 * Optional outputs: the symbol trace ((ctx << 1) | d per decision) and the number of symbols of every pass. */
int orc_t1_encode_cblk_sty(const uint32_t* smag, int w, int h, int orient, uint32_t maxmag, int cblksty,
                           uint8_t* out, int cap, int* numbps, int* npasses, int* rates, int* nmsedec,
                           uint8_t* trace, int trace_cap, uint32_t* pass_nsym) {
    if (cblksty & ~0x136) return -2;                       /* RESET 0x02, TERMALL 0x04, PTERM 0x10, SEGSYM 0x20 (+ 0x100, see below) */
    nm_init();
    t1state_t s;
    if (t1s_init(&s, w, h)) return -1;
    int worst = w * h * 8 + 64;
    uint8_t* buf = (uint8_t*)calloc(1, (size_t)worst + 2);
    mqenc_t q;
    mqe_init(&q, buf + 1);
    int nb = maxmag ? floorlog2(maxmag) + 1 - 6 : 0;
    if (nb < 0) nb = 0;
    *numbps = nb;
    g_trace = trace; g_trace_cap = trace_cap; g_trace_n = 0;
    int bpno = nb - 1, passtype = 2, passno = 0;
    for (; bpno >= 0; ++passno) {
        int nm = 0, before = g_trace_n, correction = 3;
        if (passtype == 0) enc_sigpass(&s, &q, smag, bpno, orient, &nm);
        else if (passtype == 1) enc_refpass(&s, &q, smag, bpno, &nm);
        else {
            enc_clnpass(&s, &q, smag, bpno, orient, &nm);
            if (cblksty & 0x20) {                               /* SEGSYM: 1, 0, 1, 0 in the UNI context after every clean-up pass */
                uint8_t* keep = g_trace; g_trace = NULL;        /* (t1.cpp:1598-1600, mqc.cpp:550-558); not part of the modelling trace */
                for (int i = 1; i < 5; ++i) mqe_encode(&q, CTX_UNI, (uint32_t)(i % 2));
                g_trace = keep;
            }
        }
        nmsedec[passno] = nm;
        if (pass_nsym) pass_nsym[passno] = (uint32_t)(g_trace_n - before);
        if (cblksty & 0x04) {                                   /* t1.cpp:1618-1647 */
            mqe_flush(&q);
            correction = 1;
            q.a = 0x8000; q.c = 0; q.ct = 12; q.bp--;           /* opj_mqc_restart_init_enc, mqc.cpp:521-532 */
            if (q.bp >= q.start - 1 && *q.bp == 0xff) q.ct = 13;
        }
        if (++passtype == 3) { passtype = 0; bpno--; }
        rates[passno] = mqe_numbytes(&q);
        if (rates[passno] > 0) rates[passno] += correction;
        /* REFERENCE DEFECT kept only on request (cblksty bit 0x100): after a termination, numbytes == 0 also means "one byte,
         * at index 0", and the reference then reports length 0 for the pass; its own decoder cannot read such a block.  The
         * correct length of a terminated pass is everything up to and including the byte the restart stepped back to. */
        else if ((cblksty & 0x04) && !(cblksty & 0x100)) rates[passno] = (int)(q.bp - q.start) + 1;
        if (cblksty & 0x02) mq_reset(q.st, q.mps);              /* t1.cpp:1654-1656 */
    }
    g_trace = NULL;
    if (cblksty & 0x10) {                                       /* PTERM: opj_mqc_erterm_enc, mqc.cpp:534-548 */
        int k = (int)(11 - q.ct + 1);
        while (k > 0) { q.c <<= q.ct; q.ct = 0; mqe_byteout(&q); k -= (int)q.ct; }
        if (*q.bp != 0xff) mqe_byteout(&q);
    } else mqe_flush(&q);                                       /* t1.cpp:1659-1663 */
    *npasses = passno;
    int total = mqe_numbytes(&q);
    const uint8_t* data = buf + 1;
    for (int i = 0; i < passno; ++i) {                          /* t1.cpp:1667-1677 */
        if (rates[i] > total) rates[i] = total;
        if (rates[i] > 1 && data[rates[i] - 1] == 0xFF) rates[i]--;
    }
    int used = passno ? rates[passno - 1] : 0;
    int rc = used;
    if (used > cap) rc = -1; else memcpy(out, data, (size_t)used);
    free(buf); t1s_free(&s);
    return rc;
}

/* ------------------------------------------------------------------------------------------ */
/* MQ decoder: mqc.cpp:336-378, 560-595                                                        */
typedef struct {
    uint32_t a, c, ct;
    const uint8_t *bp, *end;
    uint8_t st[NCTX], mps[NCTX];
} mqdec_t;
static void mqd_bytein(mqdec_t* q) {
    if (q->bp != q->end) {
        uint32_t c = (q->bp + 1 != q->end) ? q->bp[1] : 0xff;
        if (*q->bp == 0xff) {
            if (c > 0x8f) { q->c += 0xff00; q->ct = 8; }
            else { q->bp++; q->c += c << 9; q->ct = 7; }
        } else { q->bp++; q->c += c << 8; q->ct = 8; }
    } else { q->c += 0xff00; q->ct = 8; }
}
static void mqd_init(mqdec_t* q, const uint8_t* bp, int len) {
    mq_reset(q->st, q->mps);
    q->bp = bp; q->end = bp + len;
    q->c = len == 0 ? (0xffu << 16) : ((uint32_t)*bp << 16);
    mqd_bytein(q);
    q->c <<= 7; q->ct -= 7; q->a = 0x8000;
}
static void mqd_renorm(mqdec_t* q) {
    do {
        if (q->ct == 0) mqd_bytein(q);
        q->a <<= 1; q->c <<= 1; q->ct--;
    } while (q->a < 0x8000);
}
static int mqd_decode(mqdec_t* q, int ctx) {
    uint32_t s = q->st[ctx], qe = MQ_QE[s];
    int d;
    q->a -= qe;
    if ((q->c >> 16) < qe) {
        if (q->a < qe) { d = q->mps[ctx]; q->st[ctx] = MQ_NMPS[s]; }
        else { d = 1 - q->mps[ctx]; if (MQ_SW[s]) q->mps[ctx] ^= 1; q->st[ctx] = MQ_NLPS[s]; }
        q->a = qe;
        mqd_renorm(q);
    } else {
        q->c -= qe << 16;
        if ((q->a & 0x8000) == 0) {
            if (q->a < qe) { d = 1 - q->mps[ctx]; if (MQ_SW[s]) q->mps[ctx] ^= 1; q->st[ctx] = MQ_NLPS[s]; }
            else { d = q->mps[ctx]; q->st[ctx] = MQ_NMPS[s]; }
            mqd_renorm(q);
        } else d = q->mps[ctx];
    }
    return d;
}

/* t1.cpp:419-445,586-614 (sig), 645-665,732-757 (ref), 845-1112 (cleanup), for cblksty == 0 */
int orc_t1_decode_cblk(const uint8_t* data, int len, int w, int h, int orient, int numbps,
                       int npasses, int32_t* out) {
    memset(out, 0, sizeof(int32_t) * (size_t)w * h);
    if (npasses == 0 || len == 0) return 0;
    t1state_t s;
    if (t1s_init(&s, w, h)) return -1;
    mqdec_t q;
    mqd_init(&q, data, len);
    int bp1 = numbps, passtype = 2;          /* bpno_plus_one */
    for (int passno = 0; passno < npasses && bp1 >= 1; ++passno) {
        int32_t one = 1 << bp1, half = one >> 1, oneplushalf = one | half;
        if (passtype == 0) {
            for (int k = 0; k < h; k += 4) for (int x = 0; x < w; ++x) for (int y = k; y < k + 4 && y < h; ++y) {
                size_t p = IDX(&s, x, y);
                if (s.sig[p] || s.vis[p] || !any_sig_neighbour(&s, x, y)) continue;
                if (mqd_decode(&q, zc_ctx_at(&s, x, y, orient))) {
                    int ctx, xb; sc_ctx_at(&s, x, y, &ctx, &xb);
                    int neg = mqd_decode(&q, ctx) ^ xb;
                    out[y * w + x] = neg ? -oneplushalf : oneplushalf;
                    s.sig[p] = 1; s.neg[p] = (uint8_t)neg;
                }
                s.vis[p] = 1;
            }
        } else if (passtype == 1) {
            int32_t poshalf = one >> 1, neghalf = -poshalf;   /* bpno_plus_one >= 1 always */
            for (int k = 0; k < h; k += 4) for (int x = 0; x < w; ++x) for (int y = k; y < k + 4 && y < h; ++y) {
                size_t p = IDX(&s, x, y);
                if (!s.sig[p] || s.vis[p]) continue;
                int ctx = s.ref[p] ? CTX_MAG + 2 : (any_sig_neighbour(&s, x, y) ? CTX_MAG + 1 : CTX_MAG);
                int32_t t = mqd_decode(&q, ctx) ? poshalf : neghalf;
                out[y * w + x] += out[y * w + x] < 0 ? -t : t;
                s.ref[p] = 1;
            }
        } else {
            for (int k = 0; k < h; k += 4) for (int x = 0; x < w; ++x) {
                int y = k, partial = 0;
                int agg = (k + 4 <= h);
                for (int j = k; agg && j < k + 4; ++j) {
                    size_t p = IDX(&s, x, j);
                    if (s.sig[p] || s.vis[p] || any_sig_neighbour(&s, x, j)) agg = 0;
                }
                if (agg) {
                    if (!mqd_decode(&q, CTX_AGG)) continue;
                    int runlen = mqd_decode(&q, CTX_UNI);
                    runlen = (runlen << 1) | mqd_decode(&q, CTX_UNI);
                    y = k + runlen; partial = 1;
                }
                for (; y < k + 4 && y < h; ++y) {
                    size_t p = IDX(&s, x, y);
                    if (partial || (!s.sig[p] && !s.vis[p])) {
                        int bit = 1;
                        if (!partial) bit = mqd_decode(&q, zc_ctx_at(&s, x, y, orient));
                        partial = 0;
                        if (bit) {
                            int ctx, xb; sc_ctx_at(&s, x, y, &ctx, &xb);
                            int neg = mqd_decode(&q, ctx) ^ xb;
                            out[y * w + x] = neg ? -oneplushalf : oneplushalf;
                            s.sig[p] = 1; s.neg[p] = (uint8_t)neg;
                        }
                    }
                }
                for (int j = k; j < k + 4 && j < h; ++j) s.vis[IDX(&s, x, j)] = 0;
            }
        }
        if (++passtype == 3) { passtype = 0; bp1--; }
    }
    t1s_free(&s);
    return 0;
}
