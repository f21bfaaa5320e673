// Tier-1 decoder core, shared between the CUDA kernel (t1_dec.cu) and a host-compiled test harness
// (tests/native/t1dec_host.cpp) that runs the very same block routine with one "lane".  See t1_dec.cu.
#pragma once
#include "device.h"
#include "t1_tables.cuh"
#ifdef __CUDACC__
#define J2K_HD __host__ __device__
#else
#define J2K_HD
#endif
#if defined(__CUDA_ARCH__)
#define J2K_FFSLL(v) __ffsll((long long)(v))
#define J2K_SYNCWARP() __syncwarp()
#define J2K_FMUL(a, b) __fmul_rn((a), (b))
#else
#define J2K_FFSLL(v) __builtin_ffsll((long long)(v))
#define J2K_SYNCWARP() ((void)0)
#define J2K_FMUL(a, b) ((a) * (b))
#endif

namespace j2k {
namespace t1dec {

typedef unsigned long long u64;

struct DRows {                // index = row + 1
    u64 sig[66], sgn[66], vis[66], ref[66], bit[66], last[66];
};

// MQ decoder (mqc.cpp:336-378, 560-595).  The byte stream is consumed through two 8-byte words held in
// registers; the word after them is fetched when the window slides, i.e. dozens of decisions before
// any of its bytes is needed.  The segment must start 8-byte aligned, be followed by two 0xFF bytes
// (which reproduces the reference's behaviour at and past the end of the data, mqc.cpp:336-365) and
// by at least 24 more readable bytes.
struct MqDec {
    uint32_t a, c, ct;
    const u64* wp;            // next word to fetch
    u64 w0, w1;               // bytes [8k, 8k+8) and [8k+8, 8k+16) of the segment
    uint32_t off;             // position of the current byte inside w0
    J2K_HD inline void advance() {
        if (++off == 8) { off = 0; w0 = w1; w1 = *wp++; }
    }
    J2K_HD inline void bytein() {
        const uint32_t cur = (uint32_t)(w0 >> (8 * off)) & 0xffu;
        const uint32_t nxt = off < 7 ? (uint32_t)(w0 >> (8 * off + 8)) & 0xffu : (uint32_t)w1 & 0xffu;
        if (cur == 0xffu) {
            if (nxt > 0x8fu) { c += 0xff00u; ct = 8; }
            else { advance(); c += nxt << 9; ct = 7; }
        } else { advance(); c += nxt << 8; ct = 8; }
    }
    J2K_HD void init(const uint8_t* p) {                  // mqc.cpp:560-575
        wp = reinterpret_cast<const u64*>(p);
        w0 = *wp++; w1 = *wp++;
        off = 0;
        c = ((uint32_t)w0 & 0xffu) << 16;
        bytein();
        c <<= 7; ct -= 7; a = 0x8000u;
    }
    J2K_HD inline void renorm() {
        do {
            if (ct == 0) bytein();
            a <<= 1; c <<= 1; --ct;
        } while (a < 0x8000u);
    }
    // st: per-context state (index*2 + mps), tab: T1Tables::mq
    J2K_HD inline uint32_t decode(uint8_t* st, const uint32_t* tab, uint32_t ctx) {   // mqc.cpp:577-595
        const uint32_t s = st[ctx], e = tab[s], qe = e & 0xffffu;
        uint32_t d = s & 1u;
        a -= qe;
        if ((c >> 16) < qe) {
            if (a < qe) st[ctx] = (uint8_t)((e >> 16) & 0xffu);
            else { d ^= 1u; st[ctx] = (uint8_t)(e >> 24); }
            a = qe;
            renorm();
        } else {
            c -= qe << 16;
            if ((a & 0x8000u) == 0) {
                if (a < qe) { d ^= 1u; st[ctx] = (uint8_t)(e >> 24); }
                else st[ctx] = (uint8_t)((e >> 16) & 0xffu);
                renorm();
            }
        }
        return d;
    }
};

J2K_HD inline uint32_t w3(u64 m, int x) {          // bits x-1, x, x+1 of a row
    return (uint32_t)(x ? (m >> (x - 1)) : (m << 1)) & 7u;
}

struct Stripe {               // rows q = 0..5 <-> y = 4*stripe - 1 + q
    u64 S[6], G[6];           // significance, sign
    u64 V[4], F[4], B[4], L[4];   // visited, refined, this plane's bits, coded in this plane
};

// 3-bit significance windows (columns x-1, x, x+1) of the six rows around the stripe, for the column
// being scanned; kept current as samples of that column become significant
struct Col { uint32_t s[6]; };

J2K_HD inline void col_load(const Stripe& st, int x, Col& c) {
#pragma unroll
    for (int q = 0; q < 6; ++q) c.s[q] = w3(st.S[q], x);
}
// zero-coding neighbourhood index: bits 0 NW, 1 N, 2 NE, 3 W, 4 E, 5 SW, 6 S, 7 SE
J2K_HD inline uint32_t nbhood(const Col& c, int q) {
    return c.s[q - 1] | ((c.s[q] & 1u) << 3) | ((c.s[q] >> 2) << 4) | (c.s[q + 1] << 5);
}

J2K_HD inline void decode_sign(Stripe& s, Col& c, MqDec& mq, uint8_t* st, const uint32_t* tab, const uint8_t* sc, int q, int x) {
    const uint32_t mg = w3(s.G[q], x);
    const uint32_t gn = (w3(s.G[q - 1], x) >> 1) & 1u, gs = (w3(s.G[q + 1], x) >> 1) & 1u;
    const uint32_t sw = c.s[q] & 1u, se = c.s[q] >> 2, sn = (c.s[q - 1] >> 1) & 1u, ss = (c.s[q + 1] >> 1) & 1u;
    const uint32_t idx = sw | (se << 1) | (sn << 2) | (ss << 3) | ((sw & mg) << 4) | ((se & (mg >> 2)) << 5) | ((sn & gn) << 6) | ((ss & gs) << 7);
    const uint32_t e = sc[idx];
    const uint32_t neg = mq.decode(st, tab, CTX_SC + (e & 7u)) ^ (e >> 3);
    s.S[q] |= 1ull << x;
    s.G[q] |= (u64)neg << x;
    s.B[q - 1] |= 1ull << x;
    s.L[q - 1] |= 1ull << x;
    c.s[q] |= 2u;
}

// Decode one code-block.  Called by all `nlanes` lanes of a warp with identical arguments except `lane`;
// the decision loop runs redundantly on warp-uniform data, the lane index only partitions the
// column-parallel parts (fold, final pass).  R and st must be zero / reset on entry.
J2K_HD inline void decode_block(const T1DecBlock& blk, DRows& R, uint8_t* st, const uint8_t* zc_all, const uint8_t* sc,
                                const uint32_t* mqtab, const uint8_t* seg_arena, int raw_output, int lane, int nlanes) {
    const int w = blk.w, h = blk.h;
    int32_t* out = blk.coeff;
    const uint8_t* zc = zc_all + blk.orient * 256;
    const int nstripes = (h + 3) >> 2;
    const u64 colmask = w == 64 ? ~0ull : ((1ull << w) - 1);

    MqDec mq;
    mq.init(seg_arena + blk.data_off);

    int bp = blk.numbps - 1;              // 0-based bit-plane being decoded
    int passtype = 2;
    int lastplane = bp;
    bool firstfold = true;
    for (int passno = 0; passno < blk.npasses && bp >= 0; ++passno) {
        for (int sidx = 0; sidx < nstripes; ++sidx) {
            Stripe s;
            const int ib = 4 * sidx;       // smem index of row q = 0
#pragma unroll
            for (int q = 0; q < 6; ++q) { s.S[q] = R.sig[ib + q]; s.G[q] = R.sgn[ib + q]; }
            u64 valid[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                s.V[r] = R.vis[ib + 1 + r]; s.F[r] = R.ref[ib + 1 + r]; s.B[r] = R.bit[ib + 1 + r]; s.L[r] = R.last[ib + 1 + r];
                valid[r] = (ib + r < h) ? colmask : 0ull;
            }
            if (passtype == 0) {
                // significance propagation: only columns that have (or acquire, from the column just
                // scanned) a significant sample in their 3x6 window can hold candidates
                u64 active = 0;
#pragma unroll
                for (int q = 0; q < 6; ++q) active |= (s.S[q] << 1) | s.S[q] | (s.S[q] >> 1);
                active &= colmask;
                u64 donec = 0;
                while (u64 rem = active & ~donec) {
                    const int x = J2K_FFSLL(rem) - 1;
                    donec |= 1ull << x;
                    Col c;
                    col_load(s, x, c);
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const int q = r + 1;
                        if (!((valid[r] >> x) & 1ull) || (c.s[q] & 2u)) continue;
                        const uint32_t nb = nbhood(c, q);
                        if (!nb) continue;
                        if (mq.decode(st, mqtab, zc[nb])) {
                            decode_sign(s, c, mq, st, mqtab, sc, q, x);
                            active |= (2ull << x) & colmask;           // the next column may now hold candidates
                        }
                        s.V[r] |= 1ull << x;
                    }
                }
            } else if (passtype == 1) {
                // magnitude refinement: significant and not coded in this plane's first pass
                u64 memb[4], cols = 0;
#pragma unroll
                for (int r = 0; r < 4; ++r) { memb[r] = s.S[r + 1] & ~s.V[r] & valid[r]; cols |= memb[r]; }
                while (cols) {
                    const int x = J2K_FFSLL(cols) - 1;
                    cols &= cols - 1;
                    Col c;
                    col_load(s, x, c);
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const int q = r + 1;
                        if (!((memb[r] >> x) & 1ull)) continue;
                        const uint32_t ctx = ((s.F[r] >> x) & 1ull) ? CTX_MAG + 2 : (nbhood(c, q) ? CTX_MAG + 1 : CTX_MAG);
                        const uint32_t bit = mq.decode(st, mqtab, ctx);
                        s.B[r] |= (u64)bit << x;
                    }
                }
#pragma unroll
                for (int r = 0; r < 4; ++r) { s.L[r] |= memb[r]; s.F[r] |= memb[r]; }
            } else {
                // cleanup: everything not yet significant and not visited
                u64 memb[4], cols = 0;
#pragma unroll
                for (int r = 0; r < 4; ++r) { memb[r] = ~s.S[r + 1] & ~s.V[r] & valid[r]; cols |= memb[r]; }
                const u64 memall = (ib + 3 < h) ? (memb[0] & memb[1] & memb[2] & memb[3]) : 0ull;
                while (cols) {
                    const int x = J2K_FFSLL(cols) - 1;
                    cols &= cols - 1;
                    Col c;
                    col_load(s, x, c);
                    int rstart = 0;
                    bool partial = false;
                    if (((memall >> x) & 1ull) && (c.s[0] | c.s[1] | c.s[2] | c.s[3] | c.s[4] | c.s[5]) == 0) {
                        // run-length mode: the whole column and its neighbourhood are insignificant
                        if (!mq.decode(st, mqtab, CTX_AGG)) continue;
                        int runlen = (int)mq.decode(st, mqtab, CTX_UNI);
                        runlen = (runlen << 1) | (int)mq.decode(st, mqtab, CTX_UNI);
                        rstart = runlen;
                        partial = true;
                    }
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const int q = r + 1;
                        if (r < rstart || !((memb[r] >> x) & 1ull)) continue;
                        uint32_t bit = 1;
                        if (!(partial && r == rstart)) bit = mq.decode(st, mqtab, zc[nbhood(c, q)]);
                        if (bit) decode_sign(s, c, mq, st, mqtab, sc, q, x);
                    }
                }
#pragma unroll
                for (int r = 0; r < 4; ++r) s.V[r] = 0;
            }
            // write the stripe back (rows of the neighbouring stripes are unchanged)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int i = ib + 1 + r;
                R.sig[i] = s.S[r + 1]; R.sgn[i] = s.G[r + 1]; R.vis[i] = s.V[r]; R.ref[i] = s.F[r]; R.bit[i] = s.B[r]; R.last[i] = s.L[r];
            }
            J2K_SYNCWARP();
        }
        lastplane = bp;
        if (++passtype == 3 || passno + 1 == blk.npasses) {
            // fold this plane's decoded bits into the magnitudes kept in the coefficient plane
            const bool plane_done = passtype == 3;
            for (int y0 = 0; y0 < h; y0 += 8) {
                for (int x = lane; x < w; x += nlanes) {
                    int32_t v[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int y = y0 + k;
                        v[k] = 0;
                        if (!firstfold && y < h && R.bit[y + 1]) v[k] = out[(size_t)y * blk.pitch + x];
                    }
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int y = y0 + k;
                        if (y >= h) continue;
                        const u64 b = R.bit[y + 1];
                        if (!b && !firstfold) continue;
                        out[(size_t)y * blk.pitch + x] = v[k] | (int32_t)(((b >> x) & 1ull) << bp);
                    }
                }
            }
            firstfold = false;
            J2K_SYNCWARP();
            if (plane_done) {
                passtype = 0;
                if (passno + 1 < blk.npasses && bp > 0) {
                    for (int i = lane + 1; i <= h; i += nlanes) { R.bit[i] = 0; R.last[i] = 0; }
                }
                --bp;
                J2K_SYNCWARP();
            }
        }
    }
    // final pass: mid-point reconstruction, sign, dequantisation (T1Decoder.cpp:77-99)
    for (int y = 0; y < h; ++y) {
        const u64 sg = R.sig[y + 1], ng = R.sgn[y + 1], la = R.last[y + 1];
        for (int x = lane; x < w; x += nlanes) {
            int32_t* p = out + (size_t)y * blk.pitch + x;
            int32_t v = 0;
            if ((sg >> x) & 1ull) {
                const int32_t m = firstfold ? 0 : *p;
                const int qlast = ((la >> x) & 1ull) ? lastplane : lastplane + 1;
                const int32_t t1 = (m << 1) | (1 << qlast);          // one extra fractional bit, t1.cpp:419-445
                v = ((ng >> x) & 1ull) ? -t1 : t1;
            }
            if (raw_output) *p = v;
            else if (blk.qmfbid) *p = v / 2;
            else *reinterpret_cast<float*>(p) = J2K_FMUL((float)v, blk.stepsize);
        }
    }
}

}  // namespace t1dec
}  // namespace j2k
