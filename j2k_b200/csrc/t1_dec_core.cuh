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
#define J2K_FFS(v) __ffs((int)(v))
#define J2K_CLZ(v) __clz((int)(v))
#define J2K_SYNCWARP() __syncwarp()
#define J2K_FMUL(a, b) __fmul_rn((a), (b))
#define J2K_NOINLINE __noinline__
#else
#define J2K_NOINLINE
#define J2K_FFSLL(v) __builtin_ffsll((long long)(v))
#define J2K_FFS(v) __builtin_ffs((int)(v))
#define J2K_CLZ(v) __builtin_clz((unsigned)(v))
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
// The byte window of the segment lives in shared memory (one per warp), not in registers: it is touched only when the
// code register runs dry, about once per ten decisions, and that refill is ONE out-of-line routine instead of a copy at
// each of the eight places a decision is decoded -- the decision loops stay small enough for the instruction cache and
// the decoder state in registers shrinks to a, c, ct.
struct MqIo {
    const u64* wp;            // next word to fetch
    u64 w0, w1;               // bytes [8k, 8k+8) and [8k+8, 8k+16) of the segment
    uint32_t off;             // position of the current byte inside w0
    uint32_t pad;
};
J2K_HD inline void mq_io_bytein(MqIo* io, uint32_t& c, uint32_t& ct) {            // mqc.cpp:336-365
    const u64 w0 = io->w0;
    const uint32_t off = io->off;
    const uint32_t cur = (uint32_t)(w0 >> (8 * off)) & 0xffu;
    const uint32_t nxt = off < 7 ? (uint32_t)(w0 >> (8 * off + 8)) & 0xffu : (uint32_t)io->w1 & 0xffu;
    bool adv = true;
    if (cur == 0xffu) {
        if (nxt > 0x8fu) { c += 0xff00u; ct = 8; adv = false; }
        else { c += nxt << 9; ct = 7; }
    } else { c += nxt << 8; ct = 8; }
    if (adv) {
        if (off == 7) { io->off = 0; io->w0 = io->w1; io->w1 = *io->wp; io->wp = io->wp + 1; }
        else io->off = off + 1;
    }
}
// shift the code register by n > ct bits, fetching bytes as they are needed (mqc.cpp:367-378); returns c | ct << 32
J2K_HD J2K_NOINLINE u64 mq_refill(MqIo* io, uint32_t c, uint32_t ct, uint32_t n) {
    do {
        c <<= ct; n -= ct;
        mq_io_bytein(io, c, ct);
    } while (n > ct);
    c <<= n; ct -= n;
    return (u64)c | ((u64)ct << 32);
}

struct MqDec {
    uint32_t a, c, ct;        // a is kept shifted left by 16 so that it lines up with the code bits of c
    MqIo* io;
    J2K_HD void init(MqIo* io_, const uint8_t* p) {       // mqc.cpp:560-575
        io = io_;
        const u64* wp = reinterpret_cast<const u64*>(p);
        io->w0 = wp[0]; io->w1 = wp[1]; io->wp = wp + 2; io->off = 0;
        c = ((uint32_t)io->w0 & 0xffu) << 16;
        mq_io_bytein(io, c, ct);
        c <<= 7; ct -= 7; a = 0x80000000u;
    }
    J2K_HD inline void renorm() {                         // a < 0x8000: shift until the top bit is set, refilling c
        const uint32_t n = (uint32_t)J2K_CLZ(a);
        a <<= n;
        if (n > ct) { const u64 r = mq_refill(io, c, ct, n); c = (uint32_t)r; ct = (uint32_t)(r >> 32); }
        else { c <<= n; ct -= n; }
    }
    // st: per-context state WORD = the probability-table entry of the context's current state:
    //   qe << 16 | next-if-LPS << 9 | next-if-MPS << 2 | mps   (next-* are byte offsets into the table, tab = mq_state_words).
    // One shared-memory look-up per decision, a second only when the state moves.
    J2K_HD inline uint32_t decode(uint32_t* st, const uint32_t* tab, uint32_t ctx) {   // mqc.cpp:577-595
        const uint32_t e = st[ctx], qe = e & 0xffff0000u;
        const uint32_t d = e & 1u;
        a -= qe;
        const bool lower = c < qe;                       // the code lies in the LPS sub-interval
        if (!lower) {
            c -= qe;
            if ((int32_t)a < 0) return d;                // MPS, no renormalisation: the common case
        }
        // one copy of the state move and the renormalisation for the three remaining cases: the symbol is the LPS when
        // exactly one of "code in the lower sub-interval" and "conditional exchange (a < qe)" holds
        const uint32_t flip = (uint32_t)lower ^ (uint32_t)(a < qe);
        st[ctx] = *reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(tab) + ((flip ? (e >> 7) : e) & 0x1fcu));
        if (lower) a = qe;
        renorm();
        return d ^ flip;
    }
    // The same step with the state word of the context used last held in a register (lc, le): runs of one context --
    // the refinement pass of a dense block codes almost everything with MAG + 2 -- skip the shared-memory store and
    // the load that would have to wait for it.  st[lc] is stale until flush_cached().
    J2K_HD inline uint32_t decode_cached(uint32_t* st, const uint32_t* tab, uint32_t ctx, uint32_t& lc, uint32_t& le) {
        if (ctx != lc) { st[lc] = le; le = st[ctx]; lc = ctx; }
        const uint32_t e = le, qe = e & 0xffff0000u;
        const uint32_t d = e & 1u;
        a -= qe;
        const bool lower = c < qe;
        if (!lower) {
            c -= qe;
            if ((int32_t)a < 0) return d;
        }
        const uint32_t flip = (uint32_t)lower ^ (uint32_t)(a < qe);
        le = *reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(tab) + ((flip ? (e >> 7) : e) & 0x1fcu));
        if (lower) a = qe;
        renorm();
        return d ^ flip;
    }
};

// state word table from T1Tables::mq (entry s = state index * 2 + mps: qe | next-if-MPS << 16 | next-if-LPS << 24)
J2K_HD inline uint32_t mq_state_word(const uint32_t* mq, int s) {
    const uint32_t m = mq[s];
    // qe << 16 | next-if-LPS << 9 | next-if-MPS << 2 | mps: the next-state fields are BYTE offsets into the table (index * 4)
    return ((m & 0xffffu) << 16) | ((m >> 24) << 9) | (((m >> 16) & 0x7fu) << 2) | (uint32_t)(s & 1);
}

J2K_HD inline uint32_t w3(u64 m, int x) {          // bits x-1, x, x+1 of a row
    return (uint32_t)(x ? (m >> (x - 1)) : (m << 1)) & 7u;
}

struct Stripe {               // rows q = 0..5 <-> y = 4*stripe - 1 + q
    u64 S[6], G[6];           // significance, sign
    u64 V[4], F[4], B[4], L[4];   // visited, refined, this plane's bits, coded in this plane
};

// The column being scanned is worked on in packed form so that the row loop can stay rolled (one copy
// of the MQ decoder per call site keeps the kernel inside the instruction cache):
//   win / gwin  3-bit windows (columns x-1, x, x+1) of the six rows around the stripe, row q at bits
//               3q..3q+2, for significance and sign; the 9 bits starting at 3(q-1) are the 3x3
//               neighbourhood of the sample in row q (the layout the reference's packed flags use,
//               t1_opt.h:57-168), which indexes the zero-coding table directly
//   nibbles     one bit per stripe row
J2K_HD inline uint32_t pack_win(const u64* rows, int x) {
    uint32_t w = 0;
#pragma unroll
    for (int q = 0; q < 6; ++q) w |= w3(rows[q], x) << (3 * q);
    return w;
}
J2K_HD inline uint32_t nibble(const u64* rows, int x) {
    return (uint32_t)((rows[0] >> x) & 1ull) | ((uint32_t)((rows[1] >> x) & 1ull) << 1) |
           ((uint32_t)((rows[2] >> x) & 1ull) << 2) | ((uint32_t)((rows[3] >> x) & 1ull) << 3);
}
J2K_HD inline void scatter(u64* rows, uint32_t nib, int x) {
#pragma unroll
    for (int r = 0; r < 4; ++r) rows[r] |= (u64)((nib >> r) & 1u) << x;
}

// Sign-coding table index from the 9-bit significance and sign neighbourhoods (bits 1 N, 3 W, 5 E, 7 S): the
// significance bits go to the even positions, the sign bits stay on the odd ones, so nothing has to be gathered.
// sc_entry() gives the table in that order from T1Tables::sc (which is indexed W | E<<1 | N<<2 | S<<3 | signs << 4).
J2K_HD inline uint32_t sc_index(uint32_t s9, uint32_t g9) {
    return ((s9 >> 1) & 0x55u) | (g9 & s9 & 0xaau);
}
J2K_HD inline uint8_t sc_entry(const uint8_t* sc, int i) {
    const int sN = i & 1, gN = (i >> 1) & 1, sW = (i >> 2) & 1, gW = (i >> 3) & 1, sE = (i >> 4) & 1, gE = (i >> 5) & 1,
              sS = (i >> 6) & 1, gS = (i >> 7) & 1;
    return sc[sW | (sE << 1) | (sN << 2) | (sS << 3) | (gW << 4) | (gE << 5) | (gN << 6) | (gS << 7)];
}

// Per-warp scratch of the stripe being scanned, in column-major form: the row masks are turned into
// per-column words once per stripe by all lanes in parallel (two columns per lane), the serial decision
// loop then touches one word per column, and its results go back into the row masks with warp ballots.
//   win[x]  18-bit significance window of column x (see pack_win)
//   aux[x]  nibbles: bits 0-3 samples to code in this pass (membership / validity), 4-7 refined
//   out[x]  nibbles: 0-3 newly significant, 4-7 negative, 8-11 visited, 12-15 decoded magnitude bit
struct ColBuf {
    uint32_t gwin[64];        // sign window of column x, same layout as win (significance and cleanup passes)
    uint32_t win[64];
    uint32_t aux[64];
    uint32_t out[64];
};

// bit `b` of out[x] of all 64 columns as a row mask; o0 / o1 are this lane's out[lane] / out[lane + 32]
J2K_HD inline u64 rowmask(const uint32_t* out, uint32_t o0, uint32_t o1, int b) {
#if defined(__CUDA_ARCH__)
    (void)out;
    return (u64)__ballot_sync(0xffffffffu, (o0 >> b) & 1u) | ((u64)__ballot_sync(0xffffffffu, (o1 >> b) & 1u) << 32);
#else
    (void)o0; (void)o1;
    u64 m = 0;
    for (int x = 0; x < 64; ++x) m |= (u64)((out[x] >> b) & 1u) << x;
    return m;
#endif
}
// decisions of a column's member rows in list order -> one bit per stripe row
J2K_HD inline uint32_t mr_rows(uint32_t res, uint32_t mem4) {
    uint32_t out = 0;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const uint32_t m = (mem4 >> r) & 1u;
        out |= (res & m) << r;
        res >>= m;
    }
    return out;
}
// nibble -> the window bits of rows 1..4 at column offset 0 (bits 3, 6, 9, 12)
J2K_HD inline uint32_t spread4(uint32_t n) {
    n = (n | (n << 4)) & 0xc3u;            // bits 0 1 . . . . 2 3
    n = (n | (n << 2)) & 0x249u;           // bits 0, 3, 6, 9
    return n << 3;
}

// Decode one code-block.  Called by all `nlanes` lanes of a warp with identical arguments except `lane`;
// the decision loop runs redundantly on warp-uniform data, the lane index only partitions the
// column-parallel parts (stripe set-up and write-back, fold, final pass).  R and st must be zero / reset on entry.
// Raw ("lazy", selective arithmetic coding bypass) bit reader: raw.cpp:91-117.  A byte that follows 0xFF carries 7 bits.
struct RawDec {
    const uint8_t* p; uint32_t len, pos, c, ct;
    J2K_HD inline void init(const uint8_t* data, uint32_t n) { p = data; len = n; pos = 0; c = 0; ct = 0; }
    J2K_HD inline uint32_t get() {
        if (ct == 0) {
            ct = 8;
            if (pos == len) c = 0xff;
            else { if (c == 0xff) ct = 7; c = p[pos++]; }
        }
        --ct;
        return (c >> ct) & 1u;
    }
};
enum { CBLK_LAZY = 1, CBLK_RESET = 2, CBLK_TERMALL = 4, CBLK_VSC = 8, CBLK_PTERM = 16, CBLK_SEGSYM = 32 };   // j2k.h:77-82

J2K_HD inline void reset_contexts(uint32_t* st, const uint32_t* mqtab, int lane, int nlanes) {     // t1.cpp:1391-1394
    for (int i = lane; i < T1_NCTX; i += nlanes)
        st[i] = mqtab[i == CTX_UNI ? 46 * 2 : (i == CTX_AGG ? 3 * 2 : (i == CTX_ZC ? 4 * 2 : 0))];
    J2K_SYNCWARP();
}

// GEN = false: the plain coder (cblksty == 0, one codeword segment, no region of interest) -- what every configuration of
// the encoder writes.  GEN = true adds what foreign streams may use (t1.cpp:1337-1448): several segments per block
// (TERMALL, LAZY), raw passes (LAZY, raw.cpp), context reset after every pass (RESET), vertically causal contexts (VSC),
// the segmentation symbol (SEGSYM) and the ROI down-shift (T1Decoder.cpp:59-76).
template <bool GEN>
J2K_HD inline void decode_block(const T1DecBlock& blk, DRows& R, ColBuf& cb, MqIo* io, uint32_t* st, const uint8_t* zc_all, const uint8_t* sc,
                                const uint32_t* mqtab, const uint8_t* seg_arena, const T1DecSeg* segtab, int raw_output, int lane, int nlanes) {
    const int w = blk.w, h = blk.h;
    int32_t* out = blk.coeff;
    const uint8_t* zc = zc_all + blk.orient * 512;
    const int nstripes = (h + 3) >> 2;
    const u64 colmask = w == 64 ? ~0ull : ((1ull << w) - 1);
    const int sty = GEN ? blk.cblksty : 0;
    const bool vsc = GEN && (sty & CBLK_VSC);

    MqDec mq;
    RawDec rawd;
    bool raw = false;
    int segno = 0, seg_left = blk.npasses;
    if (!GEN) mq.init(io, seg_arena + blk.data_off);
    else { seg_left = 0; segno = -1; rawd.init(seg_arena, 0); }

    int bp = blk.numbps + (GEN ? blk.roishift : 0) - 1;              // 0-based bit-plane being decoded
    int passtype = 2;
    int lastplane = bp;
    bool firstfold = true;
    for (int passno = 0; passno < blk.npasses && bp >= 0; ++passno) {
        if (GEN && seg_left == 0) {                  // next codeword segment: restart the MQ decoder or the raw reader on its bytes
            ++segno;
            const T1DecSeg sg = blk.nsegs > 1 ? segtab[blk.seg_index + segno] : T1DecSeg{blk.data_off, blk.data_len, blk.npasses, 0};
            seg_left = (int)sg.passes;
            raw = (sty & CBLK_LAZY) && bp + 1 <= (int)blk.numbps - 4 && passtype < 2;
            if (raw) rawd.init(seg_arena + sg.off, sg.len); else mq.init(io, seg_arena + sg.off);
            J2K_SYNCWARP();
        }
        for (int sidx = 0; sidx < nstripes; ++sidx) {
            const int ib = 4 * sidx;       // smem index of row q = 0
            u64 S[6], V[4], valid[4];
#pragma unroll
            for (int q = 0; q < 6; ++q) S[q] = R.sig[ib + q];
            if (vsc) S[5] = 0;                       // vertically causal: the stripe below does not exist for the contexts
#pragma unroll
            for (int r = 0; r < 4; ++r) { V[r] = R.vis[ib + 1 + r]; valid[r] = (ib + r < h) ? colmask : 0ull; }
            // ---- which columns can hold anything to decode in this pass ----
            u64 memb[4], cols = 0, memall = 0;
            if (passtype == 0) {
                // significance propagation: only columns that hold a candidate (an insignificant sample with a
                // significant neighbour) are visited; a sample that turns significant adds the next column
                u64 Hd[6];
#pragma unroll
                for (int q = 0; q < 6; ++q) Hd[q] = (S[q] << 1) | S[q] | (S[q] >> 1);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    memb[r] = ~S[r + 1] & valid[r];
                    cols |= memb[r] & (Hd[r] | Hd[r + 2] | (S[r + 1] << 1) | (S[r + 1] >> 1));
                }
            } else if (passtype == 1) {
#pragma unroll
                for (int r = 0; r < 4; ++r) { memb[r] = S[r + 1] & ~V[r] & valid[r]; cols |= memb[r]; }
            } else {
#pragma unroll
                for (int r = 0; r < 4; ++r) { memb[r] = ~S[r + 1] & ~V[r] & valid[r]; cols |= memb[r]; }
                memall = (ib + 3 < h) ? (memb[0] & memb[1] & memb[2] & memb[3]) : 0ull;
            }
            if (cols == 0) {
                if (passtype == 2 && (V[0] | V[1] | V[2] | V[3])) {
                    if (lane < 4) R.vis[ib + 1 + lane] = 0;
                    if (nlanes == 1) { R.vis[ib + 1] = R.vis[ib + 2] = R.vis[ib + 3] = R.vis[ib + 4] = 0; }
                    J2K_SYNCWARP();
                }
                continue;
            }
            // ---- column-major set-up, two columns per lane ----
            for (int x = lane; x < 64; x += nlanes) {
                cb.win[x] = pack_win(S, x);
                if (passtype != 1) {
                    u64 G[6];
#pragma unroll
                    for (int q = 0; q < 6; ++q) G[q] = R.sgn[ib + q];
                    if (vsc) G[5] = 0;
                    cb.gwin[x] = pack_win(G, x);
                }
                uint32_t a = nibble(memb, x);
                if (passtype == 1) {
                    // magnitude refinement contexts depend on nothing decoded in this pass: the column's list of
                    // context codes (2 bits each, 1 = MAG, 2 = MAG + 1, 3 = MAG + 2, in row order) is made here
                    const uint32_t ref4 = nibble(&R.ref[ib + 1], x), w18 = cb.win[x];
                    uint32_t list = 0, k = 0;
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        if (!((a >> r) & 1u)) continue;
                        const uint32_t code = ((ref4 >> r) & 1u) ? 3u : (((w18 >> (3 * r)) & 0x1efu) ? 2u : 1u);
                        list |= code << k; k += 2;
                    }
                    a |= list << 8;
                }
                cb.aux[x] = a;
                cb.out[x] = 0;
            }
            J2K_SYNCWARP();
            bool any = false;
            int carry_x = -2;
            uint32_t carry_s = 0, carry_g = 0;
            if (passtype == 0) {
                // the 64 columns as two 32-bit halves: find-first-set and the mask updates are single instructions
                uint32_t pend = 0;
#pragma unroll 1
                for (int hh = 0; hh < 2; ++hh) {
                const uint32_t cm32 = (uint32_t)(colmask >> (32 * hh));
                uint32_t active = ((uint32_t)(cols >> (32 * hh)) | pend) & cm32, donec = 0, rem;
                pend = 0;
                while ((rem = active & ~donec) != 0) {
                    const int xl = J2K_FFS(rem) - 1, x = 32 * hh + xl;
                    donec |= 1u << xl;
                    uint32_t win = cb.win[x];
                    if (x == carry_x + 1) win |= carry_s;
                    const uint32_t mem4 = cb.aux[x];
                    // rows of this column to code: insignificant with a significant neighbour.  Bit 3q of
                    // rowany / side = any of row q's three window bits / its left or right one.
                    const uint32_t rowany = (win | (win >> 1) | (win >> 2)) & 0x9249u;
                    const uint32_t side = (win | (win >> 2)) & 0x9249u;
                    const uint32_t nb = (rowany << 3) | (rowany >> 3) | side;       // bit 3(r + 1): sample r has a neighbour
                    uint32_t todo = mem4 & (((nb >> 3) & 1u) | ((nb >> 5) & 2u) | ((nb >> 7) & 4u) | ((nb >> 9) & 8u));
                    uint32_t gwin = 0;
                    bool have_g = false;
                    uint32_t new4 = 0, neg4 = 0, vis4 = todo;
                    while (todo) {
                        const int r = J2K_FFSLL(todo) - 1;
                        todo &= todo - 1;
                        const uint32_t s9 = (win >> (3 * r)) & 0x1ffu;
                        if (GEN && raw) { if (!rawd.get()) continue; }
                        else if (!mq.decode(st, mqtab, zc[s9 & 0x1efu])) continue;
                        if (!have_g) {
                            gwin = cb.gwin[x];
                            if (x == carry_x + 1) gwin |= spread4(carry_g);
                            have_g = true;
                        }
                        uint32_t neg;
                        if (GEN && raw) neg = rawd.get();          // the sign itself, no context (t1.cpp:396-399)
                        else {
                            const uint32_t e = sc[sc_index(s9, (gwin >> (3 * r)) & 0x1ffu)];
                            neg = mq.decode(st, mqtab, CTX_SC + (e & 7u)) ^ (e >> 3);
                        }
                        win |= 2u << (3 * r + 3);
                        gwin |= (neg << 1) << (3 * r + 3);
                        new4 |= 1u << r; neg4 |= neg << r;
                        const uint32_t below = mem4 & (2u << r);                   // the row below now has a significant neighbour
                        todo |= below; vis4 |= below;
                    }
                    if (vis4) { cb.out[x] = new4 | (neg4 << 4) | (vis4 << 8); any = true; }
                    if (new4) {
                        carry_x = x; carry_s = spread4(new4); carry_g = neg4;
                        active |= (2u << xl) & cm32;                   // the next column may now hold candidates
                        if (xl == 31) pend = 1u;
                    }
                }
                }
            } else if (passtype == 1) {
                // magnitude refinement: significant and not coded in this plane's first pass; decisions come out in
                // list order and are put back on their rows by the lanes afterwards
                uint32_t mr_lc = CTX_MAG + 2, mr_le = st[CTX_MAG + 2];
#pragma unroll 1
                for (int hh = 0; hh < 2; ++hh) {
                uint32_t c32 = (uint32_t)(cols >> (32 * hh));
                while (c32) {
                    const int x = 32 * hh + J2K_FFS(c32) - 1;
                    c32 &= c32 - 1;
                    uint32_t list = cb.aux[x] >> 8, res = 0, k = 0;
                    do {
                        if (GEN && raw) res |= rawd.get() << k;
                        else res |= mq.decode_cached(st, mqtab, CTX_MAG - 1 + (list & 3u), mr_lc, mr_le) << k;
                        ++k; list >>= 2;
                    } while (list);
                    if (res) { cb.out[x] = res; any = true; }
                }
                }
                st[mr_lc] = mr_le;
            } else {
                // cleanup: everything not yet significant and not visited
#pragma unroll 1
                for (int hh = 0; hh < 2; ++hh) {
                uint32_t c32 = (uint32_t)(cols >> (32 * hh));
                const uint32_t mall32 = (uint32_t)(memall >> (32 * hh));
                while (c32) {
                    const int xl = J2K_FFS(c32) - 1, x = 32 * hh + xl;
                    c32 &= c32 - 1;
                    uint32_t win = cb.win[x];
                    if (x == carry_x + 1) win |= carry_s;
                    uint32_t gwin = 0;
                    bool have_g = false;
                    const uint32_t mem4 = cb.aux[x];
                    int rstart = 0;
                    bool partial = false;
                    if (((mall32 >> xl) & 1u) && win == 0) {
                        // run-length mode: the whole column and its neighbourhood are insignificant
                        if (!mq.decode(st, mqtab, CTX_AGG)) continue;
                        rstart = (int)mq.decode(st, mqtab, CTX_UNI);
                        rstart = (rstart << 1) | (int)mq.decode(st, mqtab, CTX_UNI);
                        partial = true;
                    }
                    uint32_t new4 = 0, neg4 = 0;
                    uint32_t todo = mem4 & (0xfu << rstart);
                    while (todo) {
                        const int r = J2K_FFSLL(todo) - 1;
                        todo &= todo - 1;
                        const uint32_t s9 = (win >> (3 * r)) & 0x1ffu;
                        if (partial) partial = false;                  // the sample that ended the run is significant
                        else if (!mq.decode(st, mqtab, zc[s9 & 0x1efu])) continue;
                        if (!have_g) {
                            gwin = cb.gwin[x];
                            if (x == carry_x + 1) gwin |= spread4(carry_g);
                            have_g = true;
                        }
                        const uint32_t e = sc[sc_index(s9, (gwin >> (3 * r)) & 0x1ffu)];
                        const uint32_t neg = mq.decode(st, mqtab, CTX_SC + (e & 7u)) ^ (e >> 3);
                        win |= 2u << (3 * r + 3);
                        gwin |= (neg << 1) << (3 * r + 3);
                        new4 |= 1u << r; neg4 |= neg << r;
                    }
                    if (new4) {
                        cb.out[x] = new4 | (neg4 << 4); any = true;
                        carry_x = x; carry_s = spread4(new4); carry_g = neg4;
                    }
                }
                }
            }
            J2K_SYNCWARP();
            // ---- results back into the row masks (rows of the neighbouring stripes are unchanged) ----
            uint32_t o0 = 0, o1 = 0;
#if defined(__CUDA_ARCH__)
            if (any) {
                o0 = cb.out[lane]; o1 = cb.out[lane + 32];
                if (passtype == 1) { o0 = mr_rows(o0, cb.aux[lane] & 15u) << 12; o1 = mr_rows(o1, cb.aux[lane + 32] & 15u) << 12; }
            }
#else
            if (any && passtype == 1)
                for (int x = 0; x < 64; ++x) cb.out[x] = mr_rows(cb.out[x], cb.aux[x] & 15u) << 12;
#endif
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int i = ib + 1 + r;
                if (passtype == 1) {
                    u64 bits = any ? rowmask(cb.out, o0, o1, 12 + r) : 0ull;
                    if (lane == 0) { R.bit[i] |= bits; R.last[i] |= memb[r]; R.ref[i] |= memb[r]; }
                } else {
                    u64 nw = 0, ng = 0, vs = 0;
                    if (any) {
                        nw = rowmask(cb.out, o0, o1, r); ng = rowmask(cb.out, o0, o1, 4 + r);
                        if (passtype == 0) vs = rowmask(cb.out, o0, o1, 8 + r);
                    }
                    if (lane == 0) {
                        if (nw) { R.sig[i] = S[r + 1] | nw; R.sgn[i] |= ng; R.bit[i] |= nw; R.last[i] |= nw; }
                        if (passtype == 0) { if (vs) R.vis[i] = V[r] | vs; }
                        else R.vis[i] = 0;
                    }
                }
            }
            J2K_SYNCWARP();
        }
        lastplane = bp;
        if (GEN) {
            if (passtype == 2 && (sty & CBLK_SEGSYM)) for (int k = 0; k < 4; ++k) mq.decode(st, mqtab, CTX_UNI);   // t1.cpp:1099-1111
            if ((sty & CBLK_RESET) && !raw) { J2K_SYNCWARP(); reset_contexts(st, mqtab, lane, nlanes); }
            --seg_left;
        }
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
                int32_t t1 = (m << 1) | (1 << qlast);                // one extra fractional bit, t1.cpp:419-445
                if (GEN && blk.roishift && t1 >= (1 << blk.roishift)) t1 >>= blk.roishift;    // T1Decoder.cpp:59-76
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
