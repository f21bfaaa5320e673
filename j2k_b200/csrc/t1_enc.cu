// EBCOT Tier-1 encoder for sm_100a, split in two kernels.
//
//  k_t1_context  one code-block per warp.  Quantises the block (T1Encoder.cpp:103-174), keeps the
//                significance / sign / visited / refined state of its up-to-64 rows as 64-bit row
//                masks (bit x = column x) built with __ballot_sync, and produces the exact sequence
//                of (context, decision) pairs the reference's three coding passes would feed to the
//                MQ coder (t1_opt.cpp:330-576), plus the per-pass distortion estimates
//                (t1.cpp:298-314).  Only the significance-propagation pass has a true scan-order
//                dependency; it is resolved as a monotone fix-point on the row masks (lane = stripe,
//                neighbour stripes exchanged with warp shuffles).  Given the membership, every
//                sample's context depends only on masks "before" and "after" the pass, so the 32
//                lanes emit disjoint (stripe, column-range) pieces of the pass in parallel at
//                offsets obtained from popcounts and a warp prefix sum.
//  k_t1_mq       one code-block per thread: the MQ arithmetic coder proper (mqc.cpp:223-452) with
//                A, C, CT and the pending byte in registers, consuming the symbol stream and
//                recording the byte count after every pass (t1_opt.cpp:711-736).
//
// Scan order (t1_opt.cpp:384-400): stripes of four rows, columns left to right inside a stripe, rows
// top to bottom inside a column.
#include "device.h"
#include "t1_tables.cuh"
#include "t1_mq_gen.cuh"
#include <cstdlib>

namespace j2k {
namespace {

typedef unsigned long long u64;
constexpr int CX_WARPS = 4;
constexpr unsigned FULL = 0xffffffffu;

struct Rows {                 // index = row + 1; entries 0 and 65 stay zero
    u64 sgn[66], sig[66], bit[66], mem[66], nsp[66], ref[66];
};
struct Luts {
    uint8_t zc[4 * 256];
    uint8_t sc[256];
    int16_t nm_sig[128], nm_sig0[128], nm_ref[128], nm_ref0[128];
};

__device__ __forceinline__ int fix_mul_t1(int a, int b) {      // opj_intmath.h:268-279
    return (int)(((long long)a * (long long)b + 4096) >> 18);
}

// sign-magnitude sample: bit 31 sign, 6 fractional bits (T1Encoder.cpp:139-171, t1_opt.h:169).  `mul` = the band's Q13 constant
// for 9/7; for 5/3 the reference takes v * 64, which is the same fixed-point product with mul = 2^24 (v * 2^24 is a multiple of
// 2^18, so the rounding term never carries): one path, no per-sample select.
__device__ __forceinline__ int smag_mul(const T1Block& b) { return b.qmfbid ? (1 << 24) : b.bandconst; }
__device__ __forceinline__ uint32_t to_smag(int v, int mul) {
    const int t = fix_mul_t1(v, mul);
    const uint32_t mag = (uint32_t)(t < 0 ? -t : t);
    return mag | ((uint32_t)t & 0x80000000u);
}
__device__ __forceinline__ uint32_t load_smag(const T1Block& b, int x, int y, int mul) {
    return to_smag(__ldg(b.coeff + (size_t)y * b.pitch + x), mul);
}

// nmsedec tables: plane 0 has tables of its own (t1_luts.h lut_nmsedec_sig0 / _ref0); they follow the general ones in Luts, so the
// choice is an offset that is fixed per bit-plane (nm_off = bpno > 0 ? 0 : 128)
__device__ __forceinline__ int nmsedec_sig(const Luts& L, uint32_t mag, int bpno, int nm_off) {
    return (&L.nm_sig[0])[nm_off + ((mag >> bpno) & 127)];
}
__device__ __forceinline__ int nmsedec_ref(const Luts& L, uint32_t mag, int bpno, int nm_off) {
    return (&L.nm_ref[0])[nm_off + ((mag >> bpno) & 127)];
}

// Symbol writer of one lane: one byte store per symbol.  (Assembling aligned 32-bit words per lane cost ten instructions per
// symbol plus a ragged first / last word per piece -- 27 % of the kernel's instructions -- for pieces that are a few symbols
// long; L2 merges the byte stores of a sector.)
struct Emitter {
    uint8_t* p;
    __device__ __forceinline__ void init(uint8_t* b, uint32_t start) { p = b + start; }
    __device__ __forceinline__ void push(uint32_t ctx, uint32_t d) { *p++ = (uint8_t)((ctx << 1) | d); }
    __device__ __forceinline__ void finish() {}
};

enum { PASS_SIG = 0, PASS_REF = 1, PASS_CLN = 2 };

__device__ __forceinline__ u64 win(u64 m, int c0) { return c0 ? (m >> (c0 - 1)) : (m << 1); }
__device__ __forceinline__ uint32_t w3(u64 m, int j) {          // bits j..j+2 of a windowed row, j <= 31
    return __funnelshift_r((uint32_t)m, (uint32_t)(m >> 32), j) & 7u;
}

// Emit one coding pass of one block.  Returns the number of symbols of the pass (warp-uniform).
// `sym` points at the block's symbol slot, `done` symbols are already there.
template <int PASS>
__device__ uint32_t emit_pass(const Rows& R, const Luts& L, const T1Block& blk, int lane, uint8_t* sym,
                              uint32_t done, uint32_t cap, bool& overflow) {
    const int w = blk.w, h = blk.h;
    const int nstripes = (h + 3) >> 2;
    int p2 = 1;
    while (p2 < nstripes) p2 <<= 1;
    const int lps = 32 / p2;                       // lanes per stripe
    const int segw = (w + lps - 1) / lps;          // <= 32
    const int stripe = lane / lps, seg = lane % lps;
    const int c0 = seg * segw;
    int ncols = min(segw, w - c0);
    if (stripe >= nstripes || ncols < 0) ncols = 0;
    const u64 colmask = w == 64 ? ~0ull : ((1ull << w) - 1);
    const uint32_t lmask = ncols >= 32 ? 0xffffffffu : ((1u << ncols) - 1u);
    const uint8_t* zc = L.zc + blk.orient * 256;

    // rows q = 0..5 <-> y = 4*stripe - 1 + q ; windowed so that local column j sits at bit j+1
    u64 A[6], B[6], SG[6];
    uint32_t MB[4], BT[4], RF[4];
    const int ybase = 4 * (stripe < nstripes ? stripe : 0);
#pragma unroll
    for (int q = 0; q < 6; ++q) {
        const int i = ybase + q;                   // = y + 1
        const u64 s0 = R.sig[i], np = R.nsp[i], mm = R.mem[i], bt = R.bit[i];
        const u64 valid = (i >= 1 && i <= h) ? colmask : 0ull;
        u64 a, b, memb;
        if (PASS == PASS_SIG) { b = s0; a = s0 | np; memb = mm; }
        else if (PASS == PASS_REF) { a = b = s0 | np; memb = s0; }
        else { b = s0 | np; memb = ~(b | mm) & valid; a = b | (memb & bt); }
        A[q] = win(a, c0); B[q] = win(b, c0); SG[q] = win(R.sgn[i], c0);
        if (q >= 1 && q <= 4) {
            MB[q - 1] = (uint32_t)(memb >> c0) & lmask;
            BT[q - 1] = (uint32_t)(bt >> c0) & lmask;
            RF[q - 1] = (uint32_t)(R.ref[i] >> c0) & lmask;
        }
    }
    uint32_t AG = 0;                               // run-length (aggregation) columns, t1_opt.cpp:487-515
    if (PASS == PASS_CLN) {
        // a column is coded in run-length mode when its four samples are all still to be coded
        // and nothing in its 3x6 neighbourhood is significant at the time the column is reached:
        // row above and left column in their state after this pass, right column and row below
        // in their state before it
        u64 up, left = 0, right = 0, own = 0, down, memall = ~0ull;
        {
            const int i0 = ybase, i5 = ybase + 5;
            const u64 b0 = R.sig[i0] | R.nsp[i0];
            const u64 valid0 = (i0 >= 1 && i0 <= h) ? colmask : 0ull;
            const u64 a0 = b0 | (~(b0 | R.mem[i0]) & valid0 & R.bit[i0]);
            up = (a0 << 1) | a0 | (a0 >> 1);
            const u64 b5 = R.sig[i5] | R.nsp[i5];
            down = (b5 << 1) | b5 | (b5 >> 1);
        }
#pragma unroll
        for (int q = 1; q <= 4; ++q) {
            const int i = ybase + q;
            const u64 b = R.sig[i] | R.nsp[i];
            const u64 valid = (i >= 1 && i <= h) ? colmask : 0ull;
            const u64 memb = ~(b | R.mem[i]) & valid;
            const u64 a = b | (memb & R.bit[i]);
            left |= a << 1; right |= b >> 1; own |= b; memall &= memb;
        }
        const bool fullstripe = ybase + 3 < h;
        u64 agg = fullstripe ? (~(up | left | right | own | down) & memall) : 0ull;
        AG = (uint32_t)(agg >> c0) & lmask;
    }

    // ---- count ----
    uint32_t cnt = 0;
    if (PASS == PASS_SIG) {
#pragma unroll
        for (int r = 0; r < 4; ++r) cnt += __popc(MB[r]) + __popc(MB[r] & BT[r]);
    } else if (PASS == PASS_REF) {
#pragma unroll
        for (int r = 0; r < 4; ++r) cnt += __popc(MB[r]);
    } else {
        const uint32_t NA = ~AG;
#pragma unroll
        for (int r = 0; r < 4; ++r) cnt += __popc(MB[r] & NA) + __popc(MB[r] & NA & BT[r]);
        const uint32_t f0 = AG & BT[0], f1 = AG & ~BT[0] & BT[1], f2 = AG & ~BT[0] & ~BT[1] & BT[2];
        const uint32_t f3 = AG & ~BT[0] & ~BT[1] & ~BT[2] & BT[3];
        const uint32_t fz = AG & ~(BT[0] | BT[1] | BT[2] | BT[3]);
        cnt += __popc(fz);                                           // AGG 0
        cnt += 4 * (__popc(f0) + __popc(f1) + __popc(f2) + __popc(f3));   // AGG 1, two UNI, sign
        cnt += 3 * __popc(f0) + 2 * __popc(f1) + __popc(f2);         // zero coding of the rows below
        cnt += __popc(f0 & BT[1]) + __popc(f0 & BT[2]) + __popc(f0 & BT[3]);   // their signs
        cnt += __popc(f1 & BT[2]) + __popc(f1 & BT[3]) + __popc(f2 & BT[3]);
    }
    if (ncols == 0) cnt = 0;
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(FULL, incl, d);
        if (lane >= d) incl += t;
    }
    const uint32_t total = __shfl_sync(FULL, incl, 31);
    if (done + total > cap) { overflow = true; return total; }
    if (cnt == 0) return total;

    Emitter E;
    E.init(sym, done + incl - cnt);
    uint32_t cols = MB[0] | MB[1] | MB[2] | MB[3];
    while (cols) {
        const int j = __ffs(cols) - 1;
        cols &= cols - 1;
        uint32_t wa[6], wb[6], ws[6];
#pragma unroll
        for (int q = 0; q < 6; ++q) { wa[q] = w3(A[q], j); wb[q] = w3(B[q], j); ws[q] = w3(SG[q], j); }
        int rstart = 0;
        bool pending_sign = false;
        if (PASS == PASS_CLN && ((AG >> j) & 1)) {
            const uint32_t bb = ((BT[0] >> j) & 1) | (((BT[1] >> j) & 1) << 1) | (((BT[2] >> j) & 1) << 2) | (((BT[3] >> j) & 1) << 3);
            if (bb == 0) { E.push(CTX_AGG, 0); continue; }
            const int r0 = __ffs(bb) - 1;
            E.push(CTX_AGG, 1);
            E.push(CTX_UNI, (uint32_t)r0 >> 1);
            E.push(CTX_UNI, (uint32_t)r0 & 1);
            rstart = r0;
            pending_sign = true;
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (r < rstart || !((MB[r] >> j) & 1)) continue;
            const int q = r + 1;
            const uint32_t bit = (BT[r] >> j) & 1;
            if (PASS == PASS_REF) {
                const uint32_t nb = wa[q - 1] | (wa[q] & 5u) | wa[q + 1];
                const uint32_t ctx = ((RF[r] >> j) & 1) ? CTX_MAG + 2 : (nb ? CTX_MAG + 1 : CTX_MAG);
                E.push(ctx, bit);
                continue;
            }
            if (!(pending_sign && r == rstart)) {
                const uint32_t above = r == 0 ? wa[0] : ((wa[q - 1] & 3u) | (wb[q - 1] & 4u));
                const uint32_t below = r == 3 ? wb[5] : ((wa[q + 1] & 1u) | (wb[q + 1] & 6u));
                const uint32_t nb = above | ((wa[q] & 1u) << 3) | (((wb[q] >> 2) & 1u) << 4) | (below << 5);
                E.push(zc[nb], bit);
            }
            if (bit) {
                const uint32_t sw = wa[q] & 1u, se = (wb[q] >> 2) & 1u;
                const uint32_t sn = (wa[q - 1] >> 1) & 1u, ss = (wb[q + 1] >> 1) & 1u;
                const uint32_t idx = sw | (se << 1) | (sn << 2) | (ss << 3) |
                                     ((sw & ws[q]) << 4) | ((se & (ws[q] >> 2)) << 5) |
                                     ((sn & (ws[q - 1] >> 1)) << 6) | ((ss & (ws[q + 1] >> 1)) << 7);
                const uint32_t e = L.sc[idx];
                E.push(CTX_SC + (e & 7u), ((ws[q] >> 1) & 1u) ^ (e >> 3));
            }
        }
    }
    E.finish();
    return total;
}

template <int MINB>
__global__ void __launch_bounds__(32 * CX_WARPS, MINB)
k_t1_context(const T1Block* __restrict__ blocks, int nblocks, const T1Tables* __restrict__ tab,
             uint8_t* __restrict__ sym_arena, uint32_t* __restrict__ pass_nsym, T1Result* __restrict__ results) {
    __shared__ Rows s_rows[CX_WARPS];
    __shared__ Luts L;
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) L.zc[i] = tab->zc[i];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) L.sc[i] = tab->sc[i];
    for (int i = threadIdx.x; i < 128; i += blockDim.x) {
        L.nm_sig[i] = tab->nm_sig[i]; L.nm_sig0[i] = tab->nm_sig0[i];
        L.nm_ref[i] = tab->nm_ref[i]; L.nm_ref0[i] = tab->nm_ref0[i];
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bi = blockIdx.x * CX_WARPS + warp;
    if (bi >= nblocks) return;
    const T1Block blk = blocks[bi];
    Rows& R = s_rows[warp];
    T1Result& res = results[bi];
    const int w = blk.w, h = blk.h;
    {
        u64* z = reinterpret_cast<u64*>(&R);
        for (int i = lane; i < (int)(sizeof(Rows) / sizeof(u64)); i += 32) z[i] = 0;
    }
    __syncwarp();
    const bool c0ok = lane < w, c1ok = lane + 32 < w;

    // ---- sweep 0: signs and the largest magnitude (T1Encoder.cpp:139-171) ----
    const int mul = smag_mul(blk);
    const int32_t* col0 = blk.coeff + (c0ok ? lane : 0);
    const int32_t* col1 = blk.coeff + (c1ok ? lane + 32 : 0);
    uint32_t mx = 0;
    for (int y = 0; y < h; y += 2) {                       // two rows per turn: four loads in flight
        const bool two = y + 1 < h;
        const size_t o0 = (size_t)y * blk.pitch, o1 = (size_t)(two ? y + 1 : y) * blk.pitch;
        const int v00 = __ldg(col0 + o0), v01 = __ldg(col1 + o0), v10 = __ldg(col0 + o1), v11 = __ldg(col1 + o1);
        const uint32_t m00 = c0ok ? to_smag(v00, mul) : 0u, m01 = c1ok ? to_smag(v01, mul) : 0u;
        const uint32_t m10 = c0ok ? to_smag(v10, mul) : 0u, m11 = c1ok ? to_smag(v11, mul) : 0u;
        mx = max(mx, max(m00 & 0x7fffffffu, m01 & 0x7fffffffu));
        const uint32_t lo0 = __ballot_sync(FULL, m00 >> 31), hi0 = __ballot_sync(FULL, m01 >> 31);
        if (lane == 0) R.sgn[y + 1] = (u64)lo0 | ((u64)hi0 << 32);
        if (two) {
            mx = max(mx, max(m10 & 0x7fffffffu, m11 & 0x7fffffffu));
            const uint32_t lo1 = __ballot_sync(FULL, m10 >> 31), hi1 = __ballot_sync(FULL, m11 >> 31);
            if (lane == 0) R.sgn[y + 2] = (u64)lo1 | ((u64)hi1 << 32);
        }
    }
    mx = __reduce_max_sync(FULL, mx);
    int numbps = mx ? (32 - __clz(mx)) - 6 : 0;               // t1_opt.cpp:680
    if (numbps < 0) numbps = 0;
    const int npasses = numbps > 0 ? 3 * numbps - 2 : 0;
    if (lane == 0) { res.numbps = numbps; res.npasses = npasses; res.status = 0; res.total = 0; res.nsym = 0; }
    if (numbps == 0) return;
    __syncwarp();

    uint8_t* sym = sym_arena + (size_t)blk.sym_off * 16;
    uint32_t* pn = pass_nsym + (size_t)bi * T1_MAX_PASSES;
    const int nstripes = (h + 3) >> 2;
    const u64 colmask = w == 64 ? ~0ull : ((1ull << w) - 1);
    uint32_t done = 0;
    int passno = 0;
    bool overflow = false;

    for (int bp = numbps - 1; bp >= 0 && !overflow; --bp) {
        const int one = bp + 6;
        const bool first = bp == numbps - 1;
        const int nm_off = bp > 0 ? 0 : 128;
        // ---- sweep A: this plane's bits; distortion of refinement and of all newly significant ----
        int d_ref = 0, d_sig = 0;
        auto plane_row = [&](int y, uint32_t m0, uint32_t m1) {
            const uint32_t b0 = (m0 >> one) & 1u, b1 = (m1 >> one) & 1u;
            const uint32_t lo = __ballot_sync(FULL, b0), hi = __ballot_sync(FULL, b1);
            if (lane == 0) R.bit[y + 1] = (u64)lo | ((u64)hi << 32);
            const u64 s0 = R.sig[y + 1];
            if (!(lo | hi) && !s0) return;                 // warp-uniform: nothing refined, nothing becomes significant in this row
            if ((s0 >> lane) & 1ull) d_ref += nmsedec_ref(L, m0, bp, nm_off); else if (b0) d_sig += nmsedec_sig(L, m0, bp, nm_off);
            if ((s0 >> (lane + 32)) & 1ull) d_ref += nmsedec_ref(L, m1, bp, nm_off); else if (b1) d_sig += nmsedec_sig(L, m1, bp, nm_off);
        };
        for (int y = 0; y < h; y += 2) {
            const bool two = y + 1 < h;
            const size_t o0 = (size_t)y * blk.pitch, o1 = (size_t)(two ? y + 1 : y) * blk.pitch;
            const int v00 = __ldg(col0 + o0), v01 = __ldg(col1 + o0), v10 = __ldg(col0 + o1), v11 = __ldg(col1 + o1);
            plane_row(y, c0ok ? (to_smag(v00, mul) & 0x7fffffffu) : 0u, c1ok ? (to_smag(v01, mul) & 0x7fffffffu) : 0u);
            if (two) plane_row(y + 1, c0ok ? (to_smag(v10, mul) & 0x7fffffffu) : 0u, c1ok ? (to_smag(v11, mul) & 0x7fffffffu) : 0u);
        }
        d_ref = __reduce_add_sync(FULL, d_ref);
        d_sig = __reduce_add_sync(FULL, d_sig);
        __syncwarp();
        int d_sp = 0;
        if (!first) {
            // ---- significance propagation membership: fix-point over the row masks ----
            u64 e[4], st[4], b[4], M[4], N[4];
            const bool act = lane < nstripes;
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int i = act ? 4 * lane + r + 1 : 0;
                const u64 s0 = R.sig[i], up = R.sig[i > 0 ? i - 1 : 0], dn = R.sig[i + 1];
                const u64 valid = (act && i <= h) ? colmask : 0ull;
                e[r] = ~s0 & valid;
                b[r] = R.bit[i];
                st[r] = (up << 1) | up | (up >> 1) | (s0 << 1) | (s0 >> 1) | (dn << 1) | dn | (dn >> 1);
                M[r] = e[r] & st[r];
                N[r] = M[r] & b[r];
            }
            bool changed;
            do {
                u64 upN = __shfl_up_sync(FULL, N[3], 1);
                if (lane == 0) upN = 0;
                changed = false;
                u64 c, m;
                c = (upN << 1) | upN | (upN >> 1) | (N[0] << 1) | (N[1] << 1);
                m = e[0] & (st[0] | c); changed |= m != M[0]; M[0] = m; N[0] = m & b[0];
                c = (N[0] << 1) | N[0] | (N[1] << 1) | (N[2] << 1);
                m = e[1] & (st[1] | c); changed |= m != M[1]; M[1] = m; N[1] = m & b[1];
                c = (N[1] << 1) | N[1] | (N[2] << 1) | (N[3] << 1);
                m = e[2] & (st[2] | c); changed |= m != M[2]; M[2] = m; N[2] = m & b[2];
                c = (N[2] << 1) | N[2] | (N[3] << 1);
                m = e[3] & (st[3] | c); changed |= m != M[3]; M[3] = m; N[3] = m & b[3];
            } while (__any_sync(FULL, changed));
            if (act) {
#pragma unroll
                for (int r = 0; r < 4; ++r) { R.mem[4 * lane + r + 1] = M[r]; R.nsp[4 * lane + r + 1] = N[r]; }
            }
            __syncwarp();
            // ---- pass: significance propagation ----
            uint32_t n = emit_pass<PASS_SIG>(R, L, blk, lane, sym, done, blk.sym_cap, overflow);
            if (overflow) break;
            done += n;
            // distortion of the samples that became significant in this pass
            for (int y = 0; y < h; ++y) {
                const u64 np = R.nsp[y + 1];
                if (!np) continue;
                if ((np >> lane) & 1ull) d_sp += nmsedec_sig(L, load_smag(blk, lane, y, mul) & 0x7fffffffu, bp, nm_off);
                if ((np >> (lane + 32)) & 1ull) d_sp += nmsedec_sig(L, load_smag(blk, lane + 32, y, mul) & 0x7fffffffu, bp, nm_off);
            }
            d_sp = __reduce_add_sync(FULL, d_sp);
            if (lane == 0) { pn[passno] = n; res.nmsedec[passno] = d_sp; }
            ++passno;
            // ---- pass: magnitude refinement ----
            n = emit_pass<PASS_REF>(R, L, blk, lane, sym, done, blk.sym_cap, overflow);
            if (overflow) break;
            done += n;
            if (lane == 0) { pn[passno] = n; res.nmsedec[passno] = d_ref; }
            ++passno;
        }
        // ---- pass: cleanup ----
        {
            uint32_t n = emit_pass<PASS_CLN>(R, L, blk, lane, sym, done, blk.sym_cap, overflow);
            if (overflow) break;
            done += n;
            if (lane == 0) { pn[passno] = n; res.nmsedec[passno] = d_sig - d_sp; }
            ++passno;
        }
        __syncwarp();
        // ---- state for the next plane ----
        for (int i = lane + 1; i <= h; i += 32) {
            const u64 s0 = R.sig[i], b1 = s0 | R.nsp[i];
            const u64 ncl = ~(b1 | R.mem[i]) & colmask & R.bit[i];
            if (!first) R.ref[i] = s0;
            R.sig[i] = b1 | ncl;
            R.mem[i] = 0; R.nsp[i] = 0;
        }
        __syncwarp();
    }
    if (lane == 0) { res.nsym = done; if (overflow) res.status = 1; }
}

// ---- MQ coder ----------------------------------------------------------------------------------
constexpr int MQ_THREADS = 128;

// Encoder registers of one code-block (mqc.cpp:91-103 keeps them in a struct in memory).  Branches
// cost a warp whose lanes code different blocks the sum of both sides, so the byte output is
// written without any: the carry goes into the pending byte unless that byte is 0xFF, and a pending
// 0xFF makes the next byte a 7-bit one.
struct MqEnc {
    uint32_t a, c, ct;
    int nb;                   // index of the pending byte (bp - start); -1 = none yet
    uint32_t last;            // the pending byte, still in a register (a later carry may bump it)
    uint8_t* out;
    uint32_t cap;
    uint32_t overflow;
    __device__ __forceinline__ void byteout() {                // mqc.cpp:223-257
        const uint32_t carry = last == 0xffu ? 0u : (c >> 27); // before the first byte: last = 0 and c < 2^27
        last += carry;
        c -= carry << 27;
        const uint32_t seven = last == 0xffu ? 1u : 0u;
        if ((uint32_t)nb < cap) out[nb] = (uint8_t)last;       // nb = -1 fails the unsigned test
        else overflow |= (uint32_t)(nb >= 0);
        ++nb;
        const uint32_t sh = 19u + seven;
        last = c >> sh;
        c &= (1u << sh) - 1u;
        ct = 8u - seven;
    }
    __device__ __forceinline__ void renorm() {                 // mqc.cpp:259-269, shift count taken at once
        uint32_t n = __clz(a) - 16;
        a <<= n;
        while (n >= ct) {
            c <<= ct; n -= ct;
            byteout();
        }
        c <<= n; ct -= n;
    }
};

// Block indices by decreasing symbol count (512-symbol buckets); order[nblocks] = how many blocks are at least
// half as long as the longest (at most max_long): those get a warp each in k_t1_mq; order[nblocks + 1] = blocks
// per warp for the rest.  One CTA.
__global__ void __launch_bounds__(1024) k_t1_order(const T1Result* __restrict__ results, int nblocks, int max_long, int lanes,
                                                  uint32_t* __restrict__ order) {
    __shared__ uint32_t s_hist[256], s_off[256];
    __shared__ uint32_t s_max;
    if (threadIdx.x < 256) s_hist[threadIdx.x] = 0;
    if (threadIdx.x == 0) s_max = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < nblocks; i += blockDim.x) {
        const uint32_t n = results[i].npasses > 0 ? results[i].nsym : 0u;
        atomicAdd(&s_hist[min(n >> 9, 255u)], 1u);
        atomicMax(&s_max, n);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0, nlong = 0;
        const uint32_t half = min((s_max / 2) >> 9, 255u);
        for (int b = 255; b >= 0; --b) {
            s_off[b] = run; run += s_hist[b];
            if ((uint32_t)b >= half && b > 0) nlong = run;
        }
        // many blocks about as long as the longest (noise-like content): the kernel is then bound by issue slots
        // rather than by one chain, and full warps are the better shape
        const bool flat = nlong * 4 > (uint32_t)nblocks;
        order[nblocks] = flat ? 0u : min(nlong, (uint32_t)max(max_long, 0));
        order[nblocks + 1] = flat ? 32u : (uint32_t)lanes;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nblocks; i += blockDim.x) {
        const uint32_t n = results[i].npasses > 0 ? results[i].nsym : 0u;
        order[atomicAdd(&s_off[min(n >> 9, 255u)], 1u)] = (uint32_t)i;
    }
}

// One code-block per thread, `lanes` threads of a warp in use.  The kernel is bound by the dependent
// chain of its longest block (a lone warp per scheduler at the end), so the decision step is written
// without branches: every branch costs such a warp a pipeline bubble several times the arithmetic it
// skips.  What is left on the loop-carried chain is  a -= qe -> compare -> select -> clz -> shift  and,
// when one context repeats, one table look-up.  Context states are kept as the probability-table WORD of
// the current state (qe | mps << 15 | next-if-MPS << 16 | next-if-LPS << 24); the word of the next
// symbol's context is fetched before the current one is updated and patched when both are the same.
// Only two rarely taken branches remain per decision: byte output and pass boundary.
__global__ void __launch_bounds__(MQ_THREADS)
k_t1_mq(const T1Block* __restrict__ blocks, int nblocks, const T1Tables* __restrict__ tab,
        const uint8_t* __restrict__ sym_arena, const uint32_t* __restrict__ pass_nsym,
        uint8_t* __restrict__ out_arena, T1Result* __restrict__ results, int lanes,
        const uint32_t* __restrict__ order) {
    __shared__ uint32_t s_mq[96];
    __shared__ uint32_t s_ctx[32][MQ_THREADS];
    if (threadIdx.x < 94) s_mq[threadIdx.x] = tab->mq[threadIdx.x] | ((threadIdx.x & 1u) << 15);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int gwarp = blockIdx.x * (MQ_THREADS / 32) + (threadIdx.x >> 5);
    // slots [0, nlong): one block per warp (the longest blocks, whose chain bounds the kernel: no lane of
    // another block to wait for); the rest: `lanes` blocks per warp.  order[] = blocks by decreasing work.
    const int nlong = order ? (int)order[nblocks] : 0;
    if (order) lanes = (int)order[nblocks + 1];
    int slot;
    if (gwarp < nlong) { if (lane) return; slot = gwarp; }
    else { if (lane >= lanes) return; slot = nlong + (gwarp - nlong) * lanes + lane; }
    if (slot >= nblocks) return;
    const int bi = order ? (int)order[slot] : slot;
    T1Result& res = results[bi];
    const int npasses = res.npasses;
    if (npasses == 0 || res.status != 0) { res.total = 0; return; }
    const T1Block blk = blocks[bi];
    const int t = threadIdx.x;
#pragma unroll
    for (int k = 0; k < 32; ++k) s_ctx[k][t] = s_mq[0];
    s_ctx[CTX_UNI][t] = s_mq[46 * 2]; s_ctx[CTX_AGG][t] = s_mq[3 * 2]; s_ctx[CTX_ZC][t] = s_mq[4 * 2];   // t1_opt.cpp:686-690

    MqEnc q;
    q.a = 0x8000u; q.c = 0; q.ct = 12; q.nb = -1; q.last = 0;                           // mqc.cpp:406-420
    q.out = out_arena + blk.out_off; q.cap = blk.out_cap; q.overflow = 0;
    // The symbol stream ((ctx << 1) | d, one byte per decision) is read 8 bytes at a time, two words ahead
    // (the arena is padded past the last symbol).
    const u64* sp = reinterpret_cast<const u64*>(sym_arena + (size_t)blk.sym_off * 16);
    const uint32_t* pn = pass_nsym + (size_t)bi * T1_MAX_PASSES;
    const uint32_t nsym = res.nsym;
    int pass = 0;
    uint32_t pend = pn[0];
    u64 cur = __ldg(sp), nxt = __ldg(sp + 1);
    uint32_t sym = (uint32_t)cur & 0xffu, ctx = (sym >> 1) & 31u;
    uint32_t e = s_ctx[ctx][t];
    // one decision, mqc.cpp:271-297 without branches; nsymb = the next decision's symbol byte
#define J2K_MQ_STEP(nsymb_expr)                                                                        \
    {                                                                                                  \
        const uint32_t nsymb = (nsymb_expr);                                                           \
        const uint32_t nctx = (nsymb >> 1) & 31u;                                                      \
        uint32_t ne = s_ctx[nctx][t];              /* fetched before this decision's update */          \
        const uint32_t qe = e & 0x7fffu;                                                               \
        const uint32_t lps = (sym ^ (e >> 15)) & 1u;                                                   \
        uint32_t a = q.a - qe;                                                                         \
        const bool lt = a < qe;                                                                        \
        const bool ren = lps || !(a & 0x8000u);    /* every path that renormalises also moves the state */ \
        const bool xch = (lps != 0) != lt;         /* interval becomes qe: MPS with a < qe, LPS with a >= qe */ \
        const uint32_t newe = s_mq[lps ? (e >> 24) : ((e >> 16) & 0xffu)];                             \
        a = xch ? qe : a;                                                                              \
        q.c += xch ? 0u : qe;                                                                          \
        if (ren) s_ctx[ctx][t] = newe;                                                                 \
        if (ren && nctx == ctx) ne = newe;                                                             \
        uint32_t n = __clz(a) - 16;                /* mqc.cpp:259-269, shift count taken at once */      \
        q.a = a << n;                                                                                  \
        if (n >= q.ct) {                                                                               \
            do { q.c <<= q.ct; n -= q.ct; q.byteout(); } while (n >= q.ct);                            \
        }                                                                                              \
        q.c <<= n; q.ct -= n;                                                                          \
        sym = nsymb; ctx = nctx; e = ne;                                                               \
    }
    for (uint32_t base = 0; base < nsym; base += 8) {
        const u64 w0 = cur, w1 = nxt;
        cur = nxt;
        nxt = __ldg(sp + (base >> 3) + 2);
        if (base + 8 <= nsym && pend - base >= 8) {
            // whole word inside one pass: no per-decision bookkeeping
#pragma unroll
            for (int k = 0; k < 8; ++k) J2K_MQ_STEP(k < 7 ? (uint32_t)(w0 >> (8 * (k + 1))) & 0xffu : (uint32_t)w1 & 0xffu)
        } else {
#pragma unroll 1
            for (int k = 0; k < 8; ++k) {
                const uint32_t pos = base + k;
                if (pos >= nsym) break;
                while (pass < npasses && pos == pend) {         // t1_opt.cpp:719
                    res.rate[pass] = (q.nb < 0 ? 0 : q.nb) + 3;
                    if (++pass < npasses) pend += pn[pass];
                }
                J2K_MQ_STEP(k < 7 ? (uint32_t)(w0 >> (8 * (k + 1))) & 0xffu : (uint32_t)w1 & 0xffu)
            }
        }
    }
#undef J2K_MQ_STEP
    while (pass < npasses) {
        res.rate[pass] = (q.nb < 0 ? 0 : q.nb) + 3;
        ++pass;
    }
    // flush, mqc.cpp:299-306, 441-452
    {
        const uint32_t tempc = q.c + q.a;
        q.c |= 0xffffu;
        if (q.c >= tempc) q.c -= 0x8000u;
        q.c <<= q.ct; q.byteout();
        q.c <<= q.ct; q.byteout();
    }
    if (q.nb >= 0) { if ((uint32_t)q.nb < q.cap) q.out[q.nb] = (uint8_t)q.last; else q.overflow = 1; }
    int total = q.nb < 0 ? 0 : q.nb;
    if (q.last != 0xffu) ++total;
    if (q.overflow) { res.status = 2; res.total = 0; return; }
    int lastrate = 0;
    for (int pass = 0; pass < npasses; ++pass) {               // t1_opt.cpp:727-736
        int r = res.rate[pass];
        if (r > total) r = total;
        if (r > 1 && q.out[r - 1] == 0xffu) --r;
        res.rate[pass] = r;
        lastrate = r;
    }
    res.total = lastrate;
}

// ---- MQ coder for the code-block styles RESET / TERMALL (row a21): the reference's generic loop, t1_mq_gen.cuh ------------
// One block per thread, no ordering: streams with these styles are rare, and the routine is the byte-pointer form that
// reproduces the reference's termination arithmetic rather than the branch-free form above.
__global__ void __launch_bounds__(MQ_THREADS)
k_t1_mq_gen(const T1Block* __restrict__ blocks, int nblocks, const T1Tables* __restrict__ tab, const uint8_t* __restrict__ sym_arena,
            const uint32_t* __restrict__ pass_nsym, uint8_t* __restrict__ out_arena, T1Result* __restrict__ results, int cblksty) {
    __shared__ uint32_t s_mq[96];
    if (threadIdx.x < 94) s_mq[threadIdx.x] = tab->mq[threadIdx.x];
    __syncthreads();
    const int bi = blockIdx.x * MQ_THREADS + threadIdx.x;
    if (bi >= nblocks) return;
    T1Result& res = results[bi];
    const int npasses = res.npasses;
    if (npasses == 0 || res.status != 0) { res.total = 0; return; }
    const T1Block blk = blocks[bi];
    int32_t total = 0;
    const int rc = mqgen::encode_block(sym_arena + (size_t)blk.sym_off * 16, pass_nsym + (size_t)bi * T1_MAX_PASSES, npasses, cblksty, s_mq,
                                       out_arena + blk.out_off, (int)blk.out_cap, res.rate, &total);
    res.total = total;
    if (rc) res.status = rc;
}

// ---- compaction: exclusive scan of the kept byte counts, then one warp copies one block --------
__global__ void __launch_bounds__(1024) k_t1_scan(const T1Result* __restrict__ results, int nblocks, uint32_t* __restrict__ offsets) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < nblocks; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < nblocks ? (uint32_t)results[i].total : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += t; }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane], wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(FULL, wi, d); if (lane >= d) wi += t; }
            s_warp[lane] = wi - w;
        }
        __syncthreads();
        const uint32_t excl = s_carry + s_warp[warp] + incl - v;
        if (i < nblocks) offsets[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) offsets[nblocks] = s_carry;
}

__global__ void __launch_bounds__(256) k_t1_gather(const T1Block* __restrict__ blocks, int nblocks, const uint8_t* __restrict__ out_arena,
                                                  const T1Result* __restrict__ results, const uint32_t* __restrict__ offsets,
                                                  uint8_t* __restrict__ packed) {
    const int bi = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (bi >= nblocks) return;
    const int n = results[bi].total;
    const uint8_t* src = out_arena + blocks[bi].out_off;
    uint8_t* dst = packed + offsets[bi];
    for (int i = lane; i < n; i += 32) dst[i] = src[i];
}

}  // namespace

void launch_t1_context(const T1Block* d_blocks, int nblocks, uint8_t* sym_arena, uint32_t* pass_nsym,
                       T1Result* results, cudaStream_t st) {
    if (nblocks <= 0) return;
    // resident CTAs per SM the register allocation aims at (4: 126 registers, 5: 96, 6: 80 with a few spills).  Measured on B200, 4K RGBA16:
    // 1.57 / 1.49 / 1.46 ms alone, 2276 / 2236 / 2413 MPix/s encode with 16 frames in flight: 6 is the default; J2KCUDA_CX_MINB to compare
    static const int minb = [] { const char* e = getenv("J2KCUDA_CX_MINB"); const int v = e ? atoi(e) : 6; return v < 4 ? 4 : v > 6 ? 6 : v; }();
    const dim3 grid((nblocks + CX_WARPS - 1) / CX_WARPS), block(32 * CX_WARPS);
    if (minb == 4) k_t1_context<4><<<grid, block, 0, st>>>(d_blocks, nblocks, t1_tables_device(), sym_arena, pass_nsym, results);
    else if (minb == 5) k_t1_context<5><<<grid, block, 0, st>>>(d_blocks, nblocks, t1_tables_device(), sym_arena, pass_nsym, results);
    else k_t1_context<6><<<grid, block, 0, st>>>(d_blocks, nblocks, t1_tables_device(), sym_arena, pass_nsym, results);
}

void launch_t1_mq(const T1Block* d_blocks, int nblocks, const uint8_t* sym_arena, const uint32_t* pass_nsym,
                  uint8_t* out_arena, T1Result* results, int lanes_per_warp, uint32_t* d_order, int max_long, cudaStream_t st, int cblksty) {
    if (nblocks <= 0) return;
    if (cblksty) {                                   // RESET / TERMALL: the generic coder (the caller has refused every other bit)
        k_t1_mq_gen<<<(nblocks + MQ_THREADS - 1) / MQ_THREADS, MQ_THREADS, 0, st>>>(d_blocks, nblocks, t1_tables_device(), sym_arena, pass_nsym,
                                                                                  out_arena, results, cblksty);
        return;
    }
    if (lanes_per_warp < 1) lanes_per_warp = 1;
    if (lanes_per_warp > 32) lanes_per_warp = 32;
    if (max_long < 0 || !d_order) max_long = 0;
    if (max_long > nblocks) max_long = nblocks;
    if (d_order) k_t1_order<<<1, 1024, 0, st>>>(results, nblocks, max_long, lanes_per_warp, d_order);
    // the grid covers the worst case (max_long single-block warps); warps past the real count exit at once
    const int warps = max_long + (nblocks + lanes_per_warp - 1) / lanes_per_warp;
    const int wpc = MQ_THREADS / 32;
    k_t1_mq<<<(warps + wpc - 1) / wpc, MQ_THREADS, 0, st>>>(d_blocks, nblocks, t1_tables_device(), sym_arena, pass_nsym, out_arena, results,
                                                         lanes_per_warp, d_order);
}

void launch_t1_compact(const T1Block* d_blocks, int nblocks, const uint8_t* out_arena, const T1Result* results,
                       uint32_t* d_offsets, uint8_t* packed, cudaStream_t st) {
    if (nblocks <= 0) return;
    k_t1_scan<<<1, 1024, 0, st>>>(results, nblocks, d_offsets);
    k_t1_gather<<<(nblocks + 7) / 8, 256, 0, st>>>(d_blocks, nblocks, out_arena, results, d_offsets, packed);
}

}  // namespace j2k
