// Tile geometry: see geometry.h.  Reference behaviour: ext/grok/src/lib/openjp2/tcd.cpp:508-842
// (bounds), t1.cpp:1478-1501 (Mallat placement of code-blocks), dwt.cpp:401-408,573-593 (steps).
#include "geometry.h"
#include <algorithm>
#include <cmath>

namespace j2k {

static const double kNorms53[4][10] = {      // dwt.cpp:193-198
    {1.000, 1.500, 2.750, 5.375, 10.68, 21.34, 42.67, 85.33, 170.7, 341.3},
    {1.038, 1.592, 2.919, 5.703, 11.33, 22.64, 45.25, 90.48, 180.9, 0},
    {1.038, 1.592, 2.919, 5.703, 11.33, 22.64, 45.25, 90.48, 180.9, 0},
    {.7186, .9218, 1.586, 3.043, 6.019, 12.01, 24.00, 47.97, 95.93, 0}};
static const double kNorms97[4][10] = {      // dwt.cpp:203-208
    {1.000, 1.965, 4.177, 8.403, 16.90, 33.84, 67.69, 135.3, 270.6, 540.9},
    {2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
    {2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
    {2.080, 3.865, 8.307, 17.18, 34.71, 69.59, 139.3, 278.6, 557.2, 0}};

double dwt_norm(int level, int orient, bool real) {
    if (level > 9) level = 9;
    return real ? kNorms97[orient][level] : kNorms53[orient][level];
}
double mct_norm(int compno, bool real) {      // mct.cpp:73,78
    static const double rct[3] = {1.732, .8292, .8292}, ict[3] = {1.732, 1.805, 1.573};
    return real ? ict[compno] : rct[compno];
}
int dwt_gain(int orient, bool real) {         // dwt.cpp:531-538,556-560
    if (real) return 0;
    return orient == 0 ? 0 : (orient == 3 ? 2 : 1);
}

void calc_explicit_stepsizes(CompParams& c) { // dwt.cpp:401-408,573-593
    int numbands = 3 * c.numres - 2;
    for (int b = 0; b < numbands; ++b) {
        int resno = b == 0 ? 0 : (b - 1) / 3 + 1;
        int orient = b == 0 ? 0 : (b - 1) % 3 + 1;
        int level = c.numres - 1 - resno;
        int gain = c.qmfbid == 0 ? 0 : dwt_gain(orient, false);
        double stepsize = 1.0;
        if (c.qntsty != 0) stepsize = (double)(1 << gain) / dwt_norm(level, orient, true);
        int s = (int)std::floor(stepsize * 8192.0);
        int p = floorlog2((uint32_t)s) - 13;
        int n = 11 - floorlog2((uint32_t)s);
        c.steps[b].mant = (n < 0 ? s >> -n : s << n) & 0x7ff;
        c.steps[b].expn = (c.prec + gain) - p;
    }
}

// Upper bounds of what build_tile would allocate, in 64-bit arithmetic and without allocating: a 60-byte header that
// announces 2^30 x 2^30 samples or one-sample precincts must give an error, not std::bad_alloc.
bool geometry_within_limits(const CodingParams& cp, std::string& err) {
    const uint64_t ntiles = (uint64_t)cp.tw * cp.th;
    const uint64_t tw = std::min<uint64_t>(cp.tdx, (uint64_t)cp.x1 - cp.tx0), th = std::min<uint64_t>(cp.tdy, (uint64_t)cp.y1 - cp.ty0);
    uint64_t precs = 0, blocks = 0;
    for (int c = 0; c < cp.ncomp; ++c) {
        const CompParams& tc = cp.comps[c];
        const uint64_t cw = tw / (uint64_t)tc.dx + 1, ch = th / (uint64_t)tc.dy + 1;
        for (int r = 0; r < tc.numres; ++r) {
            const int level = tc.numres - 1 - r;
            const uint64_t rw = (cw >> level) + 1, rh = (ch >> level) + 1;
            int pdx = tc.prcw[r], pdy = tc.prch[r];
            if (r > 0 && (pdx < 1 || pdy < 1)) { err = "invalid precinct size"; return false; }
            const uint64_t pw = (rw >> pdx) + 2, ph = (rh >> pdy) + 2;
            const int cbw = std::min(tc.cblkw, r ? pdx - 1 : pdx), cbh = std::min(tc.cblkh, r ? pdy - 1 : pdy);
            const uint64_t nb = r ? 3 : 1;
            precs += pw * ph * nb;
            blocks += ((rw >> cbw) + pw + 1) * ((rh >> cbh) + ph + 1) * nb;
        }
    }
    if (blocks * ntiles > (1ull << 23) || blocks > (1ull << 22)) { err = "more than 2^23 code-blocks"; return false; }
    if (precs * ntiles * (uint64_t)std::max(cp.numlayers, 1) > (1ull << 27) || precs > (1ull << 22)) { err = "more than 2^27 packets"; return false; }
    return true;
}

void build_tile(const CodingParams& cp, int tileno, float fraction, TileGeo& t) {
    t.tileno = tileno;
    t.comps.clear();
    t.cblks.clear();
    uint32_t p = (uint32_t)tileno % cp.tw, q = (uint32_t)tileno / cp.tw;
    uint32_t tx0 = cp.tx0 + p * cp.tdx, ty0 = cp.ty0 + q * cp.tdy;
    t.r.x0 = (int)std::max(tx0, cp.x0);
    t.r.y0 = (int)std::max(ty0, cp.y0);
    t.r.x1 = (int)std::min<uint64_t>((uint64_t)tx0 + cp.tdx, cp.x1);
    t.r.y1 = (int)std::min<uint64_t>((uint64_t)ty0 + cp.tdy, cp.y1);
    t.comps.resize(cp.ncomp);
    for (int c = 0; c < cp.ncomp; ++c) {
        const CompParams& tc = cp.comps[c];
        CompGeo& cg = t.comps[c];
        cg.r.x0 = (int)ceildiv((uint32_t)t.r.x0, tc.dx);
        cg.r.y0 = (int)ceildiv((uint32_t)t.r.y0, tc.dy);
        cg.r.x1 = (int)ceildiv((uint32_t)t.r.x1, tc.dx);
        cg.r.y1 = (int)ceildiv((uint32_t)t.r.y1, tc.dy);
        cg.numres = tc.numres;
        cg.res.resize(tc.numres);
        int stepidx = 0;
        for (int resno = 0; resno < tc.numres; ++resno) {
            int level = tc.numres - 1 - resno;
            ResGeo& rg = cg.res[resno];
            rg.r.x0 = (int)ceildivpow2((uint32_t)cg.r.x0, level);
            rg.r.y0 = (int)ceildivpow2((uint32_t)cg.r.y0, level);
            rg.r.x1 = (int)ceildivpow2((uint32_t)cg.r.x1, level);
            rg.r.y1 = (int)ceildivpow2((uint32_t)cg.r.y1, level);
            int pdx = tc.prcw[resno], pdy = tc.prch[resno];
            rg.pdx = pdx; rg.pdy = pdy;
            uint32_t tlx = floordivpow2((uint32_t)rg.r.x0, pdx) << pdx;
            uint32_t tly = floordivpow2((uint32_t)rg.r.y0, pdy) << pdy;
            uint32_t brx = ceildivpow2((uint32_t)rg.r.x1, pdx) << pdx;
            uint32_t bry = ceildivpow2((uint32_t)rg.r.y1, pdy) << pdy;
            rg.pw = rg.r.x0 == rg.r.x1 ? 0 : (int)((brx - tlx) >> pdx);
            rg.ph = rg.r.y0 == rg.r.y1 ? 0 : (int)((bry - tly) >> pdy);
            int nprec = (int)std::min<int64_t>((int64_t)rg.pw * rg.ph, 1 << 22);     // geometry_within_limits() ran first
            uint32_t tlcbgx, tlcbgy;
            int cbgw, cbgh;
            if (resno == 0) { tlcbgx = tlx; tlcbgy = tly; cbgw = pdx; cbgh = pdy; rg.numbands = 1; }
            else { tlcbgx = ceildivpow2(tlx, 1); tlcbgy = ceildivpow2(tly, 1); cbgw = pdx - 1; cbgh = pdy - 1; rg.numbands = 3; }
            int cbw = std::min(tc.cblkw, cbgw), cbh = std::min(tc.cblkh, cbgh);
            for (int b = 0; b < rg.numbands; ++b) {
                BandGeo& bg = rg.bands[b];
                bg.precs.clear();
                if (resno == 0) {
                    bg.orient = 0;
                    bg.r = rg.r;
                } else {
                    bg.orient = b + 1;
                    int64_t xob = bg.orient & 1, yob = bg.orient >> 1;
                    bg.r.x0 = (int)ceildivpow2_64((int64_t)cg.r.x0 - (xob << level), level + 1);
                    bg.r.y0 = (int)ceildivpow2_64((int64_t)cg.r.y0 - (yob << level), level + 1);
                    bg.r.x1 = (int)ceildivpow2_64((int64_t)cg.r.x1 - (xob << level), level + 1);
                    bg.r.y1 = (int)ceildivpow2_64((int64_t)cg.r.y1 - (yob << level), level + 1);
                }
                const StepSize& ss = tc.steps[tc.qntsty == 1 ? 0 : stepidx];
                int expn = ss.expn, mant = ss.mant;
                if (tc.qntsty == 1 && stepidx > 0)      // derived: j2k.cpp read_SQcd_SQcc
                    expn = std::max(0, ss.expn - (stepidx - 1) / 3);
                int gain = dwt_gain(bg.orient, tc.qmfbid == 0);
                int numbps = tc.prec + gain;
                bg.stepsize = (float)((1.0 + mant / 2048.0) * std::pow(2.0, (double)(numbps - expn))) * fraction;
                bg.numbps = expn + tc.numgbits - 1;
                bg.offx = bg.offy = 0;
                if (resno > 0) {
                    const ResGeo& pres = cg.res[resno - 1];
                    if (bg.orient & 1) bg.offx = pres.r.w();
                    if (bg.orient & 2) bg.offy = pres.r.h();
                }
                bg.precs.resize(nprec);
                for (int pn = 0; pn < nprec; ++pn) {
                    PrecGeo& pg = bg.precs[pn];
                    uint32_t cbgx0 = tlcbgx + (uint32_t)(pn % rg.pw) * (1u << cbgw);
                    uint32_t cbgy0 = tlcbgy + (uint32_t)(pn / rg.pw) * (1u << cbgh);
                    uint32_t cbgx1 = cbgx0 + (1u << cbgw), cbgy1 = cbgy0 + (1u << cbgh);
                    pg.r.x0 = std::max((int)cbgx0, bg.r.x0);
                    pg.r.y0 = std::max((int)cbgy0, bg.r.y0);
                    pg.r.x1 = std::min((int)cbgx1, bg.r.x1);
                    pg.r.y1 = std::min((int)cbgy1, bg.r.y1);
                    pg.first_cblk = (int)t.cblks.size();
                    // an empty band leaves x1 < x0 possible; the reference's unsigned arithmetic
                    // then yields cw/ch from wrapped values only when the band is non-empty
                    if (pg.r.x1 <= pg.r.x0 || pg.r.y1 <= pg.r.y0) {
                        // tcd.cpp:725-734 with an empty precinct: ceil == floor -> 0 blocks,
                        // unless the empty range straddles a block boundary; clamp to zero as
                        // no sample exists either way.
                        uint32_t a = floordivpow2((uint32_t)std::max(pg.r.x0, 0), cbw) << cbw;
                        uint32_t b2 = ceildivpow2((uint32_t)std::max(pg.r.x1, 0), cbw) << cbw;
                        uint32_t c2 = floordivpow2((uint32_t)std::max(pg.r.y0, 0), cbh) << cbh;
                        uint32_t d2 = ceildivpow2((uint32_t)std::max(pg.r.y1, 0), cbh) << cbh;
                        pg.cw = b2 > a ? (int)((b2 - a) >> cbw) : 0;
                        pg.ch = d2 > c2 ? (int)((d2 - c2) >> cbh) : 0;
                    } else {
                        uint32_t a = floordivpow2((uint32_t)pg.r.x0, cbw) << cbw;
                        uint32_t b2 = ceildivpow2((uint32_t)pg.r.x1, cbw) << cbw;
                        uint32_t c2 = floordivpow2((uint32_t)pg.r.y0, cbh) << cbh;
                        uint32_t d2 = ceildivpow2((uint32_t)pg.r.y1, cbh) << cbh;
                        pg.cw = (int)((b2 - a) >> cbw);
                        pg.ch = (int)((d2 - c2) >> cbh);
                    }
                    uint32_t tlbx = floordivpow2((uint32_t)std::max(pg.r.x0, 0), cbw) << cbw;
                    uint32_t tlby = floordivpow2((uint32_t)std::max(pg.r.y0, 0), cbh) << cbh;
                    for (int k = 0; k < pg.cw * pg.ch; ++k) {
                        CblkGeo cb;
                        int bx0 = (int)tlbx + (k % pg.cw) * (1 << cbw);
                        int by0 = (int)tlby + (k / pg.cw) * (1 << cbh);
                        cb.r.x0 = std::max(bx0, pg.r.x0);
                        cb.r.y0 = std::max(by0, pg.r.y0);
                        cb.r.x1 = std::min(bx0 + (1 << cbw), pg.r.x1);
                        cb.r.y1 = std::min(by0 + (1 << cbh), pg.r.y1);
                        if (cb.r.x1 < cb.r.x0) cb.r.x1 = cb.r.x0;
                        if (cb.r.y1 < cb.r.y0) cb.r.y1 = cb.r.y0;
                        cb.comp = c; cb.res = resno; cb.band = b; cb.orient = bg.orient; cb.prec = pn;
                        cb.px = cb.r.x0 - bg.r.x0 + bg.offx;
                        cb.py = cb.r.y0 - bg.r.y0 + bg.offy;
                        t.cblks.push_back(cb);
                    }
                }
                ++stepidx;
            }
        }
    }
}

}  // namespace j2k
