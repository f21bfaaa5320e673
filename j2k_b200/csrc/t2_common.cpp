// Tag trees and progression order; shared by the Tier-2 writer and reader.  See t2.h.
#include "t2.h"
#include <algorithm>

namespace j2k {

// tgt.cpp:66-142: leaves first (raster), then each coarser level; parent links by 2x2 grouping.
void TagTree::init(int numleafsh, int numleafsv) {
    w = numleafsh; h = numleafsv;
    nodes.clear();
    if (w <= 0 || h <= 0) return;
    int nplh[32], nplv[32], numlvls = 0, total = 0, n;
    nplh[0] = w; nplv[0] = h;
    do {
        n = nplh[numlvls] * nplv[numlvls];
        nplh[numlvls + 1] = (nplh[numlvls] + 1) / 2;
        nplv[numlvls + 1] = (nplv[numlvls] + 1) / 2;
        total += n;
        ++numlvls;
    } while (n > 1);
    nodes.assign(total, Node{-1, 999, 0, 0});
    int base = 0;
    for (int l = 0; l < numlvls - 1; ++l) {
        int next = base + nplh[l] * nplv[l];
        for (int j = 0; j < nplv[l]; ++j)
            for (int i = 0; i < nplh[l]; ++i)
                nodes[base + j * nplh[l] + i].parent = next + (j >> 1) * nplh[l + 1] + (i >> 1);
        base = next;
    }
    reset();
}
void TagTree::reset() { for (auto& nd : nodes) { nd.value = 999; nd.low = 0; nd.known = 0; } }
void TagTree::setvalue(int leaf, int value) {          // tgt.cpp:265-273
    int n = leaf;
    while (n >= 0 && nodes[n].value > value) { nodes[n].value = value; n = nodes[n].parent; }
}
void TagTree::encode(BitWriter& bio, int leaf, int threshold) {   // tgt.cpp:275-314
    int stk[32], sp = 0, n = leaf;
    while (nodes[n].parent >= 0) { stk[sp++] = n; n = nodes[n].parent; }
    int low = 0;
    for (;;) {
        Node& nd = nodes[n];
        if (low > nd.low) nd.low = low; else low = nd.low;
        while (low < threshold) {
            if (low >= nd.value) {
                if (!nd.known) { bio.putbit(1); nd.known = 1; }
                break;
            }
            bio.putbit(0);
            ++low;
        }
        nd.low = low;
        if (sp == 0) break;
        n = stk[--sp];
    }
}
int TagTree::decode(BitReader& bio, int leaf, int threshold) {    // tgt.cpp:316-352
    int stk[32], sp = 0, n = leaf;
    while (nodes[n].parent >= 0) { stk[sp++] = n; n = nodes[n].parent; }
    int low = 0;
    for (;;) {
        Node& nd = nodes[n];
        if (low > nd.low) nd.low = low; else low = nd.low;
        while (low < threshold && low < nd.value) {
            if (bio.getbit()) nd.value = low; else ++low;
        }
        nd.low = low;
        if (sp == 0) break;
        n = stk[--sp];
    }
    return nodes[n].value < threshold ? 1 : 0;
}

// ---- progression orders: pi.cpp:276-567 ------------------------------------------------------
namespace {
struct PiCtx {
    const CodingParams& cp; const TileGeo& t;
    std::vector<uint8_t>& include; std::vector<PacketId>& out;
    int maxres = 0, maxprec = 0;
    size_t idx(int l, int r, int c, int p) const {
        return (((size_t)l * maxres + r) * cp.ncomp + c) * maxprec + p;
    }
    void emit(int l, int r, int c, int p) {
        size_t i = idx(l, r, c, p);
        if (i >= include.size()) include.resize(i + 1, 0);
        if (!include[i]) { include[i] = 1; out.push_back(PacketId{l, r, c, p}); }
    }
};
inline uint64_t cdiv64(uint64_t a, uint64_t b) { return (a + b - 1) / b; }

// shared position test of the RPCL/PCRL/CPRL loops; returns precinct index or -1
int prec_at(const CodingParams& cp, const TileGeo& t, int compno, int resno, uint32_t x, uint32_t y) {
    const CompParams& cpar = cp.comps[compno];
    const ResGeo& res = t.comps[compno].res[resno];
    int levelno = cpar.numres - 1 - resno;
    uint32_t tx0 = (uint32_t)t.r.x0, ty0 = (uint32_t)t.r.y0, tx1 = (uint32_t)t.r.x1, ty1 = (uint32_t)t.r.y1;
    uint32_t trx0 = (uint32_t)cdiv64(tx0, (uint64_t)cpar.dx << levelno);
    uint32_t try0 = (uint32_t)cdiv64(ty0, (uint64_t)cpar.dy << levelno);
    uint32_t trx1 = (uint32_t)cdiv64(tx1, (uint64_t)cpar.dx << levelno);
    uint32_t try1 = (uint32_t)cdiv64(ty1, (uint64_t)cpar.dy << levelno);
    int rpx = res.pdx + levelno, rpy = res.pdy + levelno;
    if (!(((uint64_t)y % ((uint64_t)cpar.dy << rpy) == 0) ||
          ((y == ty0) && (((uint64_t)try0 << levelno) % ((uint64_t)1 << rpy))))) return -1;
    if (!(((uint64_t)x % ((uint64_t)cpar.dx << rpx) == 0) ||
          ((x == tx0) && (((uint64_t)trx0 << levelno) % ((uint64_t)1 << rpx))))) return -1;
    if (res.pw == 0 || res.ph == 0) return -1;
    if (trx0 == trx1 || try0 == try1) return -1;
    uint32_t prci = ((uint32_t)cdiv64(x, (uint64_t)cpar.dx << levelno) >> res.pdx) - (trx0 >> res.pdx);
    uint32_t prcj = ((uint32_t)cdiv64(y, (uint64_t)cpar.dy << levelno) >> res.pdy) - (try0 >> res.pdy);
    return (int)(prci + prcj * res.pw);
}
void min_dxy(const CodingParams& cp, const TileGeo& t, int c0, int c1, uint32_t& dx, uint32_t& dy) {
    dx = dy = 0;
    for (int c = c0; c < c1; ++c) {
        const CompParams& cpar = cp.comps[c];
        for (int r = 0; r < cpar.numres; ++r) {
            const ResGeo& res = t.comps[c].res[r];
            uint64_t ddx = (uint64_t)cpar.dx * ((uint64_t)1 << (res.pdx + cpar.numres - 1 - r));
            uint64_t ddy = (uint64_t)cpar.dy * ((uint64_t)1 << (res.pdy + cpar.numres - 1 - r));
            if (ddx < 0xffffffffull) dx = !dx ? (uint32_t)ddx : std::min(dx, (uint32_t)ddx);
            if (ddy < 0xffffffffull) dy = !dy ? (uint32_t)ddy : std::min(dy, (uint32_t)ddy);
        }
    }
}
}  // namespace

void progression_packets(const CodingParams& cp, const TileGeo& t, int prg, int resno0, int resno1,
                         int compno0, int compno1, int layno0, int layno1,
                         std::vector<uint8_t>& include, std::vector<PacketId>& out) {
    PiCtx pc{cp, t, include, out};
    for (int c = 0; c < cp.ncomp; ++c) {
        pc.maxres = std::max(pc.maxres, cp.comps[c].numres);
        for (auto& r : t.comps[c].res) pc.maxprec = std::max(pc.maxprec, r.pw * r.ph);
    }
    if (pc.maxprec == 0) pc.maxprec = 1;
    compno1 = std::min(compno1, cp.ncomp);
    uint32_t tx0 = (uint32_t)t.r.x0, ty0 = (uint32_t)t.r.y0, tx1 = (uint32_t)t.r.x1, ty1 = (uint32_t)t.r.y1;
    uint32_t dx, dy;
    switch (prg) {
    case 0:  // LRCP
        for (int l = layno0; l < layno1; ++l)
            for (int r = resno0; r < resno1; ++r)
                for (int c = compno0; c < compno1; ++c) {
                    if (r >= cp.comps[c].numres) continue;
                    const ResGeo& res = t.comps[c].res[r];
                    for (int p = 0; p < res.pw * res.ph; ++p) pc.emit(l, r, c, p);
                }
        break;
    case 1:  // RLCP
        for (int r = resno0; r < resno1; ++r)
            for (int l = layno0; l < layno1; ++l)
                for (int c = compno0; c < compno1; ++c) {
                    if (r >= cp.comps[c].numres) continue;
                    const ResGeo& res = t.comps[c].res[r];
                    for (int p = 0; p < res.pw * res.ph; ++p) pc.emit(l, r, c, p);
                }
        break;
    case 2:  // RPCL
        min_dxy(cp, t, 0, cp.ncomp, dx, dy);
        if (!dx || !dy) break;
        for (int r = resno0; r < resno1; ++r)
            for (uint64_t y = ty0; y < ty1; y += dy - (y % dy))
                for (uint64_t x = tx0; x < tx1; x += dx - (x % dx))
                    for (int c = compno0; c < compno1; ++c) {
                        if (r >= cp.comps[c].numres) continue;
                        int p = prec_at(cp, t, c, r, (uint32_t)x, (uint32_t)y);
                        if (p < 0) continue;
                        for (int l = layno0; l < layno1; ++l) pc.emit(l, r, c, p);
                    }
        break;
    case 3:  // PCRL
        min_dxy(cp, t, 0, cp.ncomp, dx, dy);
        if (!dx || !dy) break;
        for (uint64_t y = ty0; y < ty1; y += dy - (y % dy))
            for (uint64_t x = tx0; x < tx1; x += dx - (x % dx))
                for (int c = compno0; c < compno1; ++c)
                    for (int r = resno0; r < std::min(resno1, cp.comps[c].numres); ++r) {
                        int p = prec_at(cp, t, c, r, (uint32_t)x, (uint32_t)y);
                        if (p < 0) continue;
                        for (int l = layno0; l < layno1; ++l) pc.emit(l, r, c, p);
                    }
        break;
    case 4:  // CPRL
        for (int c = compno0; c < compno1; ++c) {
            min_dxy(cp, t, c, c + 1, dx, dy);
            if (!dx || !dy) continue;
            for (uint64_t y = ty0; y < ty1; y += dy - (y % dy))
                for (uint64_t x = tx0; x < tx1; x += dx - (x % dx))
                    for (int r = resno0; r < std::min(resno1, cp.comps[c].numres); ++r) {
                        int p = prec_at(cp, t, c, r, (uint32_t)x, (uint32_t)y);
                        if (p < 0) continue;
                        for (int l = layno0; l < layno1; ++l) pc.emit(l, r, c, p);
                    }
        }
        break;
    default: break;
    }
}

}  // namespace j2k
