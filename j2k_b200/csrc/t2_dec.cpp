// Host Tier-2 reader: main header, tile-part headers, packet headers -> per-code-block segment lists.
// Reference behaviour: ext/grok/src/lib/openjp2/j2k.cpp (marker readers :2032-4340, tile-part state
// machine :7371-7640), t2.cpp:399-550, 1075-1343 (packets), tgt.cpp:316-352, bio.cpp:99-235.
#include "t2.h"
#include <algorithm>
#include <cstring>

namespace j2k {
namespace {

struct Rd {
    const uint8_t* d; size_t n; size_t pos;
    bool bad = false;                                  // a read past the end: returns 0, the caller reports the marker as damaged
    bool ok(size_t k) const { return k <= n && pos <= n - k; }
    uint32_t u(int bytes) {
        if (!ok((size_t)bytes)) { bad = true; pos = n; return 0; }
        uint32_t v = 0; for (int i = 0; i < bytes; ++i) v = (v << 8) | d[pos++]; return v;
    }
};

bool read_spcod(Rd& r, size_t end, CompParams& t, bool has_prec) {     // j2k.cpp:8372-8470
    if (r.pos + 5 > end) return false;
    t.numres = (int)r.u(1) + 1;
    t.cblkw = (int)r.u(1) + 2; t.cblkh = (int)r.u(1) + 2;
    t.cblksty = (int)r.u(1); t.qmfbid = (int)r.u(1);
    for (int i = 0; i < MAX_RES; ++i) t.prcw[i] = t.prch[i] = 15;
    if (has_prec)
        for (int i = 0; i < t.numres && i < MAX_RES; ++i) {
            if (r.pos >= end) return false;
            uint32_t v = r.u(1); t.prcw[i] = v & 15; t.prch[i] = v >> 4;
            if (i > 0 && (t.prcw[i] == 0 || t.prch[i] == 0)) return false;     // "Invalid precinct size", j2k.cpp:8445
        }
    return true;
}
bool read_sqcd(Rd& r, size_t end, CompParams& t) {                     // j2k.cpp:8662-8760
    if (r.pos + 1 > end) return false;
    uint32_t v = r.u(1);
    t.qntsty = v & 31; t.numgbits = v >> 5;
    if (t.qntsty > 2) return false;
    int nb = t.qntsty == 1 ? 1 : (t.qntsty == 0 ? (int)(end - r.pos) : (int)(end - r.pos) / 2);
    if (nb < 1 || (t.qntsty == 1 && r.pos + 2 > end)) return false;
    for (int b = 0; b < nb && b < 3 * MAX_RES; ++b) {
        if (t.qntsty == 0) { t.steps[b].expn = (int)(r.u(1) >> 3); t.steps[b].mant = 0; }
        else { uint32_t s = r.u(2); t.steps[b].expn = (int)(s >> 11); t.steps[b].mant = (int)(s & 0x7ff); }
    }
    if (t.qntsty == 1)                                                   // derived: :8735-8750
        for (int b = 1; b < 3 * MAX_RES; ++b) {
            int e = t.steps[0].expn - (b - 1) / 3;
            t.steps[b].expn = e > 0 ? e : 0; t.steps[b].mant = t.steps[0].mant;
        }
    r.pos = end;
    return true;
}

int getnumpasses(BitReader& bio) {                                      // t2.cpp:225-239
    if (!bio.getbit()) return 1;
    if (!bio.getbit()) return 2;
    uint32_t n = bio.read(2);
    if (n != 3) return 3 + (int)n;
    n = bio.read(5);
    if (n != 31) return 6 + (int)n;
    return 37 + (int)bio.read(7);
}

}  // namespace

bool find_codestream(const uint8_t* d, size_t n, size_t& off, size_t& len, bool& is_jp2) {
    is_jp2 = false;
    if (n >= 4 && d[0] == 0xff && d[1] == 0x4f && d[2] == 0xff && d[3] == 0x51) { off = 0; len = n; return true; }
    static const uint8_t sig[12] = {0, 0, 0, 12, 'j', 'P', ' ', ' ', 0x0d, 0x0a, 0x87, 0x0a};
    if (n < 12 || memcmp(d, sig, 12) != 0) return false;
    is_jp2 = true;
    size_t pos = 0;
    while (pos + 8 <= n) {                                               // jp2.cpp box walk
        uint64_t lbox = ((uint64_t)d[pos] << 24) | (d[pos + 1] << 16) | (d[pos + 2] << 8) | d[pos + 3];
        const uint8_t* t = d + pos + 4;
        size_t hdr = 8;
        if (lbox == 1) {
            if (pos + 16 > n) return false;
            lbox = 0;
            for (int i = 0; i < 8; ++i) lbox = (lbox << 8) | d[pos + 8 + i];
            hdr = 16;
        } else if (lbox == 0) lbox = n - pos;
        if (lbox < hdr || pos + lbox > n) return false;
        if (t[0] == 'j' && t[1] == 'p' && t[2] == '2' && t[3] == 'c') { off = pos + hdr; len = (size_t)lbox - hdr; return true; }
        pos += (size_t)lbox;
    }
    return false;
}

bool DecoderHost::parse_main_header(const uint8_t* d, size_t n, size_t& pos, std::string& err) {
    Rd r{d, n, pos};
    if (!r.ok(2) || r.u(2) != 0xff4f) { err = "SOC marker missing"; return false; }
    bool have_siz = false, have_cod = false, have_qcd = false;
    bool coc_seen[MAX_COMPS] = {false}, qcc_seen[MAX_COMPS] = {false};
    cp = CodingParams();
    for (;;) {
        if (!r.ok(4)) { err = "truncated main header"; return false; }
        uint32_t m = r.u(2);
        if (m == 0xff90) { r.pos -= 2; break; }                          // SOT
        uint32_t L = r.u(2);
        if (L < 2 || !r.ok(L - 2)) { err = "bad marker segment length"; return false; }
        size_t end = r.pos + L - 2;
        switch (m) {
        case 0xff51: {                                                   // SIZ, j2k.cpp:2032-2262
            if (L < 38) { err = "bad SIZ"; return false; }
            cp.rsiz = (int)r.u(2);
            cp.x1 = r.u(4); cp.y1 = r.u(4); cp.x0 = r.u(4); cp.y0 = r.u(4);
            cp.tdx = r.u(4); cp.tdy = r.u(4); cp.tx0 = r.u(4); cp.ty0 = r.u(4);
            uint32_t nc = r.u(2);
            if (nc < 1 || nc > MAX_COMPS) { err = "unsupported number of components"; return false; }
            if (L != 38 + 3 * nc || !cp.tdx || !cp.tdy || cp.x1 <= cp.x0 || cp.y1 <= cp.y0 || cp.tx0 > cp.x0 || cp.ty0 > cp.y0) { err = "bad SIZ"; return false; }
            cp.ncomp = (int)nc;
            for (uint32_t c = 0; c < nc; ++c) {
                uint32_t s = r.u(1);
                cp.comps[c].prec = (int)(s & 0x7f) + 1; cp.comps[c].sgnd = (int)(s >> 7);
                cp.comps[c].dx = (int)r.u(1); cp.comps[c].dy = (int)r.u(1);
                if (!cp.comps[c].dx || !cp.comps[c].dy || cp.comps[c].prec > 31) { err = "bad SIZ component"; return false; }
                cp.comps[c].dc_shift = cp.comps[c].sgnd ? 0 : 1 << (cp.comps[c].prec - 1);
            }
            // the first tile must meet the image (j2k.cpp:2098-2106), and the sizes must stay allocatable
            if ((uint64_t)cp.tx0 + cp.tdx <= cp.x0 || (uint64_t)cp.ty0 + cp.tdy <= cp.y0) { err = "bad SIZ: the first tile lies outside the image"; return false; }
            if (cp.x1 - cp.x0 > (1u << 24) || cp.y1 - cp.y0 > (1u << 24) || (uint64_t)(cp.x1 - cp.x0) * (cp.y1 - cp.y0) > (1ull << 30)) { err = "image larger than 2^30 samples per component"; return false; }
            cp.tw = ceildiv(cp.x1 - cp.tx0, cp.tdx); cp.th = ceildiv(cp.y1 - cp.ty0, cp.tdy);
            if ((uint64_t)cp.tw * cp.th > 65535) { err = "too many tiles"; return false; }
            have_siz = true;
            break;
        }
        case 0xff52: {                                                   // COD, j2k.cpp:2422-2507
            if (!have_siz) { err = "COD before SIZ"; return false; }
            if (L < 12) { err = "bad COD"; return false; }
            cp.csty = (int)r.u(1); cp.prog = (int)r.u(1); cp.numlayers = (int)r.u(2); cp.mct = (int)r.u(1);
            CompParams t = cp.comps[0];
            if (!read_spcod(r, end, t, cp.csty & 1)) { err = "bad COD"; return false; }
            if (cp.numlayers < 1) { err = "bad COD: no layers"; return false; }
            for (int c = 0; c < cp.ncomp; ++c) {
                if (coc_seen[c]) continue;
                CompParams& o = cp.comps[c];
                o.numres = t.numres; o.cblkw = t.cblkw; o.cblkh = t.cblkh; o.cblksty = t.cblksty; o.qmfbid = t.qmfbid;
                o.csty = cp.csty & 1;
                memcpy(o.prcw, t.prcw, sizeof(t.prcw)); memcpy(o.prch, t.prch, sizeof(t.prch));
            }
            have_cod = true;
            break;
        }
        case 0xff53: {                                                   // COC, j2k.cpp:2618-2698
            if (!have_siz) { err = "COC before SIZ"; return false; }
            if (L < 9) { err = "bad COC"; return false; }
            uint32_t c = r.u(cp.ncomp <= 256 ? 1 : 2);
            if ((int)c >= cp.ncomp) { err = "bad COC"; return false; }
            cp.comps[c].csty = (int)r.u(1) & 1;
            if (!read_spcod(r, end, cp.comps[c], cp.comps[c].csty & 1)) { err = "bad COC"; return false; }
            coc_seen[c] = true;
            break;
        }
        case 0xff5c: {                                                   // QCD
            if (!have_siz) { err = "QCD before SIZ"; return false; }
            CompParams t;
            if (L < 4 || !read_sqcd(r, end, t)) { err = "bad QCD"; return false; }
            for (int c = 0; c < cp.ncomp; ++c) {
                if (qcc_seen[c]) continue;
                cp.comps[c].qntsty = t.qntsty; cp.comps[c].numgbits = t.numgbits;
                memcpy(cp.comps[c].steps, t.steps, sizeof(t.steps));
            }
            have_qcd = true;
            break;
        }
        case 0xff5d: {                                                   // QCC
            if (!have_siz) { err = "QCC before SIZ"; return false; }
            uint32_t c = r.u(cp.ncomp <= 256 ? 1 : 2);
            if ((int)c >= cp.ncomp) { err = "bad QCC"; return false; }
            if (!read_sqcd(r, end, cp.comps[c])) { err = "bad QCC"; return false; }
            qcc_seen[c] = true;
            break;
        }
        case 0xff5f: {                                                   // POC, j2k.cpp:3071-3170
            cp.pocs.clear();
            while (r.pos + 7 <= end) {
                PocEntry pe;
                pe.resno0 = (int)r.u(1); pe.compno0 = (int)r.u(1); pe.layno1 = (int)r.u(2);
                pe.resno1 = (int)r.u(1); pe.compno1 = (int)r.u(1); pe.prg = (int)r.u(1);
                cp.pocs.push_back(pe);
            }
            break;
        }
        case 0xff5e: {                                                   // RGN, j2k.cpp:3998-4075: implicit ROI up-shift per component
            if (!have_siz) { err = "RGN before SIZ"; return false; }
            if (L < 5) { err = "bad RGN"; return false; }
            const uint32_t c = r.u(cp.ncomp <= 256 ? 1 : 2), style = r.u(1), shift = r.u(1);
            if ((int)c >= cp.ncomp || style != 0 || shift > 30) { err = "bad RGN"; return false; }
            cp.comps[c].roishift = (int)shift;
            break;
        }
        default: break;                                                  // TLM, PLM, PPM(!), CRG, COM ...
        }
        if (m == 0xff60) { err = "packed packet headers (PPM) are not supported"; return false; }
        if (r.bad) { err = "marker segment shorter than its contents"; return false; }
        r.pos = end;
    }
    if (!have_siz || !have_cod || !have_qcd) { err = "main header lacks SIZ, COD or QCD"; return false; }
    for (int c = 0; c < cp.ncomp; ++c) {
        const CompParams& t = cp.comps[c];
        if (t.numres < 1 || t.numres > MAX_RES || t.cblkw > 10 || t.cblkh > 10 || t.cblkw + t.cblkh > 12) { err = "bad coding style"; return false; }
        if (t.cblksty & ~0x3f) { err = "unknown code-block style bits"; return false; }
    }
    pos = r.pos;
    return true;
}

static void init_dec_tile(const CodingParams& cp, int tileno, DecTile& t) {
    build_tile(cp, tileno, 0.5f, t.geo);                                 // tcd.cpp:888: decoder step sizes are halved
    t.blocks.assign(t.geo.cblks.size(), DecBlock());
    t.prec_base.clear(); t.incl.clear(); t.imsb.clear();
    for (int c = 0; c < cp.ncomp; ++c)
        for (int r = 0; r < cp.comps[c].numres; ++r) {
            const ResGeo& rg = t.geo.comps[c].res[r];
            t.prec_base.push_back((int)t.incl.size());
            for (int b = 0; b < rg.numbands; ++b)
                for (auto& pg : rg.bands[b].precs) {
                    TagTree a, m;
                    a.init(pg.cw, pg.ch); m.init(pg.cw, pg.ch);
                    t.incl.push_back(std::move(a)); t.imsb.push_back(std::move(m));
                }
        }
}

static int dec_tree_index(const CodingParams& cp, const DecTile& t, int c, int r, int b, int p) {
    int k = 0;
    for (int cc = 0; cc < c; ++cc) k += cp.comps[cc].numres;
    const ResGeo& rg = t.geo.comps[c].res[r];
    return t.prec_base[k + r] + b * (rg.pw * rg.ph) + p;
}

// t2.cpp:1075-1343 (header) and :1345-1560 (body): one packet at d[pos..end)
static bool read_packet(const CodingParams& cp, DecTile& t, const PacketId& pk, const uint8_t* d, size_t& pos, size_t end,
                        std::string& err) {
    const ResGeo& res = t.geo.comps[pk.compno].res[pk.resno];
    if (cp.csty & 2) {                                                   // SOP, optional in the stream
        if (pos + 6 <= end && d[pos] == 0xff && d[pos + 1] == 0x91) pos += 6;
    }
    BitReader bio;
    bio.init(d + pos, end - pos);
    struct Inc { int blk; uint32_t len; };
    Inc inc[4096];
    int ninc = 0;
    std::vector<Inc> big;
    const uint32_t present = bio.getbit();
    if (present) {
        for (int b = 0; b < res.numbands; ++b) {
            const BandGeo& band = res.bands[b];
            if (band.r.w() <= 0 || band.r.h() <= 0) continue;
            const PrecGeo& prc = band.precs[pk.precno];
            const int ti = dec_tree_index(cp, t, pk.compno, pk.resno, b, pk.precno);
            const int nb = prc.cw * prc.ch;
            for (int k = 0; k < nb; ++k) {
                DecBlock& cb = t.blocks[prc.first_cblk + k];
                uint32_t included;
                if (!cb.included) included = (uint32_t)t.incl[ti].decode(bio, k, pk.layno + 1);
                else included = bio.getbit();
                if (!included) continue;
                if (!cb.included) {
                    // The reference asks the tag tree "is the value below i?" for i = 1, 2, ... (t2.cpp:1226-1236), walking the tree
                    // once per question.  A node's bits are only read once all its ancestors are known, so one walk with the largest
                    // threshold the loop would reach reads the very same bits in the same order and leaves the value behind.
                    t.imsb[ti].decode(bio, k, 76);
                    const int zbp = t.imsb[ti].nodes[k].value;
                    if (zbp >= 75 || bio.overrun) { err = "bad zero bit-plane count"; return false; }
                    const int i = zbp + 1;
                    cb.numbps = band.numbps + 1 - i;
                    cb.numlenbits = 3;
                    cb.included = true;
                }
                const int np = getnumpasses(bio);
                int incr = 0;
                while (bio.getbit()) { if (++incr > 32) { err = "bad Lblock"; return false; } }
                cb.numlenbits += incr;
                // codeword segments (t2.cpp:1262-1295, 1525-1563): TERMALL ends one after every pass, LAZY after the first ten
                // passes and then after every two / one; otherwise the block is one segment
                const int sty = cp.comps[pk.compno].cblksty;
                auto new_seg = [&](bool first) {
                    DecSeg sg;
                    if (sty & 4) sg.maxpasses = 1;
                    else if (sty & 1) sg.maxpasses = first ? 10 : ((cb.segs.back().maxpasses == 1 || cb.segs.back().maxpasses == 10) ? 2 : 1);
                    cb.segs.push_back(sg);
                };
                if (cb.segs.empty()) new_seg(true);
                else if (cb.segs.back().numpasses == cb.segs.back().maxpasses) new_seg(false);
                uint32_t len = 0;
                for (int n = np; n > 0;) {
                    DecSeg& sg = cb.segs.back();
                    const int take = std::min(sg.maxpasses - sg.numpasses, n);
                    const int nbits = cb.numlenbits + floorlog2((uint32_t)take);
                    if (nbits > 31) { err = "bad segment length field"; return false; }     // a 32-bit length used to wrap to -1
                    const uint32_t sl = bio.read(nbits);
                    if (sl > end - pos || len + sl > end - pos) { err = "code-block segment longer than the tile-part"; return false; }
                    sg.numpasses += take; sg.len += (int)sl;
                    len += sl;
                    n -= take;
                    if (n > 0) { if (cb.segs.size() > 200) { err = "too many codeword segments"; return false; } new_seg(false); }
                }
                cb.npasses += np;
                if (ninc < 4096) inc[ninc++] = Inc{prc.first_cblk + k, len};
                else big.push_back(Inc{prc.first_cblk + k, len});
            }
        }
    }
    bio.inalign();
    if (bio.overrun) { err = "packet header runs past the tile-part"; return false; }
    pos = (size_t)(bio.bp - d);
    if (cp.csty & 4) {                                                   // EPH
        if (pos + 2 <= end && d[pos] == 0xff && d[pos + 1] == 0x92) pos += 2;
    }
    auto body = [&](const Inc& in) {
        if (pos > end || (size_t)in.len > end - pos) return false;
        DecBlock& cb = t.blocks[in.blk];
        if ((uint64_t)cb.len + in.len > (1u << 30)) return false;
        if (in.len) cb.chunks.emplace_back((uint32_t)pos, in.len);
        cb.len += (int)in.len;
        pos += (size_t)in.len;
        return true;
    };
    for (int i = 0; i < ninc; ++i) if (!body(inc[i])) { err = "packet body runs past the tile-part"; return false; }
    for (auto& in : big) if (!body(in)) { err = "packet body runs past the tile-part"; return false; }
    return true;
}

static bool same_geometry(const CodingParams& a, const CodingParams& b) {
    return a.x0 == b.x0 && a.y0 == b.y0 && a.x1 == b.x1 && a.y1 == b.y1 && a.tx0 == b.tx0 && a.ty0 == b.ty0 && a.tdx == b.tdx &&
           a.tdy == b.tdy && a.ncomp == b.ncomp && a.csty == b.csty && a.numlayers == b.numlayers &&
           memcmp(a.comps, b.comps, sizeof(a.comps)) == 0;
}

bool DecoderHost::parse(const uint8_t* d, size_t n, std::string& err) {
    size_t pos = 0;
    have_prev = have_prev && !tiles.empty();
    if (!parse_main_header(d, n, pos, err)) { have_prev = false; return false; }
    if (!geometry_within_limits(cp, err)) { have_prev = false; return false; }
    const int nt = (int)(cp.tw * cp.th);
    const bool reuse = have_prev && (int)tiles.size() == nt && prev_first == tile_first && prev_count == tile_count && same_geometry(prev_cp, cp);
    have_prev = false;                               // set again only when this parse succeeds
    if (!reuse) tiles.assign(nt, DecTile());
    else
        for (auto& t : tiles) {
            t.parts.clear();
            for (auto& b : t.blocks) { b.numbps = 0; b.npasses = 0; b.numlenbits = 3; b.included = false; b.chunks.clear(); b.len = 0; b.segs.clear(); }
            for (auto& tr : t.incl) tr.reset();
            for (auto& tr : t.imsb) tr.reset();
        }
    std::vector<std::vector<PocEntry>> tile_pocs(nt);
    // ---- tile-parts: j2k.cpp:7371-7640 ----
    for (;;) {
        if (pos + 2 > n) break;                                          // tolerate a missing EOC
        uint32_t m = ((uint32_t)d[pos] << 8) | d[pos + 1];
        if (m == 0xffd9) break;
        if (m != 0xff90) { err = "expected SOT or EOC"; return false; }
        if (pos + 12 > n) { err = "truncated SOT"; return false; }
        const size_t sot = pos;
        Rd r{d, n, pos + 2};
        if (r.u(2) != 10) { err = "bad SOT"; return false; }
        const uint32_t tileno = r.u(2);
        uint32_t psot = r.u(4);
        r.u(1); r.u(1);
        if ((int)tileno >= nt) { err = "tile index out of range"; return false; }
        size_t part_end = psot ? sot + psot : n - ((n >= 2 && d[n - 2] == 0xff && d[n - 1] == 0xd9) ? 2 : 0);
        if (part_end > n || part_end < r.pos) { err = "tile-part length runs past the codestream"; return false; }
        for (;;) {                                                       // tile-part header up to SOD
            if (r.pos + 2 > part_end) { err = "SOD missing"; return false; }
            uint32_t mk = r.u(2);
            if (mk == 0xff93) break;
            if (r.pos + 2 > part_end) { err = "truncated tile-part header"; return false; }
            uint32_t L = r.u(2);
            if (L < 2 || r.pos + L - 2 > part_end) { err = "bad tile-part marker length"; return false; }
            size_t end = r.pos + L - 2;
            if (mk == 0xff5f) {
                while (r.pos + 7 <= end) {
                    PocEntry pe;
                    pe.resno0 = (int)r.u(1); pe.compno0 = (int)r.u(1); pe.layno1 = (int)r.u(2);
                    pe.resno1 = (int)r.u(1); pe.compno1 = (int)r.u(1); pe.prg = (int)r.u(1);
                    tile_pocs[tileno].push_back(pe);
                }
            } else if (mk == 0xff52 || mk == 0xff53 || mk == 0xff5c || mk == 0xff5d) {
                err = "tile-specific COD/COC/QCD/QCC are not supported"; return false;
            } else if (mk == 0xff61) { err = "packed packet headers (PPT) are not supported"; return false; }
            r.pos = end;
        }
        tiles[tileno].parts.emplace_back(r.pos, part_end - r.pos);
        pos = part_end;
    }
    // ---- packets ----
    int maxres = 0;
    for (int c = 0; c < cp.ncomp; ++c) maxres = std::max(maxres, cp.comps[c].numres);
    for (int tno = 0; tno < nt; ++tno) {
        DecTile& t = tiles[tno];
        if (!tile_in_range(tno)) { t.parts.clear(); continue; }     // another GPU's tile: no geometry, no packets
        if (!reuse) init_dec_tile(cp, tno, t);
        if (t.parts.empty()) continue;
        std::vector<PocEntry> pocs = tile_pocs[tno].empty() ? cp.pocs : tile_pocs[tno];
        if (pocs.empty()) pocs.push_back(PocEntry{0, 0, cp.numlayers, maxres, cp.ncomp, cp.prog});
        std::vector<uint8_t> include;
        std::vector<PacketId> pk;
        for (auto& pe : pocs)
            progression_packets(cp, t.geo, pe.prg, pe.resno0, std::min(pe.resno1, maxres), pe.compno0, pe.compno1, 0,
                                std::min(pe.layno1, cp.numlayers), include, pk);
        size_t part = 0, p = t.parts[0].first, pend = p + t.parts[0].second;
        for (auto& id : pk) {
            while (p >= pend && part + 1 < t.parts.size()) { ++part; p = t.parts[part].first; pend = p + t.parts[part].second; }
            if (p >= pend) break;                                        // truncated stream: remaining packets absent
            if (!read_packet(cp, t, id, d, p, pend, err)) return false;
        }
    }
    prev_cp = cp; prev_first = tile_first; prev_count = tile_count; have_prev = true;
    return true;
}

}  // namespace j2k
