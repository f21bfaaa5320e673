// Host Tier-2 writer: coding-parameter setup, layer formation (PCRD), packet headers, markers.
// Reference behaviour: ext/grok/src/lib/openjp2/j2k.cpp (setup_encoder :5842-6216, cinema
// :5623-5778, update_rates :4355-4482, markers :1871-4340, tile parts :9983-10336),
// tcd.cpp:165-417 (pcrd_bisect / makelayer), t2.cpp:241-384,614-829 (packets).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include "t2.h"
#include "../../include/j2kcuda.h"
#include <algorithm>
#include <cfloat>
#include <cmath>

namespace j2k {

static const char* kCommentGrok = "Created by Grok     version 2.1.1";
static const char* kCommentOpj25 = "Created by OpenJPEG version 2.5.4";
static const char* kCommentOpj253 = "Created by OpenJPEG version 2.5.3";

// What the reference refuses before it encodes (opj_compress parameter checks and opj_j2k_encoding_validation,
// j2k.cpp:6640-6680), plus the limits of this library.  nullptr = fine.
const char* check_params(const ::j2kcuda_params& p) {
    if (!p.width || !p.height) return "empty image";
    if (p.ncomp < 1 || p.ncomp > 4) return "unsupported image geometry (1-4 components, 1-16 bits)";
    if (p.prec < 1 || p.prec > 16) return "unsupported image geometry (1-4 components, 1-16 bits)";
    uint32_t numres = p.numresolution ? p.numresolution : 6;
    const bool cinema = p.profile == J2KCUDA_PROFILE_CINEMA_2K || p.profile == J2KCUDA_PROFILE_CINEMA_4K;
    if (cinema) numres = p.profile == J2KCUDA_PROFILE_CINEMA_2K ? std::min(numres, 6u) : std::min(std::max(numres, 1u), 7u);
    if (numres > 33) return "Invalid number of resolutions : not in range [1,33]";
    const bool tiled = p.tile_w && p.tile_h && !cinema;
    const uint64_t tdx = tiled ? p.tile_w : p.width, tdy = tiled ? p.tile_h : p.height;
    if (tdx < ((uint64_t)1 << (numres - 1)) || tdy < ((uint64_t)1 << (numres - 1)))
        return "Number of resolutions is too high in comparison to the size of tiles";
    const uint32_t cw = p.cblk_w ? p.cblk_w : 64, ch = p.cblk_h ? p.cblk_h : 64;
    if (cw < 4 || ch < 4 || cw > 1024 || ch > 1024 || (cw & (cw - 1)) || (ch & (ch - 1)) || (uint64_t)cw * ch > 4096)
        return "code-block size must be a power of two between 4 and 1024 with at most 4096 samples";
    if (p.numlayers > 100) return "too many layers";
    if (p.prog_order > J2KCUDA_CPRL) return "unknown progression order";
    if (p.cblksty & ~0x26u) return "the encoder supports the code-block styles RESET (0x02), TERMALL (0x04) and SEGSYM (0x20); LAZY, VSC and PTERM are refused";
    return nullptr;
}

void setup_coding_params(const ::j2kcuda_params& p, CodingParams& cp) {
    cp = CodingParams();
    cp.x0 = cp.y0 = 0; cp.x1 = p.width; cp.y1 = p.height;
    cp.ncomp = (int)p.ncomp;
    cp.compat = (int)p.compat;
    int numres = p.numresolution ? (int)p.numresolution : 6;
    int cbw = p.cblk_w ? (int)p.cblk_w : 64, cbh = p.cblk_h ? (int)p.cblk_h : 64;
    int irreversible = p.irreversible ? 1 : 0;
    int numlayers = p.numlayers ? (int)p.numlayers : 1;
    int prog = (int)p.prog_order;
    int csty = 0;
    int mct = (int)p.mct;
    bool tile_on = p.tile_w && p.tile_h;
    uint32_t tdx = p.tile_w, tdy = p.tile_h;
    double rates[16];
    for (int i = 0; i < 16; ++i) rates[i] = p.rates[i];
    uint64_t max_cs = p.max_cs_size;
    uint32_t max_comp = p.max_comp_size;
    int rsiz = (int)p.profile;
    int res_spec = 0; int prc_init[MAX_RES];

    // j2k.cpp:5866-5892: the codestream cap limits the layer rates
    double raw_bits = (double)p.ncomp * p.width * p.height * p.prec;
    if (max_cs == 0) {
        if (numlayers > 0 && rates[numlayers - 1] > 0)
            max_cs = (uint64_t)std::floor(raw_bits / (rates[numlayers - 1] * 8.0));
    } else {
        double temp_rate = raw_bits / ((double)max_cs * 8);
        for (int i = 0; i < numlayers; ++i) if (rates[i] < temp_rate) rates[i] = temp_rate;
    }
    if (rsiz == J2KCUDA_PROFILE_CINEMA_2K || rsiz == J2KCUDA_PROFILE_CINEMA_4K) {   // :5643-5778
        tile_on = false;
        cp.tp_on = true; cp.tp_flag = 'C';
        cbw = cbh = 32;
        irreversible = 1;
        if (numlayers > 1) { rates[0] = rates[numlayers - 1]; numlayers = 1; }
        if (rsiz == J2KCUDA_PROFILE_CINEMA_2K) { if (numres > 6) numres = 6; }
        else { if (numres < 2) numres = 1; else if (numres > 7) numres = 7; }
        csty |= 1;
        res_spec = numres - 1;
        for (int i = 0; i < res_spec; ++i) prc_init[i] = 256;
        prog = J2KCUDA_CPRL;
        if (rsiz == J2KCUDA_PROFILE_CINEMA_4K) {
            cp.pocs.push_back(PocEntry{0, 0, 1, numres - 1, 3, J2KCUDA_CPRL});
            cp.pocs.push_back(PocEntry{numres - 1, 0, 1, numres, 3, J2KCUDA_CPRL});
        }
        if (max_cs == 0 || max_cs > 1302083) max_cs = 1302083;
        if (max_comp == 0 || max_comp > 1041666) max_comp = 1041666;
        rates[0] = raw_bits / ((double)max_cs * 8);
        // :5780-5840 compliance: 3 comps, 12-bit unsigned, size limits
        bool ok = p.ncomp == 3 && p.prec == 12 && !p.sgnd;
        if (rsiz == J2KCUDA_PROFILE_CINEMA_2K && (p.width > 2048 || p.height > 1080)) ok = false;
        if (rsiz == J2KCUDA_PROFILE_CINEMA_4K && (p.width > 4096 || p.height > 2160)) ok = false;
        if (!ok) rsiz = 0;
    }
    cp.rsiz = rsiz;
    cp.max_cs_size = max_cs;
    cp.max_comp_size = max_comp;
    cp.tx0 = cp.ty0 = 0;
    if (tile_on) {
        cp.tdx = tdx; cp.tdy = tdy;
        cp.tw = ceildiv(cp.x1 - cp.tx0, cp.tdx);
        cp.th = ceildiv(cp.y1 - cp.ty0, cp.tdy);
    } else {
        cp.tdx = cp.x1 - cp.tx0; cp.tdy = cp.y1 - cp.ty0;
        cp.tw = cp.th = 1;
    }
    cp.tile_first = p.tile_count ? p.tile_first : 0;
    cp.tile_count = p.tile_count;
    if (cp.tile_count && cp.tile_first + cp.tile_count > cp.tw * cp.th)
        cp.tile_count = cp.tile_first < cp.tw * cp.th ? cp.tw * cp.th - cp.tile_first : 0, cp.tile_first = cp.tile_count ? cp.tile_first : 0;
    cp.numlayers = numlayers;
    for (int i = 0; i < 16; ++i) cp.rates[i] = i < numlayers ? rates[i] : 0;
    cp.csty = csty;
    cp.prog = prog;
    cp.mct = (mct && cp.ncomp >= 3) ? 1 : 0;
    cp.comment = cp.compat == J2KCUDA_COMPAT_OPENJPEG_2_5 ? kCommentOpj25 : (cp.compat == J2KCUDA_COMPAT_OPENJPEG_2_5_3 ? kCommentOpj253 : kCommentGrok);
    for (int c = 0; c < cp.ncomp; ++c) {
        CompParams& t = cp.comps[c];
        t.prec = (int)p.prec; t.sgnd = p.sgnd ? 1 : 0; t.dx = t.dy = 1;
        t.dc_shift = t.sgnd ? 0 : 1 << (t.prec - 1);
        t.csty = csty & 1;
        t.numres = numres;
        t.cblkw = floorlog2((uint32_t)cbw); t.cblkh = floorlog2((uint32_t)cbh);
        t.cblksty = (int)p.cblksty;
        t.qmfbid = irreversible ? 0 : 1;
        t.qntsty = irreversible ? 2 : 0;
        t.numgbits = 2;
        t.roishift = 0;
        if (csty & 1) {                                   // :6155-6198
            int pp = 0;
            for (int r = numres - 1; r >= 0; --r, ++pp) {
                int size = pp < res_spec ? prc_init[pp] : (res_spec > 0 ? prc_init[res_spec - 1] >> (pp - (res_spec - 1)) : 32768);
                t.prcw[r] = t.prch[r] = size < 1 ? 1 : floorlog2((uint32_t)size);
            }
        } else {
            for (int r = 0; r < numres; ++r) t.prcw[r] = t.prch[r] = 15;
        }
        calc_explicit_stepsizes(t);
    }
}

void init_enc_tile(const CodingParams& cp, int tileno, EncTile& t) {
    build_tile(cp, tileno, 1.0f, t.geo);
    t.blocks.assign(t.geo.cblks.size(), EncBlock());
    for (auto& b : t.blocks) b.layers.assign(cp.numlayers, EncLayer());
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
    t.distotile = 0;
    t.packno = 0;
}

// index of the tag trees of (comp, res, band, prec)
static inline int tree_index(const CodingParams& cp, const EncTile& t, int c, int r, int b, int p) {
    int base = 0;
    int k = 0;
    for (int cc = 0; cc < c; ++cc) k += cp.comps[cc].numres;
    base = t.prec_base[k + r];
    const ResGeo& rg = t.geo.comps[c].res[r];
    return base + b * (rg.pw * rg.ph) + p;
}

// ---- layer formation: tcd.cpp:321-417 ----------------------------------------------------------
static void makelayer(const CodingParams& cp, EncTile& t, int layno, double thresh, bool final,
                      bool include_all, bool lossless = false) {
    for (auto& cblk : t.blocks) {
        EncLayer& layer = cblk.layers[layno];
        if (layno == 0) cblk.included_other = 0;
        int n = cblk.included_other;
        if (lossless && !include_all) {
            // every pass whose rate grows is in (no slope can be below the tile's smallest), a pass with no new bytes only when
            // it changes the distortion: by induction a pass is left out exactly when its cumulative (rate, distortion) equals
            // the previous pass's, so the layer ends at the last pass that differs from its predecessor -- found from the end
            int p = cblk.totalpasses - 1;
            const int32_t r0 = n ? cblk.rates[n - 1] : 0; const double d0 = n ? cblk.dist[n - 1] : 0.0;
            while (p >= n) {
                const int32_t rp = p > n ? cblk.rates[p - 1] : r0; const double dp = p > n ? cblk.dist[p - 1] : d0;
                if (cblk.rates[p] != rp || cblk.dist[p] != dp) break;
                --p;
            }
            n = p + 1 > n ? p + 1 : n;
        } else
        for (int passno = cblk.included_other; passno < cblk.totalpasses; ++passno) {
            uint32_t dr; double dd;
            if (n == 0) { dr = (uint32_t)cblk.rates[passno]; dd = cblk.dist[passno]; }
            else { dr = (uint32_t)(cblk.rates[passno] - cblk.rates[n - 1]); dd = cblk.dist[passno] - cblk.dist[n - 1]; }
            if (include_all) { n = passno + 1; continue; }     // OpenJPEG 2.5.x lossless layer
            if (!dr) { if (dd != 0) n = passno + 1; continue; }
            if (lossless) { n = passno + 1; continue; }         // thresh = the tile's smallest slope: every slope passes, no division
            double slope = dd / dr;
            if (thresh - slope < DBL_EPSILON) n = passno + 1;
        }
        layer.numpasses = n - cblk.included_other;
        if (!layer.numpasses) { layer.disto = 0; continue; }
        if (cblk.included_other == 0) {
            layer.len = cblk.rates[n - 1];
            layer.data_off = 0;
            layer.disto = cblk.dist[n - 1];
        } else {
            layer.len = cblk.rates[n - 1] - cblk.rates[cblk.included_other - 1];
            layer.data_off = cblk.rates[cblk.included_other - 1];
            layer.disto = cblk.dist[n - 1] - cblk.dist[cblk.included_other - 1];
        }
        if (final) cblk.included_other = n;
    }
    (void)cp;
}

// ---- packets: t2.cpp:193-223, 614-829 ----------------------------------------------------------
static void putcommacode(BitWriter& bio, int n) { while (--n >= 0) bio.putbit(1); bio.putbit(0); }
static void putnumpasses(BitWriter& bio, uint32_t n) {
    if (n == 1) bio.write(0, 1);
    else if (n == 2) bio.write(2, 2);
    else if (n <= 5) bio.write(0xc | (n - 3), 4);
    else if (n <= 36) bio.write(0x1e0 | (n - 6), 9);
    else if (n <= 164) bio.write(0xff80 | (n - 37), 16);
}

// Writes one packet (out != null) or only counts its bytes (simulate: t2.cpp:832-1000; the
// simulation walks the same header logic on the same mutable block state).
static uint64_t encode_packet(const CodingParams& cp, EncTile& t, const PacketId& pk,
                              ByteBuf* out) {
    const int compno = pk.compno, resno = pk.resno, precno = pk.precno, layno = pk.layno;
    const ResGeo& res = t.geo.comps[compno].res[resno];
    uint64_t nbytes = 0;
    if (cp.csty & 2) {                                    // SOP
        if (out) {
            const uint8_t sop[6] = {0xff, 0x91, 0, 4, (uint8_t)(t.packno >> 8), (uint8_t)t.packno};
            out->insert(out->end(), sop, sop + 6);
        }
        nbytes += 6;
    }
    if (!layno) {
        for (int b = 0; b < res.numbands; ++b) {
            const BandGeo& band = res.bands[b];
            if (band.r.w() <= 0 || band.r.h() <= 0) continue;          // see the note above encode_packet
            const PrecGeo& prc = band.precs[precno];
            int ti = tree_index(cp, t, compno, resno, b, precno);
            t.incl[ti].reset(); t.imsb[ti].reset();
            for (int k = 0; k < prc.cw * prc.ch; ++k) {
                EncBlock& cblk = t.blocks[prc.first_cblk + k];
                cblk.numpasses = 0;
                t.imsb[ti].setvalue(k, band.numbps - cblk.numbps);
            }
        }
    }
    BitWriter bio; bio.out = out;
    bio.putbit(1);
    for (int b = 0; b < res.numbands; ++b) {
        if (res.bands[b].r.w() <= 0 || res.bands[b].r.h() <= 0) continue;
        const PrecGeo& prc = res.bands[b].precs[precno];
        int ti = tree_index(cp, t, compno, resno, b, precno);
        int nb = prc.cw * prc.ch;
        for (int k = 0; k < nb; ++k) {
            EncBlock& cblk = t.blocks[prc.first_cblk + k];
            if (!cblk.numpasses && cblk.layers[layno].numpasses) t.incl[ti].setvalue(k, layno);
        }
        for (int k = 0; k < nb; ++k) {
            EncBlock& cblk = t.blocks[prc.first_cblk + k];
            const EncLayer& layer = cblk.layers[layno];
            if (!cblk.numpasses) t.incl[ti].encode(bio, k, layno + 1);
            else bio.putbit(layer.numpasses != 0);
            if (!layer.numpasses) continue;
            if (!cblk.numpasses) { cblk.numlenbits = 3; t.imsb[ti].encode(bio, k, 999); }
            putnumpasses(bio, (uint32_t)layer.numpasses);
            if (!(cp.comps[compno].cblksty & 4)) {
                // no pass of the contribution is terminated (cblksty 0 or RESET): one codeword segment, one length
                int len = layer.len;
                int increment = std::max(0, floorlog2((uint32_t)len) + 1 - (cblk.numlenbits + floorlog2((uint32_t)layer.numpasses)));
                putcommacode(bio, increment);
                cblk.numlenbits += increment;
                bio.write((uint32_t)len, cblk.numlenbits + floorlog2((uint32_t)layer.numpasses));
            } else {
                // TERMALL: every pass ends a codeword segment, so every pass has a length of its own (t2.cpp:738-790); Lblock
                // grows by the largest increment any of them needs
                const int p0 = cblk.numpasses, p1 = cblk.numpasses + layer.numpasses;
                auto passlen = [&](int p) { return cblk.rates[p] - (p ? cblk.rates[p - 1] : 0); };
                int increment = 0;
                for (int pno = p0; pno < p1; ++pno)
                    increment = std::max(increment, floorlog2((uint32_t)passlen(pno)) + 1 - cblk.numlenbits);   // one pass per segment: floorlog2(1) = 0
                putcommacode(bio, increment);
                cblk.numlenbits += increment;
                for (int pno = p0; pno < p1; ++pno) bio.write((uint32_t)passlen(pno), cblk.numlenbits);
            }
        }
    }
    bio.flush();
    nbytes += bio.n;
    if (cp.csty & 4) {                                    // EPH
        if (out) { out->push_back(0xff); out->push_back(0x92); }
        nbytes += 2;
    }
    // bodies: the encoder's packed arena holds the blocks of a band back to back in this very order, so the contributions of
    // neighbouring blocks are usually one run of memory: copied as runs, not block by block
    const uint8_t* run = nullptr; size_t runlen = 0;
    for (int b = 0; b < res.numbands; ++b) {
        if (res.bands[b].r.w() <= 0 || res.bands[b].r.h() <= 0) continue;
        const PrecGeo& prc = res.bands[b].precs[precno];
        for (int k = 0; k < prc.cw * prc.ch; ++k) {
            EncBlock& cblk = t.blocks[prc.first_cblk + k];
            const EncLayer& layer = cblk.layers[layno];
            if (!layer.numpasses) continue;
            if (out && layer.len > 0) {
                const uint8_t* src = cblk.data + layer.data_off;
                if (run && src == run + runlen) runlen += (size_t)layer.len;
                else { if (runlen) out->insert(out->end(), run, run + runlen); run = src; runlen = (size_t)layer.len; }
            }
            cblk.numpasses += layer.numpasses;
            nbytes += (uint64_t)layer.len;
        }
    }
    if (out && runlen) out->insert(out->end(), run, run + runlen);
    return nbytes;
}

// Tile parts: one progression (POC) x one split unit.  With tp_on and flag 'C' under CPRL each
// component of each POC is its own tile part (j2k.cpp:1737-1790); otherwise one part per POC.
struct TilePart { std::vector<PacketId> packets; int pino; };
static void tile_parts(const CodingParams& cp, const TileGeo& geo, std::vector<TilePart>& parts) {
    parts.clear();
    std::vector<uint8_t> include;
    std::vector<PocEntry> pocs = cp.pocs;
    int maxres = 0;
    for (int c = 0; c < cp.ncomp; ++c) maxres = std::max(maxres, cp.comps[c].numres);
    if (pocs.empty()) pocs.push_back(PocEntry{0, 0, cp.numlayers, maxres, cp.ncomp, cp.prog});
    for (size_t pi = 0; pi < pocs.size(); ++pi) {
        const PocEntry& pe = pocs[pi];
        bool split_c = cp.tp_on && cp.tp_flag == 'C' && pe.prg == 4;
        if (split_c) {
            for (int c = pe.compno0; c < std::min(pe.compno1, cp.ncomp); ++c) {
                TilePart tp; tp.pino = (int)pi;
                progression_packets(cp, geo, pe.prg, pe.resno0, pe.resno1, c, c + 1, 0, pe.layno1, include, tp.packets);
                parts.push_back(std::move(tp));
            }
        } else {
            TilePart tp; tp.pino = (int)pi;
            progression_packets(cp, geo, pe.prg, pe.resno0, pe.resno1, pe.compno0, pe.compno1, 0, pe.layno1, include, tp.packets);
            parts.push_back(std::move(tp));
        }
    }
}

// t2.cpp:319-384: would layers [0, maxlayers) fit in maxlen (and each component in max_comp_size)?
static bool simulate(const CodingParams& cp, EncTile& t, const std::vector<TilePart>& parts,
                     int maxlayers, uint64_t maxlen) {
    // the reference simulates component by component when max_comp_size is set, POC by POC
    int ncomp_iter = cp.max_comp_size > 0 ? cp.ncomp : 1;
    std::vector<PocEntry> pocs = cp.pocs;
    int maxres = 0;
    for (int c = 0; c < cp.ncomp; ++c) maxres = std::max(maxres, cp.comps[c].numres);
    if (pocs.empty()) pocs.push_back(PocEntry{0, 0, cp.numlayers, maxres, cp.ncomp, cp.prog});
    size_t npoc = cp.rsiz == J2KCUDA_PROFILE_CINEMA_4K ? 2 : 1;
    npoc = std::min(npoc, pocs.size());
    std::vector<uint8_t> include;
    int64_t remaining = (int64_t)maxlen;
    for (int c = 0; c < ncomp_iter; ++c) {
        uint64_t comp_len = 0;
        for (size_t pi = 0; pi < npoc; ++pi) {
            const PocEntry& pe = pocs[pi];
            std::vector<PacketId> pk;
            if (cp.max_comp_size > 0 && cp.tp_on)
                progression_packets(cp, t.geo, pe.prg, pe.resno0, pe.resno1, c, c + 1, 0, pe.layno1, include, pk);
            else
                progression_packets(cp, t.geo, pe.prg, pe.resno0, pe.resno1, pe.compno0, pe.compno1, 0, pe.layno1, include, pk);
            for (auto& id : pk) {
                if (id.layno >= maxlayers) continue;
                uint64_t n = encode_packet(cp, t, id, nullptr);
                if ((int64_t)n > remaining) return false;
                comp_len += n; remaining -= (int64_t)n;
            }
            if (cp.max_comp_size && comp_len > cp.max_comp_size) return false;
        }
    }
    (void)parts;
    return true;
}

// ---- rate-capped single layer: the bisection without redoing what cannot have changed ----------
// The reference redoes layer formation for every block and the Tier-2 simulation of every packet in each of the
// ~40 bisection steps (tcd.cpp:253-317): for the 26 280 blocks of a DCI 4K frame that is the critical path of the whole
// encode.  The decisions of one block are comparisons of the threshold with a handful of slopes, so
// its outcome is known to stay the same while the threshold stays between the largest slope it refused and the smallest
// one it accepted; a packet's byte count stays the same while none of its blocks changed.  Only the rest is recomputed.
// Every pass/fail answer is the one the full computation would give, and the final layer is formed by the plain
// routine, so the codestream is unchanged (tests/test_t2_host.py, tests/test_gpu_configs.py compare it with the reference).
struct BisectCache {
    struct Visit { int packet; int kind; };          // kind 0: packet; 1: start of a component; 2: end of a progression
    std::vector<PacketId> packets;
    std::vector<Visit> order;
    std::vector<uint64_t> size;
    std::vector<uint8_t> dirty;
    std::vector<int> blk_packet;
    std::vector<double> lo, hi;                       // per block: thresholds for which its cached layer holds
    std::vector<uint8_t> valid;
};

static void bisect_cache_init(const CodingParams& cp, EncTile& t, BisectCache& bc) {
    // the visiting order of simulate() below, flattened once
    int ncomp_iter = cp.max_comp_size > 0 ? cp.ncomp : 1;
    std::vector<PocEntry> pocs = cp.pocs;
    int maxres = 0;
    for (int c = 0; c < cp.ncomp; ++c) maxres = std::max(maxres, cp.comps[c].numres);
    if (pocs.empty()) pocs.push_back(PocEntry{0, 0, cp.numlayers, maxres, cp.ncomp, cp.prog});
    size_t npoc = cp.rsiz == J2KCUDA_PROFILE_CINEMA_4K ? 2 : 1;
    npoc = std::min(npoc, pocs.size());
    std::vector<uint8_t> include;
    for (int c = 0; c < ncomp_iter; ++c) {
        bc.order.push_back({-1, 1});
        for (size_t pi = 0; pi < npoc; ++pi) {
            const PocEntry& pe = pocs[pi];
            std::vector<PacketId> pk;
            if (cp.max_comp_size > 0 && cp.tp_on)
                progression_packets(cp, t.geo, pe.prg, pe.resno0, pe.resno1, c, c + 1, 0, pe.layno1, include, pk);
            else
                progression_packets(cp, t.geo, pe.prg, pe.resno0, pe.resno1, pe.compno0, pe.compno1, 0, pe.layno1, include, pk);
            for (auto& id : pk) {
                if (id.layno >= 1) continue;
                bc.order.push_back({(int)bc.packets.size(), 0});
                bc.packets.push_back(id);
            }
            bc.order.push_back({-1, 2});
        }
    }
    bc.size.assign(bc.packets.size(), 0);
    bc.dirty.assign(bc.packets.size(), 1);
    bc.blk_packet.assign(t.blocks.size(), -1);
    for (size_t i = 0; i < bc.packets.size(); ++i) {
        const PacketId& id = bc.packets[i];
        const ResGeo& res = t.geo.comps[id.compno].res[id.resno];
        for (int b = 0; b < res.numbands; ++b) {
            const PrecGeo& prc = res.bands[b].precs[id.precno];
            for (int k = 0; k < prc.cw * prc.ch; ++k) bc.blk_packet[prc.first_cblk + k] = (int)i;
        }
    }
    bc.lo.assign(t.blocks.size(), 0); bc.hi.assign(t.blocks.size(), 0);
    bc.valid.assign(t.blocks.size(), 0);
}

// makelayer(cp, t, 0, thresh, false, false) for the blocks whose cached answer may not hold at this threshold
static void makelayer_cached(EncTile& t, double thresh, BisectCache& bc) {
    const size_t nb = t.blocks.size();
    for (size_t i = 0; i < nb; ++i) {
        if (bc.valid[i] && bc.lo[i] <= thresh && thresh <= bc.hi[i]) continue;
        EncBlock& cblk = t.blocks[i];
        EncLayer& layer = cblk.layers[0];
        cblk.included_other = 0;
        int n = 0;
        double min_in = DBL_MAX, max_out = -DBL_MAX;
        for (int passno = 0; passno < cblk.totalpasses; ++passno) {
            uint32_t dr; double dd;
            if (n == 0) { dr = (uint32_t)cblk.rates[passno]; dd = cblk.dist[passno]; }
            else { dr = (uint32_t)(cblk.rates[passno] - cblk.rates[n - 1]); dd = cblk.dist[passno] - cblk.dist[n - 1]; }
            if (!dr) { if (dd != 0) n = passno + 1; continue; }
            const double slope = dd / dr;
            if (thresh - slope < DBL_EPSILON) { n = passno + 1; if (slope < min_in) min_in = slope; }
            else if (slope > max_out) max_out = slope;
        }
        // same decisions for every threshold in [max_out + 1, min_in]: accepted slopes stay >= it (difference <= 0 <
        // epsilon), refused ones stay at least 1 below it (difference >= 1 - rounding > epsilon)
        bc.lo[i] = max_out == -DBL_MAX ? -DBL_MAX : (std::fabs(max_out) < 1e15 ? max_out + 1.0 : DBL_MAX);
        bc.hi[i] = min_in;
        const int oldn = bc.valid[i] ? layer.numpasses : -1;
        bc.valid[i] = 1;
        layer.numpasses = n;
        if (!n) layer.disto = 0;
        else { layer.len = cblk.rates[n - 1]; layer.data_off = 0; layer.disto = cblk.dist[n - 1]; }
        if (oldn != n && bc.blk_packet[i] >= 0) bc.dirty[bc.blk_packet[i]] = 1;
    }
}

// simulate(cp, t, parts, 1, maxlen): the same checks in the same order on byte counts that are recomputed only for
// packets with a changed block
static bool simulate_cached(const CodingParams& cp, EncTile& t, uint64_t maxlen, BisectCache& bc) {
    for (size_t i = 0; i < bc.packets.size(); ++i)
        if (bc.dirty[i]) { bc.size[i] = encode_packet(cp, t, bc.packets[i], nullptr); bc.dirty[i] = 0; }
    int64_t remaining = (int64_t)maxlen;
    uint64_t comp_len = 0;
    for (const auto& v : bc.order) {
        if (v.kind == 1) { comp_len = 0; continue; }
        if (v.kind == 2) { if (cp.max_comp_size && comp_len > cp.max_comp_size) return false; continue; }
        const uint64_t n = bc.size[v.packet];
        if ((int64_t)n > remaining) return false;
        comp_len += n; remaining -= (int64_t)n;
    }
    return true;
}

// tcd.cpp:165-319
static void pcrd_bisect(const CodingParams& cp, EncTile& t, const std::vector<TilePart>& parts,
                        const double* rates, uint64_t len) {
    double min_slope = DBL_MAX, max_slope = 0;
    bool any_rate = false;
    for (int layno = 0; layno < cp.numlayers; ++layno) any_rate |= rates[layno] > 0.0;
    // lossless layers only (the plug-in's case): the threshold would be the smallest slope of the tile, which every pass
    // meets -- the slopes need not be computed at all (tcd.cpp:180-239, 309, 356-382)
    if (any_rate)
    for (auto& cblk : t.blocks)
        for (int passno = 0; passno < cblk.totalpasses; ++passno) {
            int32_t dr; double dd;
            if (passno == 0) { dr = cblk.rates[0]; dd = cblk.dist[0]; }
            else { dr = cblk.rates[passno] - cblk.rates[passno - 1]; dd = cblk.dist[passno] - cblk.dist[passno - 1]; }
            if (dr == 0) continue;
            double s = dd / dr;
            if (s < min_slope) min_slope = s;
            if (s > max_slope) max_slope = s;
        }
    for (int layno = 0; layno < cp.numlayers; ++layno) {
        double lo = min_slope, hi = max_slope;
        uint64_t maxlen = rates[layno] > 0.0 ? std::min<uint64_t>((uint64_t)std::ceil(rates[layno]), len) : len;
        double goodthresh = 0, old_thresh = -1;
        bool all = false;
        if (rates[layno] > 0.0) {
            double thresh = 0;
            const bool cached = cp.numlayers == 1 && !(cp.csty & 2);     // SOP packet numbers depend on the visiting history
            BisectCache bc;
            if (cached) bisect_cache_init(cp, t, bc);
            for (int i = 0; i < 128; ++i) {
                thresh = (lo + hi) / 2;
                if (cached) makelayer_cached(t, thresh, bc);
                else makelayer(cp, t, layno, thresh, false, false);
                if (std::fabs(old_thresh - thresh) < 0.001) break;
                old_thresh = thresh;
                if (!(cached ? simulate_cached(cp, t, maxlen, bc) : simulate(cp, t, parts, layno + 1, maxlen))) { lo = thresh; continue; }
                hi = thresh;
            }
            goodthresh = (hi == 0) ? thresh : hi;
        } else {
            goodthresh = min_slope;
            all = cp.compat == J2KCUDA_COMPAT_OPENJPEG_2_5 || cp.compat == J2KCUDA_COMPAT_OPENJPEG_2_5_3;
        }
        makelayer(cp, t, layno, goodthresh, true, all, !any_rate);
    }
}

// ---- markers: j2k.cpp:1871-4340 ---------------------------------------------------------------
template <class B> static void put(B& o, uint64_t v, int nbytes) {
    for (int i = nbytes - 1; i >= 0; --i) o.push_back((uint8_t)(v >> (8 * i)));
}
static void write_spcod(std::vector<uint8_t>& o, const CompParams& t) {      // :8289-8370
    put(o, t.numres - 1, 1); put(o, t.cblkw - 2, 1); put(o, t.cblkh - 2, 1);
    put(o, t.cblksty, 1); put(o, t.qmfbid, 1);
    if (t.csty & 1) for (int r = 0; r < t.numres; ++r) put(o, t.prcw[r] + (t.prch[r] << 4), 1);
}
static void write_sqcd(std::vector<uint8_t>& o, const CompParams& t) {       // :8574-8660
    int nb = t.qntsty == 1 ? 1 : 3 * t.numres - 2;
    put(o, t.qntsty + (t.numgbits << 5), 1);
    for (int b = 0; b < nb; ++b) {
        if (t.qntsty == 0) put(o, t.steps[b].expn << 3, 1);
        else put(o, (t.steps[b].expn << 11) + t.steps[b].mant, 2);
    }
}
static bool same_coding(const CompParams& a, const CompParams& b) {          // :4597-4620
    if (a.numres != b.numres || a.cblkw != b.cblkw || a.cblkh != b.cblkh || a.cblksty != b.cblksty ||
        a.qmfbid != b.qmfbid || (a.csty & 1) != (b.csty & 1)) return false;
    for (int r = 0; r < a.numres; ++r) if (a.prcw[r] != b.prcw[r] || a.prch[r] != b.prch[r]) return false;
    return true;
}
static bool same_quant(const CompParams& a, const CompParams& b) {           // :4622-4642
    if (a.qntsty != b.qntsty || a.numgbits != b.numgbits) return false;
    int nb = a.qntsty == 1 ? 1 : 3 * a.numres - 2;
    for (int i = 0; i < nb; ++i) {
        if (a.steps[i].expn != b.steps[i].expn) return false;
        if (a.qntsty != 0 && a.steps[i].mant != b.steps[i].mant) return false;
    }
    return true;
}

static void write_main_header(const CodingParams& cp, int total_tile_parts, ByteBuf& o,
                              size_t& tlm_pos) {
    put(o, 0xff4f, 2);                                                        // SOC
    put(o, 0xff51, 2); put(o, 38 + 3 * cp.ncomp, 2); put(o, cp.rsiz, 2);       // SIZ :1938-2030
    put(o, cp.x1, 4); put(o, cp.y1, 4); put(o, cp.x0, 4); put(o, cp.y0, 4);
    put(o, cp.tdx, 4); put(o, cp.tdy, 4); put(o, cp.tx0, 4); put(o, cp.ty0, 4);
    put(o, cp.ncomp, 2);
    for (int c = 0; c < cp.ncomp; ++c) {
        put(o, cp.comps[c].prec - 1 + (cp.comps[c].sgnd << 7), 1);
        put(o, cp.comps[c].dx, 1); put(o, cp.comps[c].dy, 1);
    }
    {                                                                         // COD :2338-2420
        std::vector<uint8_t> sp; write_spcod(sp, cp.comps[0]);
        put(o, 0xff52, 2); put(o, 2 + 1 + 4 + sp.size(), 2);   // Lcod = len + Scod + SGcod + SPcod
        put(o, cp.csty, 1); put(o, cp.prog, 1); put(o, cp.numlayers, 2); put(o, cp.mct, 1);
        o.insert(o.end(), sp.begin(), sp.end());
    }
    {                                                                         // QCD :2700-2788
        std::vector<uint8_t> sq; write_sqcd(sq, cp.comps[0]);
        put(o, 0xff5c, 2); put(o, 2 + sq.size(), 2);
        o.insert(o.end(), sq.begin(), sq.end());
    }
    for (int c = 1; c < cp.ncomp; ++c)                                        // COC :2509-2616
        if (!same_coding(cp.comps[0], cp.comps[c])) {
            std::vector<uint8_t> sp; write_spcod(sp, cp.comps[c]);
            put(o, 0xff53, 2); put(o, 2 + 1 + 1 + sp.size(), 2);
            put(o, c, 1); put(o, cp.comps[c].csty, 1);
            o.insert(o.end(), sp.begin(), sp.end());
        }
    for (int c = 1; c < cp.ncomp; ++c)                                        // QCC :2790-2880
        if (!same_quant(cp.comps[0], cp.comps[c])) {
            std::vector<uint8_t> sq; write_sqcd(sq, cp.comps[c]);
            put(o, 0xff5d, 2); put(o, 2 + 1 + sq.size(), 2); put(o, c, 1);
            o.insert(o.end(), sq.begin(), sq.end());
        }
    tlm_pos = 0;
    bool cinema = cp.rsiz == J2KCUDA_PROFILE_CINEMA_2K || cp.rsiz == J2KCUDA_PROFILE_CINEMA_4K;
    if (cinema) {                                                             // TLM :3772-3824
        put(o, 0xff55, 2); put(o, 4 + 5 * total_tile_parts, 2); put(o, 0, 1); put(o, 0x50, 1);
        tlm_pos = o.size();
        for (int k = 0; k < 5 * total_tile_parts; ++k) o.push_back(0);
        if (cp.rsiz == J2KCUDA_PROFILE_CINEMA_4K && !cp.pocs.empty()) {       // POC :2947-3069
            put(o, 0xff5f, 2); put(o, 2 + 7 * cp.pocs.size(), 2);
            for (auto& pe : cp.pocs) {
                put(o, pe.resno0, 1); put(o, pe.compno0, 1); put(o, pe.layno1, 2);
                put(o, pe.resno1, 1); put(o, pe.compno1, 1); put(o, pe.prg, 1);
            }
        }
    }
    if (!cp.comment.empty()) {                                                // COM :2264-2335
        put(o, 0xff64, 2); put(o, 4 + cp.comment.size(), 2); put(o, 1, 2);
        o.insert(o.end(), cp.comment.begin(), cp.comment.end());
    }
}

bool EncoderHost::assemble(std::vector<EncTile>& tiles, ByteBuf& out, std::string& err) {
    const int nt = (int)(cp.tw * cp.th);
    auto tS = std::chrono::steady_clock::now();
    if ((int)tiles.size() != nt) { err = "tile count mismatch"; return false; }
    std::vector<std::vector<TilePart>> parts(nt);
    int total_tp = 0;
    const bool fragment = cp.tile_count != 0;
    for (int i = 0; i < nt; ++i) {
        if (!cp.tile_in_range(i) && fragment) {
            // tile-part counts of the other ranks' tiles are still needed for the header's size
            TileGeo g; build_tile(cp, i, 1.0f, g); tile_parts(cp, g, parts[i]);
        } else tile_parts(cp, tiles[i].geo, parts[i]);
        total_tp += (int)parts[i].size();
    }
    size_t tlm_pos = 0;
    out.clear();
    {
        // one allocation for the whole codestream: kept code-block bytes + headers
        size_t est = 65536;
        for (auto& t : tiles) {
            for (auto& b : t.blocks) if (b.totalpasses > 0) est += (size_t)b.rates[b.totalpasses - 1] + 8;
            est += t.blocks.size() * 4 + 4096;
        }
        (void)fragment;
        out.reserve(est);
    }
    write_main_header(cp, total_tp, out, tlm_pos);
    const bool cinema = tlm_pos != 0;
    if (cinema && fragment) { err = "tile ranges cannot be combined with the cinema profiles (TLM)"; return false; }
    const size_t header_len = out.size();
    // j2k.cpp:4355-4449: layer rates (compression ratios) -> byte budgets per tile
    const double sot_remove = (double)out.size() / nt;
    std::vector<uint8_t> tlm;
    for (int i = 0; i < nt; ++i) {
        if (!cp.tile_in_range(i)) continue;
        EncTile& t = tiles[i];
        double rates[16] = {0};
        double offset = (cp.tp_on ? (double)(float)((parts[i].size() - 1) * 14) : 0.0) / cp.numlayers;
        double bits_empty = 8.0 * cp.comps[0].dx * cp.comps[0].dy;
        double size_pixel = (double)cp.ncomp * cp.comps[0].prec;
        for (int k = 0; k < cp.numlayers; ++k) {
            rates[k] = cp.rates[k];
            if (rates[k] > 0.0)
                rates[k] = (size_pixel * (double)(uint32_t)t.geo.r.w() * (double)(uint32_t)t.geo.r.h()) / (rates[k] * bits_empty) - offset;
        }
        {
            int k = 0;
            if (rates[0] > 0.0) { rates[0] -= sot_remove; if (rates[0] < 30.0) rates[0] = 30.0; }
            int last = cp.numlayers - 1;
            for (k = 1; k < last; ++k)
                if (rates[k] > 0.0) { rates[k] -= sot_remove; if (rates[k] < rates[k - 1] + 10.0) rates[k] = rates[k - 1] + 20.0; }
            // the reference's pointer has advanced to layer max(1, last) here (:4441-4446); with a
            // single layer that is one past the array, which it leaves untouched
            if (last >= 1 && rates[last] > 0.0) {
                rates[last] -= (sot_remove + 2.0);
                if (rates[last] < rates[last - 1] + 10.0) rates[last] = rates[last - 1] + 20.0;
            }
        }
        // encoded tile buffer size (:4451-4462) bounds the layer length
        uint64_t tile_size = 0;
        for (int c = 0; c < cp.ncomp; ++c)
            tile_size += (uint64_t)ceildiv(cp.tdx, cp.comps[c].dx) * ceildiv(cp.tdy, cp.comps[c].dy) * cp.comps[c].prec;
        tile_size = (uint64_t)(tile_size * 0.1625) + 12 * parts[i].size();
        uint64_t len = tile_size > 16 ? tile_size - 16 : 0;
        bool any_rate = false;
        for (int k = 0; k < cp.numlayers; ++k) any_rate |= rates[k] > 0.0;
        if (!any_rate) len = ~(uint64_t)0 >> 1;
        auto tA = std::chrono::steady_clock::now();
        pcrd_bisect(cp, t, parts[i], rates, len);
        if (getenv("J2KCUDA_T2_TIMING")) fprintf(stderr, "pcrd %.2f ms\n", std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - tA).count());
        t.packno = 0;
        for (size_t tp = 0; tp < parts[i].size(); ++tp) {
            size_t sot = out.size();
            put(out, 0xff90, 2); put(out, 10, 2); put(out, i, 2); put(out, 0, 4);   // SOT :3826-3862
            put(out, tp, 1); put(out, parts[i].size(), 1);
            if (tp == 0 && !cinema && !cp.pocs.empty()) {                          // :10196-10203
                put(out, 0xff5f, 2); put(out, 2 + 7 * cp.pocs.size(), 2);
                for (auto& pe : cp.pocs) {
                    put(out, pe.resno0, 1); put(out, pe.compno0, 1); put(out, pe.layno1, 2);
                    put(out, pe.resno1, 1); put(out, pe.compno1, 1); put(out, pe.prg, 1);
                }
            }
            put(out, 0xff93, 2);                                                   // SOD
            for (auto& pk : parts[i][tp].packets) {
                if (pk.layno >= cp.numlayers) continue;
                encode_packet(cp, t, pk, &out);
                ++t.packno;
            }
            uint64_t psot = out.size() - sot;
            for (int b = 0; b < 4; ++b) out[sot + 6 + b] = (uint8_t)(psot >> (8 * (3 - b)));
            if (cinema) { put(tlm, i, 1); put(tlm, psot, 4); }                     // :974-992
        }
    }
    if (getenv("J2KCUDA_T2_TIMING")) fprintf(stderr, "assemble total %.2f ms\n", std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - tS).count());
    if (fragment) { out.erase(out.begin(), out.begin() + (long)header_len); return true; }   // tile-parts alone
    put(out, 0xffd9, 2);                                                           // EOC
    if (cinema) std::copy(tlm.begin(), tlm.end(), out.begin() + (long)tlm_pos);
    return true;
}

// One frame from the tile-part fragments of several encoders (tile ranges): main header, the
// fragments ordered by the tile index of their first SOT, EOC.
bool join_tile_streams(const CodingParams& cp_in, const uint8_t* const* frags, const size_t* lens, int nfrags,
                       ByteBuf& out, std::string& err) {
    CodingParams cp = cp_in;
    cp.tile_first = cp.tile_count = 0;
    const int nt = (int)(cp.tw * cp.th);
    int total_tp = 0;
    for (int i = 0; i < nt; ++i) {
        TileGeo g; std::vector<TilePart> parts;
        build_tile(cp, i, 1.0f, g); tile_parts(cp, g, parts);
        total_tp += (int)parts.size();
    }
    size_t tlm_pos = 0;
    out.clear();
    size_t total = 65536;
    for (int i = 0; i < nfrags; ++i) total += lens[i];
    out.reserve(total);
    write_main_header(cp, total_tp, out, tlm_pos);
    if (tlm_pos) { err = "cinema profiles (TLM) cannot be joined from tile fragments"; return false; }
    std::vector<std::pair<int, int>> order;          // (first tile index, fragment)
    std::vector<char> seen((size_t)nt, 0);
    for (int i = 0; i < nfrags; ++i) {
        if (lens[i] == 0) continue;
        // walk the tile-parts of the fragment: SOT (0xff90), Lsot = 10, Isot, Psot
        size_t pos = 0; int first = -1;
        while (pos < lens[i]) {
            const uint8_t* q = frags[i] + pos;
            if (lens[i] - pos < 12 || q[0] != 0xff || q[1] != 0x90) { err = "fragment does not start at a tile-part"; return false; }
            const int isot = (q[4] << 8) | q[5];
            const size_t psot = ((size_t)q[6] << 24) | ((size_t)q[7] << 16) | ((size_t)q[8] << 8) | q[9];
            if (isot >= nt || psot < 14 || psot > lens[i] - pos) { err = "bad tile-part in fragment"; return false; }
            if (q[10] == 0) { if (seen[(size_t)isot]) { err = "tile present in two fragments"; return false; } seen[(size_t)isot] = 1; }
            if (first < 0) first = isot;
            pos += psot;
        }
        order.push_back({first, i});
    }
    for (int t = 0; t < nt; ++t) if (!seen[(size_t)t]) { err = "tile " + std::to_string(t) + " is missing from the fragments"; return false; }
    std::sort(order.begin(), order.end());
    for (auto& o : order) out.insert(out.end(), frags[o.second], frags[o.second] + lens[o.second]);
    put(out, 0xffd9, 2);
    return true;
}

}  // namespace j2k
