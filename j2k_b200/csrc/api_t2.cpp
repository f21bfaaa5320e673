// C-ABI entry points that are host-only: parameter defaults, Tier-2 assembly from externally
// supplied block results (the grok plug-in hand-off), Tier-2 parse, probe.  See include/j2kcuda.h.
#include "../../include/j2kcuda.h"
#include "t2.h"
#include "common.h"
#include <cstdlib>
#include <cstring>

using namespace j2k;

extern "C" {

int j2kcuda_version(void) { return j2k::guarded([&]() -> int { return J2KCUDA_VERSION; }); }
const char* j2kcuda_last_error(void) { return j2k::last_error().c_str(); }
void j2kcuda_free(void* p) { j2k::out_free(p); }

void j2kcuda_default_params(j2kcuda_params* p) {   // openjpeg.cpp:706-728
    if (!p) return;
    memset(p, 0, sizeof(*p));
    p->numresolution = 6;
    p->cblk_w = p->cblk_h = 64;
    p->prog_order = J2KCUDA_LRCP;
    p->numlayers = 1;
    p->device = -1;
}

int j2kcuda_t2_block_count(const j2kcuda_params* p) { return j2k::guarded([&]() -> int {
    if (!p) return -1;
    if (const char* why = check_params(*p)) { set_error(J2KCUDA_ERR_PARAM, why); return -1; }
    CodingParams cp;
    setup_coding_params(*p, cp);
    int n = 0;
    for (int t = 0; t < (int)(cp.tw * cp.th); ++t) {
        if (!cp.tile_in_range(t)) continue;
        TileGeo g;
        build_tile(cp, t, 1.0f, g);
        n += (int)g.cblks.size();
    }
    return n;
}); }

int j2kcuda_t2_block_geometry(const j2kcuda_params* p, int32_t* geo, int max_blocks) { return j2k::guarded([&]() -> int {
    if (!p || !geo) return -1;
    if (const char* why = check_params(*p)) { set_error(J2KCUDA_ERR_PARAM, why); return -1; }
    CodingParams cp;
    setup_coding_params(*p, cp);
    int n = 0;
    for (int t = 0; t < (int)(cp.tw * cp.th); ++t) {
        if (!cp.tile_in_range(t)) continue;
        TileGeo g;
        build_tile(cp, t, 1.0f, g);
        for (auto& cb : g.cblks) {
            if (n >= max_blocks) return n;
            const BandGeo& band = g.comps[cb.comp].res[cb.res].bands[cb.band];
            int32_t* o = geo + 12 * (size_t)n;
            o[0] = t; o[1] = cb.comp; o[2] = cb.px; o[3] = cb.py; o[4] = cb.r.w(); o[5] = cb.r.h();
            o[6] = cb.orient; o[7] = cp.comps[cb.comp].numres - 1 - cb.res; o[8] = band.numbps;
            o[9] = band_const(band.stepsize);
            memcpy(&o[10], &band.stepsize, 4);
            o[11] = cb.res;
            ++n;
        }
    }
    return n;
}); }

// Where each code-block sits in the tile -> component -> resolution -> band -> precinct hierarchy (tcd.cpp:592-842), 8 int32 per
// block in the order of j2kcuda_t2_block_geometry: band index within the resolution, precinct index, block index within the
// precinct, precincts of the resolution (pw * ph, empty ones included), then the block's bounds in band coordinates
// x0, y0, x1, y1.  What a filler of grok's opj_plugin_tile_t needs (openjpeg.h:1308-1360).
int j2kcuda_t2_block_structure(const j2kcuda_params* p, int32_t* out, int max_blocks) { return j2k::guarded([&]() -> int {
    if (!p || !out) return -1;
    if (const char* why = check_params(*p)) { set_error(J2KCUDA_ERR_PARAM, why); return -1; }
    CodingParams cp;
    setup_coding_params(*p, cp);
    int n = 0;
    for (int t = 0; t < (int)(cp.tw * cp.th); ++t) {
        if (!cp.tile_in_range(t)) continue;
        TileGeo g;
        build_tile(cp, t, 1.0f, g);
        for (size_t k = 0; k < g.cblks.size(); ++k) {
            if (n >= max_blocks) return n;
            const CblkGeo& cb = g.cblks[k];
            const ResGeo& rg = g.comps[cb.comp].res[cb.res];
            const PrecGeo& pg = rg.bands[cb.band].precs[cb.prec];
            int32_t* o = out + 8 * (size_t)n;
            o[0] = cb.band; o[1] = cb.prec; o[2] = (int32_t)k - pg.first_cblk; o[3] = rg.pw * rg.ph;
            o[4] = cb.r.x0; o[5] = cb.r.y0; o[6] = cb.r.x1; o[7] = cb.r.y1;
            ++n;
        }
    }
    return n;
}); }

int j2kcuda_t2_assemble(const j2kcuda_params* p, const j2kcuda_block_result* blocks, int nblocks,
                        uint8_t** out, size_t* out_len) { return j2k::guarded([&]() -> int {
    if (!p || !blocks || !out || !out_len) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    if (const char* why = check_params(*p)) return set_error(J2KCUDA_ERR_PARAM, why);
    EncoderHost host;
    setup_coding_params(*p, host.cp);
    int nt = (int)(host.cp.tw * host.cp.th);
    std::vector<EncTile> tiles(nt);
    int k = 0;
    for (int t = 0; t < nt; ++t) {
        if (!host.cp.tile_in_range(t)) continue;
        init_enc_tile(host.cp, t, tiles[t]);
        for (auto& b : tiles[t].blocks) {
            if (k >= nblocks) return set_error(J2KCUDA_ERR_PARAM, "too few block results");
            b.data = blocks[k].data; b.numbps = blocks[k].numbps; b.totalpasses = blocks[k].npasses;
            b.rates = blocks[k].rates; b.dist = blocks[k].dist;
            ++k;
        }
    }
    if (k != nblocks) return set_error(J2KCUDA_ERR_PARAM, "block count mismatch");
    ByteBuf cs;
    std::string err;
    if (!host.assemble(tiles, cs, err)) return set_error(J2KCUDA_ERR_PARAM, err);
    *out_len = cs.size();
    *out = cs.release();
    return J2KCUDA_OK;
}); }

int j2kcuda_join_tile_streams(const j2kcuda_params* p, const uint8_t* const* frags, const size_t* lens, int nfrags,
                              uint8_t** out, size_t* out_len) { return j2k::guarded([&]() -> int {
    if (!p || !frags || !lens || !out || !out_len || nfrags < 0) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    CodingParams cp;
    setup_coding_params(*p, cp);
    ByteBuf cs;
    std::string err;
    if (!join_tile_streams(cp, frags, lens, nfrags, cs, err)) return set_error(J2KCUDA_ERR_PARAM, err);
    *out_len = cs.size();
    *out = cs.release();
    return J2KCUDA_OK;
}); }

}  // extern "C"
