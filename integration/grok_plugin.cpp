// libgrok_plugin.so -- libj2kcuda in the accelerator slot of the reference's own codec.
//
// ext/grok loads "lib<OPENJPEG_PLUGIN_NAME>.so" (openjpeg.cpp:1077-1090), asks it to encode (plugin_encode,
// plugin_interface.h:78) and gets back, through a callback, an opj_image_t plus an opj_plugin_tile_t (openjpeg.h:1308-1360):
// per code-block the bytes, the bit-plane count and per pass {rate, distortion decrease}.  The application then calls
// opj_encode_with_plugin(codec, tile, stream): opj_tcd_encode_tile skips DC shift, MCT, DWT and Tier-1 (tcd.cpp:1047-1079) and
// the rate allocation pulls every block's results in (encode_synch_with_plugin, plugin_bridge.cpp:72-174) -- the reference's
// own device / host split, which is also libj2kcuda's.  Here MCT, DWT and Tier-1 run on the B200 through the C ABI
// (j2kcuda_encoder_upload / _run / _block_results); PCRD, packets and markers stay with grok.
//
// With J2KCUDA_GROK_PLUGIN_DEBUG=1 plugin_get_debug_state() answers OPJ_PLUGIN_STATE_DEBUG_ENCODE: grok then runs its own
// Tier-1 on the DWT coefficients the plug-in hands over as image data and reports every difference in pass count, bytes,
// rates and distortion (plugin_bridge.cpp:86-170): a reference-defined parity check of the CUDA path.
//
// Built by oracle/Makefile against the reference's headers into oracle/_ref/libgrok_plugin.so (test artefact; this file is
// the source a maintainer would build next to ext/grok).  Input images: binary PGM / PPM (P5 / P6), 8 or 16 bit.
#include "plugin_interface.h"
#include "j2kcuda.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#define PLUGIN_EXPORT extern "C" __attribute__((visibility("default")))

namespace {

const char* kPluginId = "j2kcuda";

int32_t exit_func() { j2kcuda_shutdown(); return 0; }
void* create(minpf_object_params*) { return 0; }
int32_t destroy(void*) { return 0; }

bool debug_mode() { const char* e = getenv("J2KCUDA_GROK_PLUGIN_DEBUG"); return e && e[0] == '1'; }

// ---- binary PNM reader: P5 (one component) / P6 (three), maxval < 256: 1 byte per sample, else 2 bytes big-endian ----
struct Pnm { int w = 0, h = 0, nc = 0, maxval = 0, prec = 0; std::vector<uint16_t> pix; };

bool read_token(FILE* f, int& v) {
    int c = fgetc(f);
    for (;;) {
        while (c == ' ' || c == '\n' || c == '\r' || c == '\t') c = fgetc(f);
        if (c != '#') break;
        while (c != '\n' && c != EOF) c = fgetc(f);
    }
    if (c < '0' || c > '9') return false;
    v = 0;
    while (c >= '0' && c <= '9') { v = v * 10 + (c - '0'); c = fgetc(f); }
    return true;                                                     // the single whitespace after the token is consumed
}

bool read_pnm(const char* path, Pnm& im) {
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    bool ok = false;
    const int m0 = fgetc(f), m1 = fgetc(f);
    if (m0 == 'P' && (m1 == '5' || m1 == '6') && read_token(f, im.w) && read_token(f, im.h) && read_token(f, im.maxval) &&
        im.w > 0 && im.h > 0 && im.maxval > 0 && im.maxval < 65536) {
        im.nc = m1 == '5' ? 1 : 3;
        im.prec = 1;
        while ((1 << im.prec) <= im.maxval) ++im.prec;
        const size_t n = (size_t)im.w * im.h * im.nc;
        im.pix.resize(n);
        if (im.maxval < 256) {
            std::vector<uint8_t> b(n);
            ok = fread(b.data(), 1, n, f) == n;
            for (size_t i = 0; i < n; ++i) im.pix[i] = b[i];
        } else {
            std::vector<uint8_t> b(2 * n);
            ok = fread(b.data(), 1, 2 * n, f) == 2 * n;
            for (size_t i = 0; i < n; ++i) im.pix[i] = (uint16_t)((b[2 * i] << 8) | b[2 * i + 1]);
        }
    }
    fclose(f);
    return ok;
}

// ---- the tile handed to grok: owns its nested arrays ----
struct TileStore {
    opj_plugin_tile_t tile;
    std::vector<opj_plugin_tile_component_t> comps;
    std::vector<opj_plugin_tile_component_t*> comp_ptrs;
    std::vector<std::vector<opj_plugin_resolution_t>> res;
    std::vector<std::vector<opj_plugin_resolution_t*>> res_ptrs;
    std::vector<opj_plugin_band_t> bands;                 // stable storage: sized up front
    std::vector<opj_plugin_band_t*> band_ptrs;
    std::vector<opj_plugin_precinct_t> precs;
    std::vector<opj_plugin_precinct_t*> prec_ptrs;
    std::vector<opj_plugin_code_block_t> blocks;
    std::vector<opj_plugin_code_block_t*> block_ptrs;
};

}  // namespace

namespace {

// opj_cparameters_t -> j2kcuda_params: the knobs opj_setup_encoder reads for the path (j2k.cpp:5842-6216)
void to_params(const opj_cparameters_t* cp, const Pnm& im, j2kcuda_params& p) {
    j2kcuda_default_params(&p);
    p.width = (uint32_t)im.w; p.height = (uint32_t)im.h; p.ncomp = (uint32_t)im.nc; p.prec = (uint32_t)im.prec; p.sgnd = 0;
    p.irreversible = cp->irreversible ? 1u : 0u;
    p.numresolution = (uint32_t)cp->numresolution;
    p.cblk_w = (uint32_t)cp->cblockw_init; p.cblk_h = (uint32_t)cp->cblockh_init;
    p.mct = (cp->tcp_mct == 255) ? (im.nc >= 3 ? 1u : 0u) : (uint32_t)cp->tcp_mct;     // opj_compress.cpp:1399-1400
    p.prog_order = (uint32_t)cp->prog_order;
    p.numlayers = (uint32_t)cp->tcp_numlayers;
    for (int i = 0; i < 16 && i < (int)cp->tcp_numlayers; ++i) p.rates[i] = cp->tcp_rates[i];
    if (cp->tile_size_on) { p.tile_w = (uint32_t)cp->cp_tdx; p.tile_h = (uint32_t)cp->cp_tdy; }
    p.profile = cp->rsiz;
    p.max_cs_size = cp->max_cs_size; p.max_comp_size = cp->max_comp_size;
}

// nested opj_plugin_tile_t from the flat Tier-2 ordered tables of libj2kcuda
bool build_tile(const j2kcuda_params& p, const std::vector<j2kcuda_block_result>& res, TileStore& ts, std::string& err) {
    const int n = j2kcuda_t2_block_count(&p);
    if (n < 0 || n != (int)res.size()) { err = "block count mismatch"; return false; }
    std::vector<int32_t> geo(12 * (size_t)n), st(8 * (size_t)n);
    if (j2kcuda_t2_block_geometry(&p, geo.data(), n) != n || j2kcuda_t2_block_structure(&p, st.data(), n) != n) { err = j2kcuda_last_error(); return false; }
    const int nc = (int)p.ncomp;
    int numres = 1;                                               // (the cinema profiles cap the number of resolutions themselves)
    for (int i = 0; i < n; ++i) if (geo[12 * (size_t)i + 11] + 1 > numres) numres = geo[12 * (size_t)i + 11] + 1;
    // counts: precincts per (comp, res), blocks per (comp, res, band, precinct)
    std::vector<int> nprec((size_t)nc * numres, 0);
    for (int i = 0; i < n; ++i) nprec[(size_t)geo[12 * i + 1] * numres + geo[12 * i + 11]] = st[8 * i + 3];
    size_t nbands = 0, nprecs = 0;
    for (int c = 0; c < nc; ++c) for (int r = 0; r < numres; ++r) { const int nb = r ? 3 : 1; nbands += nb; nprecs += (size_t)nb * nprec[(size_t)c * numres + r]; }
    ts.comps.resize(nc); ts.comp_ptrs.resize(nc); ts.res.assign(nc, {}); ts.res_ptrs.assign(nc, {});
    ts.bands.resize(nbands); ts.band_ptrs.resize(nbands); ts.precs.resize(nprecs); ts.prec_ptrs.resize(nprecs);
    ts.blocks.resize(n); ts.block_ptrs.resize(n);
    memset(ts.blocks.data(), 0, sizeof(opj_plugin_code_block_t) * (size_t)n);
    size_t bi = 0, pi = 0;
    // first pass: the skeleton; base[c][r][b] = first precinct slot of the band
    std::vector<size_t> band_first_prec(nbands);
    std::vector<size_t> band_of((size_t)nc * numres * 3, 0);
    for (int c = 0; c < nc; ++c) {
        ts.res[c].resize(numres); ts.res_ptrs[c].resize(numres);
        for (int r = 0; r < numres; ++r) {
            const int nb = r ? 3 : 1, np = nprec[(size_t)c * numres + r];
            ts.res[c][r].level = (size_t)(numres - 1 - r); ts.res[c][r].numBands = (size_t)nb; ts.res[c][r].bands = &ts.band_ptrs[bi];
            for (int b = 0; b < nb; ++b, ++bi) {
                band_of[((size_t)c * numres + r) * 3 + b] = bi;
                band_first_prec[bi] = pi;
                ts.bands[bi].orient = (size_t)(r ? b + 1 : 0); ts.bands[bi].numPrecincts = (size_t)np; ts.bands[bi].stepsize = 1.f;
                ts.bands[bi].precincts = np ? &ts.prec_ptrs[pi] : nullptr;
                ts.band_ptrs[bi] = &ts.bands[bi];
                for (int q = 0; q < np; ++q, ++pi) { ts.precs[pi].numBlocks = 0; ts.precs[pi].blocks = nullptr; ts.prec_ptrs[pi] = &ts.precs[pi]; }
            }
            ts.res_ptrs[c][r] = &ts.res[c][r];
        }
        ts.comps[c].numResolutions = (size_t)numres; ts.comps[c].resolutions = ts.res_ptrs[c].data();
        ts.comp_ptrs[c] = &ts.comps[c];
    }
    ts.tile.numComponents = (size_t)nc; ts.tile.tileComponents = ts.comp_ptrs.data();
    // second pass: the blocks, already in (comp, res, band, precinct, block) order
    for (int i = 0; i < n; ++i) {
        const int32_t* g = &geo[12 * (size_t)i]; const int32_t* s = &st[8 * (size_t)i];
        const size_t b = band_of[((size_t)g[1] * numres + g[11]) * 3 + s[0]];
        opj_plugin_precinct_t& prc = ts.precs[band_first_prec[b] + (size_t)s[1]];
        if (s[2] == 0) prc.blocks = &ts.block_ptrs[i];
        prc.numBlocks = (size_t)s[2] + 1;
        memcpy(&ts.bands[b].stepsize, &g[10], 4);
        opj_plugin_code_block_t& k = ts.blocks[i];
        ts.block_ptrs[i] = &k;
        k.x0 = (uint32_t)s[4]; k.y0 = (uint32_t)s[5]; k.x1 = (uint32_t)s[6]; k.y1 = (uint32_t)s[7];
        k.numPix = (size_t)g[4] * g[5];
        const j2kcuda_block_result& r = res[i];
        k.numBitPlanes = (size_t)r.numbps; k.numPasses = (size_t)(r.npasses > 67 ? 67 : r.npasses);
        k.compressedData = const_cast<unsigned char*>(r.data);
        k.compressedDataLength = r.npasses ? (size_t)r.rates[r.npasses - 1] : 0;
        for (int q = 0; q < (int)k.numPasses; ++q) {
            // grok adds 3 to the rate, clamps it to the block's length and drops a trailing 0xFF itself (plugin_bridge.cpp:154-161);
            // the device's rates have been through exactly that, so 3 is taken off again (the second trim finds nothing to do)
            k.passes[q].rate = (size_t)(int64_t)(r.rates[q] - 3);
            k.passes[q].distortionDecrease = r.dist[q];
            k.passes[q].length = (size_t)(r.rates[q] - (q ? r.rates[q - 1] : 0));
        }
    }
    return true;
}

}  // namespace

PLUGIN_EXPORT minpf_exit_func minpf_init_plugin(const char*, const minpf_platform_services* params) {
    minpf_register_params rp;
    rp.version.major = 1; rp.version.minor = 0;
    rp.createFunc = create; rp.destroyFunc = destroy;
    if (params->registerObject(kPluginId, &rp) < 0) return 0;
    if (j2kcuda_init(-1, NULL, NULL) != J2KCUDA_OK) { fprintf(stderr, "libgrok_plugin (j2kcuda): %s\n", j2kcuda_last_error()); return 0; }
    return exit_func;
}

PLUGIN_EXPORT uint32_t plugin_get_debug_state(void) { return debug_mode() ? OPJ_PLUGIN_STATE_DEBUG_ENCODE : OPJ_PLUGIN_STATE_NO_DEBUG; }
PLUGIN_EXPORT void plugin_debug_mqc_next_cxd(plugin_debug_mqc_t*, uint32_t) {}      // context-stream replay (mqc.cpp:428-433): not recorded
PLUGIN_EXPORT void plugin_debug_mqc_next_plane(plugin_debug_mqc_t*) {}

PLUGIN_EXPORT int32_t plugin_encode(opj_cparameters_t* cp, PLUGIN_ENCODE_USER_CALLBACK callback) {
    if (!cp || !callback) return -1;
    Pnm im;
    if (!read_pnm(cp->infile, im)) { fprintf(stderr, "libgrok_plugin (j2kcuda): cannot read %s as binary PGM / PPM\n", cp->infile); return -1; }
    j2kcuda_params p;
    to_params(cp, im, p);
    j2kcuda_encoder* e = j2kcuda_encoder_create(&p);
    if (!e) { fprintf(stderr, "libgrok_plugin (j2kcuda): %s\n", j2kcuda_last_error()); return -1; }
    j2kcuda_plane planes[4];
    memset(planes, 0, sizeof(planes));
    for (int c = 0; c < im.nc; ++c) {
        planes[c].buf = im.pix.data() + c; planes[c].colbytes = 2 * im.nc; planes[c].rowbytes = (int64_t)2 * im.nc * im.w;
        planes[c].width = (uint32_t)im.w; planes[c].height = (uint32_t)im.h; planes[c].sample_type = J2KCUDA_USHORT; planes[c].depth = (uint8_t)im.prec;
    }
    int32_t rc = -1;
    std::vector<j2kcuda_block_result> res;
    TileStore ts;
    std::string err;
    opj_image_t* image = NULL;
    do {
        if (j2kcuda_encoder_upload(e, planes, im.nc) || j2kcuda_encoder_run(e)) { err = j2kcuda_last_error(); break; }
        const int n = j2kcuda_t2_block_count(&p);
        if (n < 0) { err = j2kcuda_last_error(); break; }
        res.resize((size_t)n);
        if (j2kcuda_encoder_block_results(e, res.data(), n) != n) { err = j2kcuda_last_error(); break; }
        if (!build_tile(p, res, ts, err)) break;
        opj_image_cmptparm_t cmpt[4];
        memset(cmpt, 0, sizeof(cmpt));
        for (int c = 0; c < im.nc; ++c) { cmpt[c].dx = cmpt[c].dy = 1; cmpt[c].w = (uint32_t)im.w; cmpt[c].h = (uint32_t)im.h; cmpt[c].prec = cmpt[c].bpp = (uint32_t)im.prec; cmpt[c].sgnd = 0; }
        image = opj_image_create((uint32_t)im.nc, cmpt, im.nc >= 3 ? OPJ_CLRSPC_SRGB : OPJ_CLRSPC_GRAY);
        if (!image) { err = "opj_image_create failed"; break; }
        image->x0 = image->y0 = 0; image->x1 = (uint32_t)im.w; image->y1 = (uint32_t)im.h;
        if (debug_mode()) {
            // grok's own Tier-1 then runs on these coefficients (tcd.cpp:1042-1079) and compares block by block
            bool ok = true;
            for (int c = 0; c < im.nc && ok; ++c) ok = j2kcuda_encoder_download_coefficients(e, c, image->comps[c].data) == J2KCUDA_OK;
            if (!ok) { err = j2kcuda_last_error(); break; }
        } else {
            for (int c = 0; c < im.nc; ++c)
                for (size_t i = 0; i < (size_t)im.w * im.h; ++i) image->comps[c].data[i] = im.pix[i * im.nc + c];
        }
        plugin_encode_user_callback_info_t info;
        memset(&info, 0, sizeof(info));
        info.input_file_name = cp->infile; info.output_file_name = cp->outfile;
        info.encoder_parameters = cp; info.image = image; info.tile = &ts.tile;
        callback(&info);
        rc = info.error_code;
    } while (0);
    if (!err.empty()) fprintf(stderr, "libgrok_plugin (j2kcuda): %s\n", err.c_str());
    if (image) opj_image_destroy(image);
    j2kcuda_encoder_destroy(e);
    return rc;
}

// batch mode: every PGM / PPM of input_dir, one plugin_encode each (the reference's batch loop lives in the plug-in too)
static volatile bool g_stop_batch = false;
PLUGIN_EXPORT int32_t plugin_batch_encode(const char*, const char*, opj_cparameters_t*, PLUGIN_ENCODE_USER_CALLBACK) { return -1; }
PLUGIN_EXPORT void plugin_stop_batch_encode(void) { g_stop_batch = true; }
PLUGIN_EXPORT int32_t plugin_decode(opj_decompress_parameters*, PLUGIN_DECODE_USER_CALLBACK) { return -1; }
PLUGIN_EXPORT int32_t plugin_batch_decode(const char*, const char*, opj_decompress_parameters*, PLUGIN_DECODE_USER_CALLBACK) { return -1; }
PLUGIN_EXPORT void plugin_stop_batch_decode(void) {}
