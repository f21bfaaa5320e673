/*
 * ref_shim.cpp -- TEST INFRASTRUCTURE ONLY.
 *
 * A thin C-ABI over the reference's own ext/grok sources (compiled where they lie under
 * /root/reference by oracle/Makefile into oracle/_ref/libgrokref.so; nothing from the
 * reference is copied into this repository).  It exposes
 *   - the whole codec through grok's public API (openjpeg.h) with in-memory streams, the way
 *     src/common/j2k_openjpeg_codec.cpp:590-758 (write) and :452-586 (read) drive it, and
 *   - the individual hot-path stage functions (mct.h, dwt.h, t1_opt.h) that have external
 *     linkage, so the CUDA kernels and the C restatement in j2k_oracle.c can be compared
 *     stage by stage.
 * Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) load it.
 */
#include "opj_includes.h"
#include <vector>
#include <cstdio>

extern "C" {

struct ref_params {
    int32_t irreversible;      /* 0 = 5/3 lossless, 1 = 9/7 */
    int32_t numresolution;     /* default 6 */
    int32_t cblockw, cblockh;  /* default 64 */
    int32_t tcp_mct;           /* 0 / 1 */
    int32_t prog_order;        /* OPJ_PROG_ORDER */
    int32_t tile_w, tile_h;    /* 0 = untiled */
    int32_t numlayers;         /* 0 -> 1 */
    int32_t disto_alloc;
    int32_t rsiz;              /* 0, 3 (cinema 2K), 4 (cinema 4K) */
    int32_t max_comp_size;
    int64_t max_cs_size;
    int32_t num_threads;
    double  rates[8];
};

struct MemOut { std::vector<uint8_t> buf; size_t pos = 0; };

static size_t mo_write(void* p, size_t n, void* ud) {
    MemOut* m = (MemOut*)ud;
    if (m->pos + n > m->buf.size()) m->buf.resize(m->pos + n);
    memcpy(m->buf.data() + m->pos, p, n);
    m->pos += n;
    return n;
}
static size_t mo_read(void* p, size_t n, void* ud) {
    MemOut* m = (MemOut*)ud;
    if (m->pos >= m->buf.size()) return (size_t)-1;
    size_t k = std::min(n, m->buf.size() - m->pos);
    memcpy(p, m->buf.data() + m->pos, k);
    m->pos += k;
    return k;
}
static int64_t mo_skip(int64_t n, void* ud) {
    MemOut* m = (MemOut*)ud;
    if (m->pos + n > m->buf.size()) m->buf.resize(m->pos + n);
    m->pos += (size_t)n;
    return n;
}
static bool mo_seek(int64_t n, void* ud) {
    MemOut* m = (MemOut*)ud;
    if ((size_t)n > m->buf.size()) m->buf.resize((size_t)n);
    m->pos = (size_t)n;
    return true;
}
static int g_verbose = 0;
static void msg_cb(const char* msg, void*) { if (g_verbose) fprintf(stderr, "[grok] %s", msg); }
static void quiet_cb(const char*, void*) {}

static bool g_init = false;
static void ensure_init() {
    if (!g_init) {
        /* grok prints "Unable to open library" through stdio when no plug-in is found; harmless */
        opj_initialize(NULL);
        g_init = true;
    }
}

void ref_set_verbose(int v) { g_verbose = v; }
/* JP2 wrapper for the next ref_encode calls: on = write a JP2 file instead of a raw codestream,
 * alpha_comp >= 0 marks that component as opacity (-> cdef box, jp2.cpp:2166-2230). */
static int g_jp2 = 0, g_alpha = -1;
void ref_set_jp2(int on, int alpha_comp) { g_jp2 = on; g_alpha = alpha_comp; }
const char* ref_version(void) { return opj_version(); }
/* Extra encoder switches for the next ref_encode calls: code-block style (opj_cparameters_t::mode: 1 LAZY, 2 RESET,
 * 4 TERMALL, 8 VSC, 16 PTERM, 32 SEGSYM; j2k.h:77-82), region of interest (component, up-shift), SOP / EPH markers (csty 2 / 4),
 * sub-sampling of components 1.. (dx, dy; component 0 stays 1x1). */
static int g_mode = 0, g_roi_comp = -1, g_roi_shift = 0, g_csty = 0, g_sub_dx = 1, g_sub_dy = 1;
void ref_set_extra(int mode, int roi_comp, int roi_shift, int csty, int sub_dx, int sub_dy) {
    g_mode = mode; g_roi_comp = roi_comp; g_roi_shift = roi_shift; g_csty = csty; g_sub_dx = sub_dx > 0 ? sub_dx : 1; g_sub_dy = sub_dy > 0 ? sub_dy : 1;
}

/* Whole-frame encode through grok's public API.  planes[c] = w*h int32 samples (unsigned
 * values, as CopyBuffer would leave them).  Returns 0 on success. */
int ref_encode(const int32_t* const* planes, int ncomp, int w, int h, int prec, int sgnd,
               const ref_params* p, uint8_t* out, size_t cap, size_t* out_len)
{
    ensure_init();
    opj_image_cmptparm_t cp[8];
    memset(cp, 0, sizeof(cp));
    for (int i = 0; i < ncomp; ++i) {
        cp[i].dx = i ? g_sub_dx : 1; cp[i].dy = i ? g_sub_dy : 1;
        cp[i].w = (w + cp[i].dx - 1) / cp[i].dx; cp[i].h = (h + cp[i].dy - 1) / cp[i].dy; cp[i].prec = prec;
        cp[i].bpp = prec; cp[i].sgnd = sgnd;
    }
    opj_image_t* image = opj_image_create(ncomp, cp, ncomp >= 3 ? OPJ_CLRSPC_SRGB : OPJ_CLRSPC_GRAY);
    if (!image) return 1;
    image->x0 = image->y0 = 0; image->x1 = w; image->y1 = h;
    for (int i = 0; i < ncomp; ++i) memcpy(image->comps[i].data, planes[i], sizeof(int32_t) * (size_t)cp[i].w * cp[i].h);
    if (g_jp2 && g_alpha >= 0 && g_alpha < ncomp) image->comps[g_alpha].alpha = 1;

    opj_cparameters_t params;
    opj_set_default_encoder_parameters(&params);
    params.irreversible = p->irreversible;
    if (p->numresolution > 0) params.numresolution = p->numresolution;
    if (p->cblockw > 0) params.cblockw_init = p->cblockw;
    if (p->cblockh > 0) params.cblockh_init = p->cblockh;
    params.tcp_mct = (uint8_t)p->tcp_mct;
    params.prog_order = (OPJ_PROG_ORDER)p->prog_order;
    params.tcp_numlayers = p->numlayers;
    params.cp_disto_alloc = p->disto_alloc;
    for (int i = 0; i < 8; ++i) params.tcp_rates[i] = p->rates[i];
    if (p->tile_w > 0) {
        params.tile_size_on = true; params.cp_tx0 = params.cp_ty0 = 0;
        params.cp_tdx = p->tile_w; params.cp_tdy = p->tile_h;
    }
    if (p->rsiz) {
        params.rsiz = (uint16_t)p->rsiz;
        params.max_comp_size = p->max_comp_size;
        params.max_cs_size = p->max_cs_size;
    }
    params.numThreads = p->num_threads > 0 ? p->num_threads : 1;
    params.mode = g_mode;
    params.csty |= g_csty;
    if (g_roi_comp >= 0) { params.roi_compno = g_roi_comp; params.roi_shift = g_roi_shift; }

    int rc = 2;
    MemOut mo;
    opj_stream_t* stream = opj_stream_create(1 << 20, false);
    opj_codec_t* codec = opj_create_compress(g_jp2 ? OPJ_CODEC_JP2 : OPJ_CODEC_J2K);
    if (stream && codec) {
        opj_stream_set_user_data(stream, &mo, NULL);
        opj_stream_set_read_function(stream, mo_read);
        opj_stream_set_write_function(stream, mo_write);
        opj_stream_set_skip_function(stream, mo_skip);
        opj_stream_set_seek_function(stream, mo_seek);
        opj_set_error_handler(codec, msg_cb, NULL);
        opj_set_warning_handler(codec, msg_cb, NULL);
        opj_set_info_handler(codec, quiet_cb, NULL);
        if (opj_setup_encoder(codec, &params, image) && opj_start_compress(codec, image, stream) &&
            opj_encode(codec, stream) && opj_end_compress(codec, stream))
            rc = 0;
    }
    if (stream) opj_stream_destroy(stream);
    if (codec) opj_destroy_codec(codec);
    opj_image_destroy(image);
    if (rc == 0) {
        *out_len = mo.buf.size();
        if (mo.buf.size() > cap) return 3;
        memcpy(out, mo.buf.data(), mo.buf.size());
    }
    return rc;
}

static int g_last_dx = 1, g_last_dy = 1;
void ref_last_subsampling(int* dx, int* dy) { *dx = g_last_dx; *dy = g_last_dy; }   /* of component 1 in the last header read */
/* Whole-stream decode.  planes[c] must hold w*h int32 (caller sized them from ref_probe; sub-sampled components hold
 * ceil(w / dx) * ceil(h / dy)). */
int ref_decode(const uint8_t* data, size_t len, int reduce, int nthreads,
               int32_t* const* planes, int max_planes, int* info /* w,h,ncomp,prec,sgnd */)
{
    ensure_init();
    int rc = 2;
    opj_stream_t* stream = opj_stream_create_buffer_stream((uint8_t*)data, len, true);
    bool jp2 = len >= 12 && data[4] == 'j' && data[5] == 'P';
    opj_codec_t* codec = opj_create_decompress(jp2 ? OPJ_CODEC_JP2 : OPJ_CODEC_J2K);
    opj_image_t* image = NULL;
    if (stream && codec) {
        opj_set_error_handler(codec, msg_cb, NULL);
        opj_set_warning_handler(codec, msg_cb, NULL);
        opj_set_info_handler(codec, quiet_cb, NULL);
        opj_dparameters_t dp;
        opj_set_default_decoder_parameters(&dp);
        dp.cp_reduce = reduce;
        dp.numThreads = nthreads > 0 ? nthreads : 1;
        if (opj_setup_decoder(codec, &dp) && opj_read_header(stream, codec, &image)) {
            info[0] = image->comps[0].w; info[1] = image->comps[0].h; info[2] = image->numcomps;
            info[3] = image->comps[0].prec; info[4] = image->comps[0].sgnd;
            g_last_dx = image->numcomps > 1 ? image->comps[1].dx : 1; g_last_dy = image->numcomps > 1 ? image->comps[1].dy : 1;
            if (!planes) rc = 0;
            else if (opj_decode(codec, stream, image) && opj_end_decompress(codec, stream)) {
                rc = 0;
                info[0] = image->comps[0].w; info[1] = image->comps[0].h;
                /* Reference quirk: a single-tile image decoded at reduced resolution is handed out as the
                   tile buffer itself (j2k.cpp:7685-7695), whose rows keep the FULL-resolution pitch
                   although comps[c].w is the reduced width; read it with the pitch it really has. */
                opj_codestream_info_v2_t* ci = opj_get_cstr_info(codec);
                bool single = ci && ci->tw * ci->th == 1;
                if (ci) opj_destroy_cstr_info(&ci);
                for (int c = 0; c < (int)image->numcomps && c < max_planes; ++c) {
                    const opj_image_comp_t& ic = image->comps[c];
                    size_t pitch = ic.w;
                    if (single && reduce > 0) pitch = (image->x1 - image->x0 + ic.dx - 1) / ic.dx;
                    for (uint32_t y = 0; y < ic.h; ++y)
                        memcpy(planes[c] + (size_t)y * ic.w, ic.data + (size_t)y * pitch, sizeof(int32_t) * ic.w);
                }
            }
        }
    }
    if (image) opj_image_destroy(image);
    if (stream) opj_stream_destroy(stream);
    if (codec) opj_destroy_codec(codec);
    return rc;
}

/* ---- stage functions ------------------------------------------------------------------ */
void ref_mct_encode(int32_t* c0, int32_t* c1, int32_t* c2, uint64_t n) { opj_mct_encode(c0, c1, c2, n); }
void ref_mct_decode(int32_t* c0, int32_t* c1, int32_t* c2, uint64_t n) { opj_mct_decode(c0, c1, c2, n); }
void ref_mct_encode_real(int32_t* c0, int32_t* c1, int32_t* c2, uint64_t n) { opj_mct_encode_real(c0, c1, c2, n); }
void ref_mct_decode_real(float* c0, float* c1, float* c2, uint64_t n) { opj_mct_decode_real(c0, c1, c2, n); }

/* Builds the minimal opj_tcd_tilecomp_t the DWT routines read: bounds of every resolution
 * (tcd.cpp:598-607: ceildivpow2 of the tile-component bounds) and a tile buffer that aliases
 * `data` with stride = tile-component width. */
struct TileComp {
    opj_tcd_tilecomp_t tc;
    std::vector<opj_tcd_resolution_t> res;
    opj_tile_buf_component_t buf;
    TileComp(int32_t* data, int x0, int y0, int x1, int y1, int numres) : res(numres) {
        memset(&tc, 0, sizeof(tc));
        tc.x0 = x0; tc.y0 = y0; tc.x1 = x1; tc.y1 = y1;
        tc.numresolutions = numres; tc.minimum_num_resolutions = numres;
        for (int r = 0; r < numres; ++r) {
            int lvl = numres - 1 - r;
            memset(&res[r], 0, sizeof(opj_tcd_resolution_t));
            res[r].x0 = opj_uint_ceildivpow2(x0, lvl); res[r].y0 = opj_uint_ceildivpow2(y0, lvl);
            res[r].x1 = opj_uint_ceildivpow2(x1, lvl); res[r].y1 = opj_uint_ceildivpow2(y1, lvl);
        }
        tc.resolutions = res.data();
        buf.data = data; buf.owns_data = false;
        buf.data_size = buf.data_size_needed = 0;
        opj_rect_init(&buf.dim, x0, y0, x1, y1);
        opj_rect_init(&buf.tile_dim, x0, y0, x1, y1);
        tc.buf = &buf;
    }
};

int ref_dwt_encode(int32_t* data, int x0, int y0, int x1, int y1, int numres, int real)
{
    TileComp t(data, x0, y0, x1, y1, numres);
    bool ok = real ? opj_dwt_encode_real(&t.tc) : opj_dwt_encode(&t.tc);
    return ok ? 0 : 1;
}
/* real: data holds float32 bit patterns in and out (dwt.cpp:927-1062) */
int ref_dwt_decode(int32_t* data, int x0, int y0, int x1, int y1, int numres, int real, int nthreads)
{
    TileComp t(data, x0, y0, x1, y1, numres);
    bool ok = real ? opj_dwt_decode_real(&t.tc, numres, nthreads > 0 ? nthreads : 1)
                   : opj_dwt_decode(&t.tc, numres, nthreads > 0 ? nthreads : 1);
    return ok ? 0 : 1;
}
double ref_dwt_getnorm(int level, int orient, int real) {
    return real ? opj_dwt_getnorm_real(level, orient) : opj_dwt_getnorm(level, orient);
}
double ref_mct_getnorm(int compno, int real) { return real ? opj_mct_getnorm_real(compno) : opj_mct_getnorm(compno); }

/* One code-block through the reference's quantiser (T1Encoder.cpp:103-174) and packed-flag
 * tier-1 encoder (t1_opt.cpp:659-739).  `coeff` = w*h DWT coefficients with row stride
 * `stride`.  Outputs: bytes, numbps, passes, per pass cumulative rate and distortion. */
int ref_t1_encode_cblk(const int32_t* coeff, int w, int h, int stride, int orient, int compno,
                       int level, int qmfbid, double stepsize, int use_mct_norms, int numcomps,
                       uint8_t* out, int cap, int* numbps, int* npasses, int* rates /*[>=100]*/,
                       double* dist /*[>=100]*/)
{
    opj_t1_opt_t* t1 = opj_t1_opt_create(true);
    if (!t1 || !opj_t1_opt_allocate_buffers(t1, w, h)) return 1;
    opj_t1_opt_init_buffers(t1, w, h);
    uint32_t max = 0;
    int32_t bandconst = qmfbid == 1 ? 0 : 8192 * 8192 / ((int32_t)floor(((float)stepsize) * 8192));
    for (int j = 0; j < h; ++j)
        for (int i = 0; i < w; ++i) {
            int32_t tmp = coeff[j * stride + i];
            if (qmfbid == 1) tmp *= (1 << T1_NMSEDEC_FRACBITS);
            else tmp = opj_int_fix_mul_t1(tmp, bandconst);
            uint32_t mag = (uint32_t)opj_int_abs(tmp);
            if (mag > max) max = mag;
            t1->data[j * w + i] = mag | ((uint32_t)(tmp < 0) << T1_DATA_SIGN_BIT_INDEX);
        }
    opj_tcd_cblk_enc_t cblk;
    memset(&cblk, 0, sizeof(cblk));
    std::vector<uint8_t> data((size_t)w * h * 4 + 16);
    std::vector<opj_tcd_pass_t> passes(100);
    cblk.data = data.data() + 1;
    cblk.passes = passes.data();
    cblk.x0 = 0; cblk.y0 = 0; cblk.x1 = w; cblk.y1 = h;
    const double* norms = NULL;
    if (use_mct_norms) norms = qmfbid == 1 ? opj_mct_get_mct_norms() : opj_mct_get_mct_norms_real();
    opj_t1_opt_encode_cblk(t1, &cblk, orient, compno, level, qmfbid, (float)stepsize, numcomps,
                           norms, 3, max);
    *numbps = cblk.numbps;
    *npasses = cblk.totalpasses;
    int total = 0;
    for (uint32_t i = 0; i < cblk.totalpasses; ++i) {
        rates[i] = passes[i].rate; dist[i] = passes[i].distortiondec;
        total = passes[i].rate;
    }
    if (total > cap) { opj_t1_opt_destroy(t1); return 2; }
    memcpy(out, cblk.data, total);
    opj_t1_opt_destroy(t1);
    return 0;
}

/* One code-block through the reference's generic decoder (t1.cpp:1337-1448): `data` = the block's bytes, made of nsegs
 * segments (seg_len bytes, seg_passes passes each; with cblksty 0 there is one).  out = w*h values in the decoder's
 * representation (one extra fractional bit), before the ROI down-shift and dequantisation of T1Decoder.cpp:59-99. */
int ref_t1_decode_cblk(const uint8_t* data, const int* seg_len, const int* seg_passes, int nsegs, int w, int h, int orient,
                       int numbps, int roishift, int cblksty, int32_t* out)
{
    opj_t1_t* t1 = opj_t1_create(false, (uint16_t)w, (uint16_t)h);
    if (!t1) return 1;
    opj_tcd_cblk_dec_t cblk;
    memset(&cblk, 0, sizeof(cblk));
    opj_vec_init(&cblk.seg_buffers, true);
    std::vector<opj_tcd_seg_t> segs(nsegs > 0 ? nsegs : 1);
    memset(segs.data(), 0, sizeof(opj_tcd_seg_t) * segs.size());
    size_t pos = 0;
    for (int i = 0; i < nsegs; ++i) {
        segs[i].dataindex = (uint32_t)pos; segs[i].len = seg_len[i];
        segs[i].numpasses = segs[i].real_num_passes = seg_passes[i]; segs[i].maxpasses = 109;
        pos += seg_len[i];
    }
    /* one buffer for the whole block keeps opj_min_buf's 16-bit lengths out of the way for blocks beyond 64 KB */
    if (pos) opj_min_buf_vec_push_back(&cblk.seg_buffers, (uint8_t*)data, (uint16_t)(pos > 65535 ? 65535 : pos));
    cblk.segs = segs.data();
    cblk.x0 = 0; cblk.y0 = 0; cblk.x1 = w; cblk.y1 = h;
    cblk.numbps = numbps; cblk.numsegs = cblk.real_num_segs = nsegs;
    memset(out, 0, sizeof(int32_t) * (size_t)w * h);
    bool ok = pos <= 65535 && opj_t1_decode_cblk(t1, &cblk, orient, roishift, cblksty);
    if (ok && pos) memcpy(out, t1->data, sizeof(int32_t) * (size_t)w * h);      /* the decoder's block buffer is contiguous (T1Decoder.cpp:77-99) */
    opj_vec_cleanup(&cblk.seg_buffers);
    opj_t1_destroy(t1);
    return ok ? 0 : 2;
}

/* One code-block through the reference's GENERIC encoder (t1.cpp:1537-1673; the one used when cblksty != 0): quantised like
 * ref_t1_encode_cblk.  Outputs per pass: cumulative rate, length and term flag; the bytes. */
int ref_t1_encode_cblk_generic(const int32_t* coeff, int w, int h, int stride, int orient, int cblksty, uint8_t* out, int cap,
                               int* numbps, int* npasses, int* rates, int* lens, int* terms)
{
    opj_t1_t* t1 = opj_t1_create(true, 0, 0);
    if (!t1 || !opj_t1_allocate_buffers(t1, w, h)) return 1;
    std::vector<int32_t> plane((size_t)w * h);
    t1->data = plane.data();                              /* the encoder codes straight from the tile buffer (t1.cpp:1496): here a copy */
    t1->data_stride = w;
    uint32_t max = 0;
    for (int j = 0; j < h; ++j)
        for (int i = 0; i < w; ++i) {
            int32_t tmp = coeff[j * stride + i] * (1 << T1_NMSEDEC_FRACBITS);
            t1->data[j * w + i] = tmp;
            uint32_t mag = (uint32_t)opj_int_abs(tmp);
            if (mag > max) max = mag;
        }
    opj_tcd_cblk_enc_t cblk;
    memset(&cblk, 0, sizeof(cblk));
    std::vector<uint8_t> data((size_t)w * h * 8 + 64);
    std::vector<opj_tcd_pass_t> passes(110);
    cblk.data = data.data() + 1;
    cblk.passes = passes.data();
    cblk.x0 = 0; cblk.y0 = 0; cblk.x1 = w; cblk.y1 = h;
    opj_t1_encode_cblk(t1, &cblk, orient, 0, 0, 1, 1.0, cblksty, 1, NULL, 0);
    *numbps = cblk.numbps; *npasses = cblk.totalpasses;
    int total = 0;
    for (uint32_t i = 0; i < cblk.totalpasses; ++i) {
        rates[i] = passes[i].rate; lens[i] = passes[i].len; terms[i] = passes[i].term;
        total = passes[i].rate;
    }
    if (total > cap) { opj_t1_destroy(t1); return 2; }
    memcpy(out, cblk.data, total);
    opj_t1_destroy(t1);
    return 0;
}

/* ---- the accelerator path of the reference (openjpeg.cpp:1077-1160): load a plug-in, let it encode `infile`, and finish the
 * codestream with opj_encode_with_plugin in the callback, as bin/jp2/opj_compress.cpp:1526-1630 does.  Returns 0 and the bytes. */
static MemOut* g_plugin_out = NULL;
static int g_plugin_rc = 0;
static void plugin_cb(opj_plugin_encode_user_callback_info_t* info) {
    g_plugin_rc = 10;
    opj_codec_t* codec = opj_create_compress(OPJ_CODEC_J2K);
    opj_stream_t* stream = opj_stream_create(1 << 20, false);
    if (codec && stream) {
        opj_stream_set_user_data(stream, g_plugin_out, NULL);
        opj_stream_set_read_function(stream, mo_read);
        opj_stream_set_write_function(stream, mo_write);
        opj_stream_set_skip_function(stream, mo_skip);
        opj_stream_set_seek_function(stream, mo_seek);
        opj_set_error_handler(codec, msg_cb, NULL);
        opj_set_warning_handler(codec, msg_cb, NULL);
        opj_set_info_handler(codec, quiet_cb, NULL);
        if (opj_setup_encoder(codec, info->encoder_parameters, info->image) && opj_start_compress(codec, info->image, stream) &&
            opj_encode_with_plugin(codec, info->tile, stream) && opj_end_compress(codec, stream))
            g_plugin_rc = 0;
    }
    if (stream) opj_stream_destroy(stream);
    if (codec) opj_destroy_codec(codec);
    info->error_code = g_plugin_rc;
}
int ref_encode_with_plugin(const char* plugin_path, const char* infile, const ref_params* p, uint8_t* out, size_t cap, size_t* out_len)
{
    static bool loaded = false;
    if (!loaded) {
        opj_plugin_init_info_t pi;
        pi.plugin_path = plugin_path;
        if (!opj_plugin_init(pi)) return 1;
        loaded = true;
    }
    opj_cparameters_t params;
    opj_set_default_encoder_parameters(&params);
    params.irreversible = p->irreversible;
    if (p->numresolution > 0) params.numresolution = p->numresolution;
    if (p->cblockw > 0) params.cblockw_init = p->cblockw;
    if (p->cblockh > 0) params.cblockh_init = p->cblockh;
    params.tcp_mct = (uint8_t)p->tcp_mct;
    params.prog_order = (OPJ_PROG_ORDER)p->prog_order;
    params.tcp_numlayers = p->numlayers;
    params.cp_disto_alloc = p->disto_alloc;
    for (int i = 0; i < 8; ++i) params.tcp_rates[i] = p->rates[i];
    if (p->rsiz) { params.rsiz = (uint16_t)p->rsiz; params.max_comp_size = p->max_comp_size; params.max_cs_size = p->max_cs_size; }
    params.numThreads = p->num_threads > 0 ? p->num_threads : 1;
    strncpy(params.infile, infile, sizeof(params.infile) - 1);
    params.outfile[0] = 0;
    params.cod_format = 0;
    MemOut mo;
    g_plugin_out = &mo;
    g_plugin_rc = 11;
    int32_t rc = opj_plugin_encode(&params, plugin_cb);
    g_plugin_out = NULL;
    if (rc || g_plugin_rc) return 20 + (rc ? 1 : 0) + g_plugin_rc;
    *out_len = mo.buf.size();
    if (mo.buf.size() > cap) return 3;
    memcpy(out, mo.buf.data(), mo.buf.size());
    return 0;
}

} /* extern "C" */
