// JP2 file wrapper: box reader and writer (host only).
// Reference behaviour: ext/grok/src/lib/openjp2/jp2.cpp -- box header :476-558, ihdr :560-644 / :646-698,
// res :703-820, bpcc :909-975, cdef :977-1025 / :1539-1632, colr :1027-1110 / :1634-1816, pclr :1317-1419,
// cmap :1421-1537, jp2h :1817-1925 / :2754-2824, ftyp :1927-1974 / :2668-2752, jp2c :1976-2011, jp :2013-2056 /
// :2621-2666, encoder defaults :2058-2262.  What the plug-in reads from it: j2k_openjpeg_codec.cpp:303-408.
#include "../../include/j2kcuda.h"
#include "common.h"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
using namespace j2k;

namespace {

constexpr uint32_t fourcc(char a, char b, char c, char d) {
    return ((uint32_t)(uint8_t)a << 24) | ((uint32_t)(uint8_t)b << 16) | ((uint32_t)(uint8_t)c << 8) | (uint32_t)(uint8_t)d;
}
constexpr uint32_t BOX_JP = fourcc('j', 'P', ' ', ' '), BOX_FTYP = fourcc('f', 't', 'y', 'p'), BOX_JP2H = fourcc('j', 'p', '2', 'h'),
                   BOX_IHDR = fourcc('i', 'h', 'd', 'r'), BOX_BPCC = fourcc('b', 'p', 'c', 'c'), BOX_COLR = fourcc('c', 'o', 'l', 'r'),
                   BOX_CDEF = fourcc('c', 'd', 'e', 'f'), BOX_PCLR = fourcc('p', 'c', 'l', 'r'), BOX_CMAP = fourcc('c', 'm', 'a', 'p'),
                   BOX_RES = fourcc('r', 'e', 's', ' '), BOX_RESC = fourcc('r', 'e', 's', 'c'), BOX_RESD = fourcc('r', 'e', 's', 'd'),
                   BOX_JP2C = fourcc('j', 'p', '2', 'c'), BRAND_JP2 = fourcc('j', 'p', '2', ' ');

inline uint32_t rd(const uint8_t* p, int nbytes) {
    uint32_t v = 0;
    for (int i = 0; i < nbytes; ++i) v = (v << 8) | p[i];
    return v;
}

struct Box { uint32_t type; size_t body, len; };   // body offset and body length inside the buffer

// one box header at pos (jp2.cpp:476-558): LBox 0 = to the end of the file, 1 = 64-bit XLBox follows
bool read_box(const uint8_t* d, size_t end, size_t pos, Box& b, size_t& next) {
    if (pos + 8 > end) return false;
    uint64_t lbox = rd(d + pos, 4);
    b.type = rd(d + pos + 4, 4);
    size_t hdr = 8;
    if (lbox == 1) {
        if (pos + 16 > end) return false;
        lbox = ((uint64_t)rd(d + pos + 8, 4) << 32) | rd(d + pos + 12, 4);
        hdr = 16;
    } else if (lbox == 0) lbox = end - pos;
    if (lbox < hdr || lbox > end - pos) return false;
    b.body = pos + hdr; b.len = (size_t)lbox - hdr;
    next = pos + (size_t)lbox;
    return true;
}

double res_value(uint32_t num, uint32_t den, int exponent) {    // jp2.cpp:696-701
    if (den == 0) return 0.0;
    return ((double)num / den) * std::pow(10.0, exponent);
}

}  // namespace

extern "C" {

int j2kcuda_probe_jp2(const uint8_t* d, size_t n, j2kcuda_jp2_info* out) { return j2k::guarded([&]() -> int {
    if (!d || !out) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    memset(out, 0, sizeof(*out));
    if (n >= 2 && d[0] == 0xff && d[1] == 0x4f) {                 // raw codestream (SOC)
        out->codestream_offset = 0; out->codestream_length = n;
        return J2KCUDA_OK;
    }
    Box b; size_t pos = 0, next = 0;
    // signature box first, then ftyp (jp2.cpp:2621-2752)
    if (!read_box(d, n, 0, b, next) || b.type != BOX_JP || b.len != 4 || rd(d + b.body, 4) != 0x0d0a870au)
        return set_error(J2KCUDA_ERR_STREAM, "not a JPEG 2000 codestream or JP2 file");
    out->is_jp2 = 1;
    pos = next;
    bool have_ftyp = false, have_jp2h = false, have_ihdr = false, have_colr = false, have_bpcc = false;
    while (pos < n && read_box(d, n, pos, b, next)) {
        if (b.type == BOX_FTYP) {
            if (b.len < 8 || (b.len & 3)) return set_error(J2KCUDA_ERR_STREAM, "bad ftyp box");
            out->brand = rd(d + b.body, 4); out->minversion = rd(d + b.body + 4, 4);
            have_ftyp = true;
        } else if (b.type == BOX_JP2H) {
            if (!have_ftyp) return set_error(J2KCUDA_ERR_STREAM, "jp2h box before ftyp");
            have_jp2h = true;
            const size_t end = b.body + b.len;
            size_t p = b.body, pn = 0;
            Box s;
            while (p < end) {
                if (!read_box(d, end, p, s, pn)) return set_error(J2KCUDA_ERR_STREAM, "truncated box inside jp2h");
                const uint8_t* q = d + s.body;
                if (s.type == BOX_IHDR) {                               // jp2.cpp:560-644
                    if (s.len != 14) return set_error(J2KCUDA_ERR_STREAM, "bad ihdr box");
                    out->height = rd(q, 4); out->width = rd(q + 4, 4); out->ncomp = rd(q + 8, 2);
                    out->bpc = q[10];
                    if (q[11] != 7) return set_error(J2KCUDA_ERR_STREAM, "ihdr: compression type is not 7");
                    out->unk_c = q[12]; out->ipr = q[13];
                    if (out->ncomp == 0 || out->ncomp > 16384) return set_error(J2KCUDA_ERR_STREAM, "ihdr: bad component count");
                    have_ihdr = true;
                } else if (s.type == BOX_BPCC) {                        // jp2.cpp:945-975
                    if (!have_ihdr || s.len != out->ncomp) return set_error(J2KCUDA_ERR_STREAM, "bad bpcc box");
                    for (uint32_t i = 0; i < out->ncomp && i < J2KCUDA_JP2_MAX_CHANNELS; ++i) out->bpcc[i] = q[i];
                    have_bpcc = true;
                } else if (s.type == BOX_COLR) {                        // jp2.cpp:1634-1816: only the first colr box counts
                    if (s.len < 3) return set_error(J2KCUDA_ERR_STREAM, "bad colr box");
                    if (!have_colr) {
                        const uint32_t meth = q[0];
                        out->precedence = q[1]; out->approx = q[2];
                        if (meth == 1) {
                            if (s.len < 7) return set_error(J2KCUDA_ERR_STREAM, "bad colr box (enumerated)");
                            out->colr_method = 1; out->enumcs = rd(q + 3, 4);
                            have_colr = true;
                        } else if (meth == 2 || meth == 3) {
                            out->colr_method = meth;
                            out->icc_offset = s.body + 3; out->icc_length = s.len - 3;
                            have_colr = true;
                        } else if (meth > 3) {
                            out->colr_method = 4;                       // vendor / reserved: box ignored, as the reference does
                        }
                    }
                } else if (s.type == BOX_CDEF) {                        // jp2.cpp:1539-1632
                    if (s.len < 2) return set_error(J2KCUDA_ERR_STREAM, "bad cdef box");
                    const uint32_t cnt = rd(q, 2);
                    if (cnt == 0 || s.len < 2 + 6 * (size_t)cnt) return set_error(J2KCUDA_ERR_STREAM, "bad cdef box");
                    if (cnt > J2KCUDA_JP2_MAX_CHANNELS) return set_error(J2KCUDA_ERR_UNSUPPORTED, "cdef box with more than 16 channels");
                    out->cdef_n = cnt;
                    for (uint32_t i = 0; i < cnt; ++i) {
                        out->cdef_cn[i] = (uint16_t)rd(q + 2 + 6 * i, 2);
                        out->cdef_typ[i] = (uint16_t)rd(q + 4 + 6 * i, 2);
                        out->cdef_asoc[i] = (uint16_t)rd(q + 6 + 6 * i, 2);
                    }
                } else if (s.type == BOX_PCLR) {                        // jp2.cpp:1317-1419
                    if (s.len < 3) return set_error(J2KCUDA_ERR_STREAM, "bad pclr box");
                    const uint32_t ne = rd(q, 2), nc = q[2];
                    if (ne == 0 || ne > 1024 || nc == 0) return set_error(J2KCUDA_ERR_STREAM, "bad pclr box");
                    if (nc > J2KCUDA_JP2_MAX_CHANNELS) return set_error(J2KCUDA_ERR_UNSUPPORTED, "palette with more than 16 columns");
                    if (s.len < 3 + (size_t)nc) return set_error(J2KCUDA_ERR_STREAM, "bad pclr box");
                    size_t rowbytes = 0;
                    for (uint32_t c = 0; c < nc; ++c) { out->pclr_depth[c] = q[3 + c]; rowbytes += ((q[3 + c] & 0x7fu) + 1 + 7) / 8; }
                    if (s.len < 3 + (size_t)nc + rowbytes * ne) return set_error(J2KCUDA_ERR_STREAM, "truncated pclr box");
                    out->pclr_entries = ne; out->pclr_channels = nc; out->pclr_offset = s.body + 3 + nc;
                } else if (s.type == BOX_CMAP) {                        // jp2.cpp:1421-1537: needs the palette first
                    if (!out->pclr_entries) return set_error(J2KCUDA_ERR_STREAM, "cmap box without a preceding pclr box");
                    const uint32_t cnt = out->pclr_channels;
                    if (s.len < 4 * (size_t)cnt) return set_error(J2KCUDA_ERR_STREAM, "bad cmap box");
                    out->cmap_n = cnt;
                    for (uint32_t i = 0; i < cnt; ++i) {
                        out->cmap_cmp[i] = (uint16_t)rd(q + 4 * i, 2); out->cmap_mtyp[i] = q[4 * i + 2]; out->cmap_pcol[i] = q[4 * i + 3];
                    }
                } else if (s.type == BOX_RES) {                         // jp2.cpp:703-820
                    size_t r = s.body, rn = 0; Box t;
                    while (r < s.body + s.len && read_box(d, s.body + s.len, r, t, rn)) {
                        if ((t.type == BOX_RESC || t.type == BOX_RESD) && t.len == 10) {
                            const uint8_t* z = d + t.body;
                            double* dst = t.type == BOX_RESC ? out->capture_res : out->display_res;
                            dst[0] = res_value(rd(z, 2), rd(z + 2, 2), (int8_t)z[8]);      // vertical
                            dst[1] = res_value(rd(z + 4, 2), rd(z + 6, 2), (int8_t)z[9]);  // horizontal
                            (t.type == BOX_RESC ? out->has_capture_res : out->has_display_res) = 1;
                        }
                        r = rn;
                    }
                }
                p = pn;
            }
            if (!have_ihdr) return set_error(J2KCUDA_ERR_STREAM, "jp2h box without ihdr");
            if (out->bpc == 255 && !have_bpcc) return set_error(J2KCUDA_ERR_STREAM, "ihdr says per-component depths but there is no bpcc box");
            if (!have_bpcc) for (uint32_t i = 0; i < out->ncomp && i < J2KCUDA_JP2_MAX_CHANNELS; ++i) out->bpcc[i] = out->bpc;
        } else if (b.type == BOX_JP2C) {
            if (!have_jp2h) return set_error(J2KCUDA_ERR_STREAM, "jp2c box before jp2h");
            out->codestream_offset = b.body; out->codestream_length = b.len;
            return J2KCUDA_OK;
        }
        pos = next;
    }
    return set_error(J2KCUDA_ERR_STREAM, "JP2 file without a codestream box");
}); }

int j2kcuda_jp2_palette(const uint8_t* d, size_t n, uint32_t* entries, size_t cap) { return j2k::guarded([&]() -> int {
    j2kcuda_jp2_info info;
    const int rc = j2kcuda_probe_jp2(d, n, &info);
    if (rc) return -rc;
    if (!entries) { set_error(J2KCUDA_ERR_PARAM, "null argument"); return -J2KCUDA_ERR_PARAM; }
    const size_t total = (size_t)info.pclr_entries * info.pclr_channels;
    if (total > cap) { set_error(J2KCUDA_ERR_PARAM, "palette buffer too small"); return -J2KCUDA_ERR_PARAM; }
    const uint8_t* q = d + info.pclr_offset;
    for (uint32_t e = 0; e < info.pclr_entries; ++e)
        for (uint32_t c = 0; c < info.pclr_channels; ++c) {
            const int nb = (int)((info.pclr_depth[c] & 0x7fu) + 1 + 7) / 8;
            entries[(size_t)e * info.pclr_channels + c] = rd(q, nb > 4 ? 4 : nb);
            q += nb;
        }
    return (int)total;
}); }

int j2kcuda_wrap_jp2(const uint8_t* cs, size_t len, const j2kcuda_jp2_params* p, uint8_t** out, size_t* out_len) { return j2k::guarded([&]() -> int {
    if (!cs || !p || !out || !out_len) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    if (p->ncomp == 0 || p->ncomp > J2KCUDA_JP2_MAX_CHANNELS) return set_error(J2KCUDA_ERR_PARAM, "1..16 components");
    if (len > 0xffffffffull - 8) return set_error(J2KCUDA_ERR_UNSUPPORTED, "codestream too long for a 32-bit box length");
    std::vector<uint8_t> v;
    auto put = [&](uint32_t x, int nbytes) { for (int i = nbytes - 1; i >= 0; --i) v.push_back((uint8_t)(x >> (8 * i))); };
    // signature and file type (jp2.cpp:2013-2056, 1927-1974; brand 'jp2 ', MinV 0, one compatibility entry)
    put(12, 4); put(BOX_JP, 4); put(0x0d0a870au, 4);
    put(20, 4); put(BOX_FTYP, 4); put(BRAND_JP2, 4); put(0, 4); put(BRAND_JP2, 4);
    // header super-box (jp2.cpp:1817-1925): ihdr, [bpcc], colr, [cdef]
    uint32_t bpc = (p->prec[0] - 1) + (p->sgnd[0] << 7);                    // jp2.cpp:2115-2124
    for (uint32_t i = 1; i < p->ncomp; ++i) if (p->prec[i] != p->prec[0]) bpc = 255;
    uint32_t enumcs = p->enumcs;
    const bool icc = p->icc && p->icc_len;
    if (icc) enumcs = 0;
    // automatic cdef for exactly one opacity channel behind the colour channels (jp2.cpp:2166-2230)
    uint32_t color_channels = 0;
    bool cdef = false;
    if (p->alpha_comp >= 0 && (uint32_t)p->alpha_comp < p->ncomp) {
        if (enumcs == 16 || enumcs == 18) color_channels = 3;
        else if (enumcs == 17) color_channels = 1;
        cdef = color_channels && p->ncomp >= color_channels + 1 && (uint32_t)p->alpha_comp >= color_channels;
    }
    const uint32_t colr_len = 11 + (icc ? (uint32_t)p->icc_len : 4);
    const uint32_t jp2h_len = 8 + 22 + (bpc == 255 ? 8 + p->ncomp : 0) + colr_len + (cdef ? 10 + 6 * p->ncomp : 0);
    put(jp2h_len, 4); put(BOX_JP2H, 4);
    put(22, 4); put(BOX_IHDR, 4); put(p->height, 4); put(p->width, 4); put(p->ncomp, 2);
    put(bpc, 1); put(7, 1); put(0, 1); put(0, 1);                            // C = 7, UnkC = 0, IPR = 0
    if (bpc == 255) {
        put(8 + p->ncomp, 4); put(BOX_BPCC, 4);
        for (uint32_t i = 0; i < p->ncomp; ++i) put((p->prec[i] - 1) + (p->sgnd[i] << 7), 1);
    }
    put(colr_len, 4); put(BOX_COLR, 4); put(icc ? 2 : 1, 1); put(0, 1); put(0, 1);   // METH, PRECEDENCE 0, APPROX 0
    if (icc) v.insert(v.end(), p->icc, p->icc + p->icc_len);
    else put(enumcs, 4);
    if (cdef) {
        put(10 + 6 * p->ncomp, 4); put(BOX_CDEF, 4); put(p->ncomp, 2);
        for (uint32_t i = 0; i < p->ncomp; ++i) {
            put(i, 2);
            if (i < color_channels) { put(0, 2); put(i + 1, 2); }
            else if ((int32_t)i == p->alpha_comp) { put(1, 2); put(0, 2); }          // opacity of the whole image
            else { put(65535, 2); put(65535, 2); }                                   // unspecified
        }
    }
    // contiguous codestream box with its real length (jp2.cpp:1976-2011)
    put((uint32_t)(len + 8), 4); put(BOX_JP2C, 4);
    uint8_t* buf = (uint8_t*)out_alloc(v.size() + len);
    if (!buf) return set_error(J2KCUDA_ERR_NOMEM, "out of memory");
    memcpy(buf, v.data(), v.size());
    memcpy(buf + v.size(), cs, len);
    *out = buf; *out_len = v.size() + len;
    return J2KCUDA_OK;
}); }

}  // extern "C"
