// TEST INFRASTRUCTURE: compiles the Tier-1 decoder core (j2k_b200/csrc/t1_dec_core.cuh) for the host and
// runs the very block routine the CUDA kernel runs, with a single "lane".  Lets the CPU suite check
// the decoder logic against the oracle without a GPU.  Never linked into libj2kcuda.
#include "../../j2k_b200/csrc/t1_dec_core.cuh"
#include <cstring>
#include <vector>

using namespace j2k;
using namespace j2k::t1dec;

extern "C" int t1dec_host_block(const uint8_t* data, int len, int w, int h, int orient, int numbps, int npasses,
                                const uint8_t* zc9, const uint8_t* sc, const uint32_t* mq, int32_t* out) {
    std::vector<uint8_t> seg(data, data + len);
    for (int i = 0; i < 40; ++i) seg.push_back(0xff);
    while (reinterpret_cast<uintptr_t>(seg.data()) & 7) return -1;
    T1DecBlock b;
    memset(&b, 0, sizeof(b));
    b.coeff = out; b.pitch = w; b.w = (uint16_t)w; b.h = (uint16_t)h; b.orient = (uint8_t)orient; b.qmfbid = 1;
    b.numbps = (uint8_t)numbps; b.npasses = (uint8_t)npasses; b.data_off = 0; b.data_len = (uint32_t)len; b.stepsize = 1.f;
    memset(out, 0, sizeof(int32_t) * (size_t)w * h);
    if (!npasses || !numbps || !w || !h) return 0;
    static DRows R;
    memset(&R, 0, sizeof(R));
    uint32_t tab[94];
    for (int i = 0; i < 94; ++i) tab[i] = mq_state_word(mq, i);
    uint32_t st[32];
    for (int i = 0; i < 32; ++i) st[i] = tab[0];
    st[CTX_UNI] = tab[46 * 2]; st[CTX_AGG] = tab[3 * 2]; st[CTX_ZC] = tab[4 * 2];
    static ColBuf cb;
    uint8_t sc2[256];
    for (int i = 0; i < 256; ++i) sc2[i] = sc_entry(sc, i);
    static MqIo io;
    decode_block<false>(b, R, cb, &io, st, zc9, sc2, tab, seg.data(), nullptr, 1, 0, 1);
    return 0;
}

// ---- the sample conversion the MCT kernels run (j2k_b200/csrc/copybuffer.cuh), compiled for the host ----
#include "../../j2k_b200/csrc/copybuffer.cuh"
extern "C" void copyconv_host(const int64_t* src, int64_t* dst, int n, int src_type, int src_depth, int src_sgnd, int dst_type,
                              int dst_depth, int dst_sgnd) {
    const j2k::CopyConv c = j2k::make_copy_conv(src_type, src_depth, src_sgnd, dst_type, dst_depth, dst_sgnd);
    for (int i = 0; i < n; ++i) dst[i] = c.plain ? j2k::cb_as(dst_type, src[i]) : j2k::copy_convert(src[i], c);
}

// ---- the generic variant (code-block styles, several codeword segments, ROI shift): decode_block<true> ----
extern "C" int t1dec_host_block_gen(const uint8_t* data, const int* seg_len, const int* seg_passes, int nsegs, int w, int h, int orient,
                                    int numbps, int cblksty, int roishift, const uint8_t* zc9, const uint8_t* sc, const uint32_t* mq, int32_t* out) {
    std::vector<uint64_t> arena64;
    std::vector<T1DecSeg> segs;
    size_t total = 0;
    for (int i = 0; i < nsegs; ++i) total += ((size_t)seg_len[i] + 2 + 7) / 8 + 1;
    arena64.assign(total + 8, ~0ull);
    uint8_t* arena = reinterpret_cast<uint8_t*>(arena64.data());
    size_t pos = 0, in = 0;
    int npasses = 0;
    for (int i = 0; i < nsegs; ++i) {
        memcpy(arena + pos, data + in, (size_t)seg_len[i]);
        segs.push_back(T1DecSeg{(uint32_t)pos, (uint32_t)seg_len[i], (uint32_t)seg_passes[i], 0});
        in += (size_t)seg_len[i]; npasses += seg_passes[i];
        pos += (((size_t)seg_len[i] + 2) + 7) & ~(size_t)7;
    }
    T1DecBlock b;
    memset(&b, 0, sizeof(b));
    b.coeff = out; b.pitch = w; b.w = (uint16_t)w; b.h = (uint16_t)h; b.orient = (uint8_t)orient; b.qmfbid = 1;
    b.numbps = (uint8_t)numbps; b.npasses = (uint8_t)npasses; b.stepsize = 1.f;
    b.data_off = segs.empty() ? 0 : segs[0].off; b.data_len = segs.empty() ? 0 : segs[0].len;
    b.nsegs = (uint16_t)nsegs; b.seg_index = 0; b.cblksty = (uint8_t)cblksty; b.roishift = (uint8_t)roishift;
    memset(out, 0, sizeof(int32_t) * (size_t)w * h);
    if (!npasses || !numbps || !w || !h) return 0;
    static DRows R;
    memset(&R, 0, sizeof(R));
    uint32_t tab[96];
    for (int i = 0; i < 94; ++i) tab[i] = mq_state_word(mq, i);
    uint32_t st[32];
    for (int i = 0; i < 32; ++i) st[i] = tab[0];
    st[CTX_UNI] = tab[46 * 2]; st[CTX_AGG] = tab[3 * 2]; st[CTX_ZC] = tab[4 * 2];
    static ColBuf cb;
    static MqIo io;
    uint8_t sc2[256];
    for (int i = 0; i < 256; ++i) sc2[i] = sc_entry(sc, i);
    decode_block<true>(b, R, cb, &io, st, zc9, sc2, tab, arena, segs.data(), 1, 0, 1);
    return 0;
}

// ---- the generic MQ encoder (RESET / TERMALL) the CUDA kernel k_t1_mq_gen runs, compiled for the host ----
#include "../../j2k_b200/csrc/t1_mq_gen.cuh"
extern "C" int mqgen_host(const uint8_t* sym, const uint32_t* pass_nsym, int npasses, int cblksty, const uint32_t* mq, uint8_t* out, int cap,
                          int32_t* rate, int32_t* total) {
    return j2k::mqgen::encode_block(sym, pass_nsym, npasses, cblksty, mq, out, cap, rate, total);
}
