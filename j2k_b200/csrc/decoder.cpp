// Persistent decoder context: host Tier-2 parse -> segment arena H2D -> Tier-1 decode (+ dequantise)
// -> inverse DWT levels -> inverse MCT + level shift + clamp + pack -> D2H.
// Mirrors opj_tcd_decode_tile (ext/grok/src/lib/openjp2/tcd.cpp:1106-1137).  No CPU pixel path.
#include "../../include/j2kcuda.h"
#include "common.h"
#include <thread>
#include "device.h"
#include "t2.h"
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

using namespace j2k;

// wait for everything queued on the stream with the calling thread asleep (cudaStreamSynchronize spins; with 16 worker
// threads per GPU and 8 GPUs that is 128 spinning threads on the box's cores)
static inline cudaError_t stream_wait(cudaStream_t st, cudaEvent_t ev) {
    cudaError_t rc = cudaEventRecord(ev, st);
    return rc != cudaSuccess ? rc : cudaEventSynchronize(ev);
}


#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return set_error(J2KCUDA_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

struct j2kcuda_decoder {
    int device = 0;
    cudaStream_t st = nullptr;
    DecoderHost host;
    unsigned reduce = 0;
    int W = 0, H = 0, pitch = 0, ncomp = 0, numres = 0, irreversible = 0;   // full-resolution geometry
    int outW = 0, outH = 0, S = 2;                                            // decoded (reduced) size, sample bytes
    size_t plane_cap = 0;
    int32_t* planeA[4] = {nullptr, nullptr, nullptr, nullptr};
    int32_t* planeB[4] = {nullptr, nullptr, nullptr, nullptr};
    // per-frame tables
    std::vector<T1DecBlock> h_blocks;
    std::vector<T1DecSeg> h_segs;                                             // codeword segments of blocks that have several
    T1DecSeg* d_segs = nullptr; size_t d_segs_cap = 0;
    bool generic = false;                                                     // some block needs the generic Tier-1 variant
    bool subsampled = false;                                                  // some component has dx or dy != 1
    // decode window in reference-grid coordinates at full resolution (opj_set_decode_area); x1 == 0: the whole image
    uint32_t win_x0 = 0, win_y0 = 0, win_x1 = 0, win_y1 = 0;
    uint64_t blocks_skipped = 0;
    T1DecBlock* d_blocks = nullptr; size_t d_blocks_cap = 0;
    uint8_t* h_seg = nullptr; size_t h_seg_cap = 0;                           // pinned
    uint8_t* d_seg = nullptr; size_t d_seg_cap = 0;
    size_t seg_bytes = 0;
    std::vector<std::vector<DwtJob>> level_jobs;
    std::vector<int> level_items, level_seg, level_off;
    DwtJob* d_jobs = nullptr; size_t d_jobs_cap = 0;
    uint8_t* d_out = nullptr; size_t d_out_cap = 0;
    uint8_t* h_out = nullptr; size_t h_out_cap = 0;                           // pinned
    uint8_t* d_stage = nullptr; size_t d_stage_cap = 0;                       // per-channel planes in the caller's sample type
    cudaEvent_t ev[8];
    cudaEvent_t ev_wait = nullptr;   // blocking-sync event: the host worker sleeps instead of spinning
    float ms[8] = {0};
    uint64_t counters[6] = {0};
    bool prepared = false, ran = false;
};

static void free_decoder(j2kcuda_decoder* d) {
    if (!d) return;
    cudaSetDevice(d->device);
    for (int c = 0; c < 4; ++c) { cudaFree(d->planeA[c]); cudaFree(d->planeB[c]); }
    cudaFree(d->d_blocks); cudaFree(d->d_segs); cudaFree(d->d_seg); cudaFree(d->d_jobs); cudaFree(d->d_out); cudaFree(d->d_stage);
    cudaFreeHost(d->h_seg); cudaFreeHost(d->h_out);
    for (auto& e : d->ev) if (e) cudaEventDestroy(e);
    if (d->ev_wait) cudaEventDestroy(d->ev_wait);
    if (d->st) cudaStreamDestroy(d->st);
    delete d;
}

template <class T> static int grow_dev(T*& p, size_t& cap, size_t need) {
    if (need <= cap) return 0;
    cudaFree(p); p = nullptr; cap = 0;
    size_t n = need + need / 4 + 256;
    if (cudaMalloc(&p, n) != cudaSuccess) return set_error(J2KCUDA_ERR_NOMEM, "cudaMalloc failed");
    cap = n;
    return 0;
}
static int grow_host(uint8_t*& p, size_t& cap, size_t need) {
    if (need <= cap) return 0;
    cudaFreeHost(p); p = nullptr; cap = 0;
    size_t n = need + need / 4 + 256;
    if (cudaMallocHost(&p, n) != cudaSuccess) return set_error(J2KCUDA_ERR_NOMEM, "cudaMallocHost failed");
    cap = n;
    return 0;
}

static void fill_info(const CodingParams& cp, j2kcuda_info* out) {
    memset(out, 0, sizeof(*out));
    out->width = cp.x1 - cp.x0; out->height = cp.y1 - cp.y0; out->ncomp = (uint32_t)cp.ncomp;
    for (int c = 0; c < cp.ncomp; ++c) {
        out->prec[c] = (uint32_t)cp.comps[c].prec; out->sgnd[c] = (uint32_t)cp.comps[c].sgnd;
        out->dx[c] = (uint32_t)cp.comps[c].dx; out->dy[c] = (uint32_t)cp.comps[c].dy;
    }
    out->reversible = cp.comps[0].qmfbid == 1; out->numresolution = (uint32_t)cp.comps[0].numres;
    out->numlayers = (uint32_t)cp.numlayers; out->mct = (uint32_t)cp.mct; out->prog_order = (uint32_t)cp.prog;
    out->tile_w = cp.tdx; out->tile_h = cp.tdy; out->tiles_x = cp.tw; out->tiles_y = cp.th;
    out->cblk_w = 1u << cp.comps[0].cblkw; out->cblk_h = 1u << cp.comps[0].cblkh; out->profile = (uint32_t)cp.rsiz;
}

static int prepare(j2kcuda_decoder* d, const uint8_t* bytes, size_t n, unsigned reduce) {
    auto t0 = std::chrono::steady_clock::now();
    size_t off = 0, len = 0; bool jp2 = false;
    if (!find_codestream(bytes, n, off, len, jp2)) return set_error(J2KCUDA_ERR_STREAM, "not a JPEG 2000 codestream or JP2 file");
    const uint8_t* cs = bytes + off;
    std::string err;
    if (!d->host.parse(cs, len, err)) return set_error(J2KCUDA_ERR_STREAM, err);
    const CodingParams& cp = d->host.cp;
    bool subsampled = false;
    for (int c = 0; c < cp.ncomp; ++c) {
        if (cp.comps[c].dx != 1 || cp.comps[c].dy != 1) subsampled = true;
        if (cp.comps[c].numres != cp.comps[0].numres || cp.comps[c].qmfbid != cp.comps[0].qmfbid)
            return set_error(J2KCUDA_ERR_UNSUPPORTED, "components with different decompositions are not supported");
        if (cp.comps[c].cblkw > 6 || cp.comps[c].cblkh > 6) return set_error(J2KCUDA_ERR_UNSUPPORTED, "code-blocks larger than 64x64 samples");
        if (cp.comps[c].prec > 16) return set_error(J2KCUDA_ERR_UNSUPPORTED, "more than 16 bits per sample");
    }
    if (subsampled) {
        // components at a lower resolution than the reference grid (chroma sub-sampling): decoded at their own size, replicated
        // into the output as Codec::CopyBuffer does it above the reference codec (src/common/j2k_codec.cpp:262-279)
        if (cp.comps[0].dx != 1 || cp.comps[0].dy != 1) return set_error(J2KCUDA_ERR_UNSUPPORTED, "sub-sampled first component");
        if (cp.mct) return set_error(J2KCUDA_ERR_UNSUPPORTED, "multi-component transform over sub-sampled components");
        if (reduce > 0 && cp.tw * cp.th > 1) return set_error(J2KCUDA_ERR_UNSUPPORTED, "reduced decode of a tiled image with sub-sampled components");
        if (d->host.tile_count) return set_error(J2KCUDA_ERR_UNSUPPORTED, "tile-range decode with sub-sampled components");
    }
    d->subsampled = subsampled;
    d->numres = cp.comps[0].numres;
    if ((int)reduce >= d->numres) return set_error(J2KCUDA_ERR_PARAM, "reduce exceeds the number of resolutions");   // j2k.cpp read_SPCod check
    d->reduce = reduce;
    d->ncomp = cp.ncomp;
    d->irreversible = cp.comps[0].qmfbid == 0;
    d->W = (int)(cp.x1 - cp.x0); d->H = (int)(cp.y1 - cp.y0);
    d->pitch = (d->W + 3) & ~3;
    d->outW = (int)(ceildivpow2(cp.x1, reduce) - ceildivpow2(cp.x0, reduce));
    d->outH = (int)(ceildivpow2(cp.y1, reduce) - ceildivpow2(cp.y0, reduce));
    int maxprec = 0;
    for (int c = 0; c < cp.ncomp; ++c) maxprec = std::max(maxprec, cp.comps[c].prec);
    d->S = maxprec <= 8 ? 1 : 2;
    CK(cudaSetDevice(d->device));
    const size_t plane_bytes = (size_t)d->pitch * d->H * 4;
    if (plane_bytes > d->plane_cap) {
        for (int c = 0; c < 4; ++c) { cudaFree(d->planeA[c]); cudaFree(d->planeB[c]); d->planeA[c] = d->planeB[c] = nullptr; }
        d->plane_cap = 0;
        for (int c = 0; c < 4; ++c) { CK(cudaMalloc(&d->planeA[c], plane_bytes)); CK(cudaMalloc(&d->planeB[c], plane_bytes)); }
        d->plane_cap = plane_bytes;
    }
    // ---- block table + segment arena ----
    const int levels = d->numres - 1;
    const int maxres = d->numres - 1 - (int)reduce;
    d->h_blocks.clear(); d->h_segs.clear(); d->blocks_skipped = 0;
    bool generic = false;
    size_t seg = 0;
    for (auto& t : d->host.tiles)
        for (size_t k = 0; k < t.geo.cblks.size(); ++k)                   // (tiles outside this context's range have no geometry)
            if (t.geo.cblks[k].res <= maxres) seg += ((((size_t)t.blocks[k].len + 2) + 7) & ~(size_t)7) + 16 * t.blocks[k].segs.size();
    if (int rc = grow_host(d->h_seg, d->h_seg_cap, seg + 64)) return rc;
    seg = 0;
    d->level_jobs.assign(levels, {});
    uint64_t nblk = 0;
    for (auto& t : d->host.tiles) {
        const TileGeo& g = t.geo;
        if (g.comps.empty()) continue;
        for (int c = 0; c < d->ncomp; ++c) {
            const CompGeo& cg = g.comps[c];
            const size_t origin = (size_t)(cg.r.y0 - (int)ceildiv(cp.y0, cp.comps[c].dy)) * d->pitch + (cg.r.x0 - (int)ceildiv(cp.x0, cp.comps[c].dx));
            for (int l = levels; l > (int)reduce; --l) {
                const ResGeo& outr = cg.res[d->numres - l];          // region this inverse level produces
                if (outr.r.w() <= 0 || outr.r.h() <= 0) continue;
                DwtJob j;
                int32_t* from = (l & 1) ? d->planeB[c] : d->planeA[c];
                int32_t* to = ((l - 1) & 1) ? d->planeB[c] : d->planeA[c];
                j.src = from + origin; j.dst = to + origin;
                j.x0 = outr.r.x0; j.y0 = outr.r.y0; j.x1 = outr.r.x1; j.y1 = outr.r.y1;
                j.pitch = d->pitch; j.strip0 = j.nstrips = j.nsegs = 0;
                d->level_jobs[l - 1].push_back(j);
            }
        }
        for (size_t k = 0; k < g.cblks.size(); ++k) {
            const CblkGeo& cb = g.cblks[k];
            if (cb.res > maxres) continue;
            DecBlock& db = t.blocks[k];
            const CompGeo& cg = g.comps[cb.comp];
            if (d->win_x1 && db.included) {
                // Decode window (dwt_region.cpp, tile_buf.cpp:233-340): a code-block whose samples cannot reach the window
                // through the synthesis filters is left at zero.  A sample of decomposition level n sits on a 2^n grid; the
                // inverse transform spreads it over fewer than 3 * 2^n (5/3) / 6 * 2^n (9/7) reference-grid samples per side.
                const int n = cb.res == 0 ? levels : d->numres - cb.res;
                const int64_t m = (int64_t)(cp.comps[cb.comp].qmfbid ? 3 : 6) << n;
                const int64_t dx = cp.comps[cb.comp].dx, dy = cp.comps[cb.comp].dy;
                const int64_t bx0 = (((int64_t)cb.r.x0 << n) - m) * dx, bx1 = (((int64_t)cb.r.x1 << n) + m) * dx;
                const int64_t by0 = (((int64_t)cb.r.y0 << n) - m) * dy, by1 = (((int64_t)cb.r.y1 << n) + m) * dy;
                if (bx1 <= (int64_t)d->win_x0 || bx0 >= (int64_t)d->win_x1 || by1 <= (int64_t)d->win_y0 || by0 >= (int64_t)d->win_y1) {
                    db.included = false; db.npasses = 0; db.len = 0; db.chunks.clear(); db.segs.clear();
                    ++d->blocks_skipped;
                }
            }
            const BandGeo& band = cg.res[cb.res].bands[cb.band];
            T1DecBlock b;
            memset(&b, 0, sizeof(b));
            const int level = cb.res == 0 ? levels : d->numres - cb.res;
            int32_t* plane = (level & 1) ? d->planeB[cb.comp] : d->planeA[cb.comp];
            b.coeff = plane + (size_t)(cg.r.y0 - (int)ceildiv(cp.y0, cp.comps[cb.comp].dy) + cb.py) * d->pitch + (cg.r.x0 - (int)ceildiv(cp.x0, cp.comps[cb.comp].dx) + cb.px);
            b.pitch = d->pitch;
            b.w = (uint16_t)cb.r.w(); b.h = (uint16_t)cb.r.h();
            b.orient = (uint8_t)cb.orient; b.qmfbid = (uint8_t)cp.comps[cb.comp].qmfbid;
            int nbps = db.included ? db.numbps : 0;
            const int roishift = cp.comps[cb.comp].roishift, sty = cp.comps[cb.comp].cblksty;
            if (nbps + roishift > 30) return set_error(J2KCUDA_ERR_STREAM, "code-block with more than 30 bit-planes");
            b.numbps = (uint8_t)std::max(nbps, 0);
            const int maxp = 3 * (nbps + roishift) - 2;
            b.npasses = (uint8_t)std::min(db.npasses, maxp > 0 ? maxp : 0);
            b.stepsize = band.stepsize;
            b.cblksty = (uint8_t)sty; b.roishift = (uint8_t)roishift;
            b.nsegs = (uint16_t)std::max<size_t>(db.segs.size(), 1);
            if (sty || roishift || db.segs.size() > 1) generic = true;
            b.data_off = (uint32_t)seg; b.data_len = (uint32_t)db.len;
            if (db.segs.size() <= 1) {
                uint8_t* p = d->h_seg + seg;
                for (auto& ch : db.chunks) { memcpy(p, cs + ch.first, ch.second); p += ch.second; }
                p[0] = 0xff; p[1] = 0xff;
                seg += (((size_t)db.len + 2) + 7) & ~(size_t)7;
            } else {
                // several codeword segments: each one 8-byte aligned and followed by 0xFF 0xFF, as the MQ reader wants them
                static thread_local std::vector<uint8_t> flat;
                flat.clear();
                for (auto& ch : db.chunks) flat.insert(flat.end(), cs + ch.first, cs + ch.first + ch.second);
                b.seg_index = (uint32_t)d->h_segs.size();
                size_t in = 0;
                int passes_left = b.npasses;
                for (auto& sg : db.segs) {
                    const size_t sl = std::min<size_t>((size_t)sg.len, flat.size() - in);
                    const int np = std::min(sg.numpasses, passes_left);
                    memcpy(d->h_seg + seg, flat.data() + in, sl);
                    d->h_seg[seg + sl] = 0xff; d->h_seg[seg + sl + 1] = 0xff;
                    d->h_segs.push_back(T1DecSeg{(uint32_t)seg, (uint32_t)sl, (uint32_t)np, 0});
                    in += sl; passes_left -= np;
                    seg += ((sl + 2) + 7) & ~(size_t)7;
                }
            }
            d->h_blocks.push_back(b);
            ++nblk;
        }
    }
    memset(d->h_seg + seg, 0xff, 32);          // the word-wise MQ reader looks up to 24 bytes past a segment
    seg += 32;
    d->seg_bytes = seg;
    // longest code-blocks first: a block's decode time follows its byte count, and the warps that get the
    // long blocks of the low-frequency bands last would otherwise run on alone at the end of the launch
    std::stable_sort(d->h_blocks.begin(), d->h_blocks.end(),
                     [](const T1DecBlock& a, const T1DecBlock& b) { return a.data_len > b.data_len; });
    // ---- DWT job tables ----
    std::vector<DwtJob> all;
    d->level_off.assign(levels, 0); d->level_items.assign(levels, 0); d->level_seg.assign(levels, 0);
    for (int l = 0; l < levels; ++l) {
        auto& v = d->level_jobs[l];
        d->level_off[l] = (int)all.size();
        if (v.empty()) continue;
        d->level_items[l] = dwt_plan_jobs(v.data(), (int)v.size());
        d->level_seg[l] = v[0].strip0;
        all.insert(all.end(), v.begin(), v.end());
    }
    d->ms[0] = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    // ---- H2D ----
    CK(cudaEventRecord(d->ev[0], d->st));
    if (int rc = grow_dev(d->d_seg, d->d_seg_cap, seg + 16)) return rc;
    if (int rc = grow_dev(d->d_blocks, d->d_blocks_cap, d->h_blocks.size() * sizeof(T1DecBlock))) return rc;
    if (int rc = grow_dev(d->d_jobs, d->d_jobs_cap, all.size() * sizeof(DwtJob))) return rc;
    if (seg) CK(cudaMemcpyAsync(d->d_seg, d->h_seg, seg, cudaMemcpyHostToDevice, d->st));
    d->generic = generic;
    if (int rc = grow_dev(d->d_segs, d->d_segs_cap, (d->h_segs.size() + 1) * sizeof(T1DecSeg))) return rc;
    if (!d->h_segs.empty()) CK(cudaMemcpyAsync(d->d_segs, d->h_segs.data(), d->h_segs.size() * sizeof(T1DecSeg), cudaMemcpyHostToDevice, d->st));
    if (!d->h_blocks.empty()) CK(cudaMemcpyAsync(d->d_blocks, d->h_blocks.data(), d->h_blocks.size() * sizeof(T1DecBlock), cudaMemcpyHostToDevice, d->st));
    if (!all.empty()) CK(cudaMemcpyAsync(d->d_jobs, all.data(), all.size() * sizeof(DwtJob), cudaMemcpyHostToDevice, d->st));
    CK(cudaEventRecord(d->ev[1], d->st));
    CK(stream_wait(d->st, d->ev_wait));          // h_blocks / all are pageable host vectors
    d->counters[1] = nblk; d->counters[2] = d->blocks_skipped; d->counters[3] = (uint64_t)len;
    d->counters[4] = seg + d->h_blocks.size() * sizeof(T1DecBlock) + all.size() * sizeof(DwtJob);
    d->prepared = true; d->ran = false;
    return J2KCUDA_OK;
}

static int run(j2kcuda_decoder* d) {
    if (!d->prepared) return set_error(J2KCUDA_ERR_PARAM, "decoder_prepare has not completed");
    CK(cudaSetDevice(d->device));
    const CodingParams& cp = d->host.cp;
    cudaStream_t st = d->st;
    uint64_t launches = 0;
    CK(cudaEventRecord(d->ev[2], st));
    launch_t1_decode(d->d_blocks, (int)d->h_blocks.size(), d->d_seg, d->d_segs, d->generic ? 1 : 0, 0, st); ++launches;
    CK(cudaEventRecord(d->ev[3], st));
    const int levels = d->numres - 1;
    for (int l = levels; l > (int)d->reduce; --l) {
        if (d->level_jobs[l - 1].empty()) continue;
        if (l == (int)d->reduce + 1) CK(cudaEventRecord(d->ev[6], st));
        launch_dwt_inverse(d->d_jobs + d->level_off[l - 1], (int)d->level_jobs[l - 1].size(), d->level_items[l - 1], d->level_seg[l - 1],
                           d->irreversible ? DWT_97_FLOAT : DWT_53, st);
        ++launches;
    }
    if (levels <= (int)d->reduce) CK(cudaEventRecord(d->ev[6], st));
    CK(cudaEventRecord(d->ev[4], st));
    // ---- inverse MCT + level shift + clamp + pack, tile by tile, into a tight interleaved image ----
    const size_t out_bytes = (size_t)d->outW * d->outH * d->ncomp * d->S;
    if (int rc = grow_dev(d->d_out, d->d_out_cap, out_bytes)) return rc;
    const int ix0 = (int)ceildivpow2(cp.x0, d->reduce), iy0 = (int)ceildivpow2(cp.y0, d->reduce);
    if (d->subsampled) {
        SubOut so;
        memset(&so, 0, sizeof(so));
        so.width = d->outW; so.height = d->outH; so.pitch = d->pitch; so.ncomp = d->ncomp; so.sample_bytes = d->S;
        so.irreversible = d->irreversible; so.sgnd = cp.comps[0].sgnd;
        for (int c = 0; c < d->ncomp; ++c) {
            so.plane[c] = (d->reduce & 1) ? d->planeB[c] : d->planeA[c];
            so.dx[c] = cp.comps[c].dx; so.dy[c] = cp.comps[c].dy;
            so.cw[c] = (int)(ceildivpow2(ceildiv(cp.x1, cp.comps[c].dx), d->reduce) - ceildivpow2(ceildiv(cp.x0, cp.comps[c].dx), d->reduce));
            so.ch[c] = (int)(ceildivpow2(ceildiv(cp.y1, cp.comps[c].dy), d->reduce) - ceildivpow2(ceildiv(cp.y0, cp.comps[c].dy), d->reduce));
            so.prec[c] = cp.comps[c].prec; so.dc_shift[c] = cp.comps[c].sgnd ? 0 : 1 << (cp.comps[c].prec - 1);
        }
        launch_output_subsampled(so, d->d_out, st);
        ++launches;
    } else
    for (auto& t : d->host.tiles) {
        if (t.geo.comps.empty()) continue;
        const CompGeo& cg = t.geo.comps[0];
        const int rx0 = (int)ceildivpow2((uint32_t)cg.r.x0, d->reduce), ry0 = (int)ceildivpow2((uint32_t)cg.r.y0, d->reduce);
        const int rx1 = (int)ceildivpow2((uint32_t)cg.r.x1, d->reduce), ry1 = (int)ceildivpow2((uint32_t)cg.r.y1, d->reduce);
        if (rx1 <= rx0 || ry1 <= ry0) continue;
        MctParams mp;
        memset(&mp, 0, sizeof(mp));
        mp.width = rx1 - rx0; mp.height = ry1 - ry0; mp.pitch = d->pitch; mp.ncomp = d->ncomp;
        mp.mct = cp.mct ? 1 : 0; mp.irreversible = d->irreversible;
        PixelDst dst;
        memset(&dst, 0, sizeof(dst));
        dst.sample_bytes = d->S; dst.sgnd = cp.comps[0].sgnd;
        const int32_t* planes[4] = {nullptr, nullptr, nullptr, nullptr};
        const size_t origin = (size_t)(cg.r.y0 - (int)cp.y0) * d->pitch + (cg.r.x0 - (int)cp.x0);
        for (int c = 0; c < d->ncomp; ++c) {
            mp.prec[c] = cp.comps[c].prec;
            mp.dc_shift[c] = cp.comps[c].sgnd ? 0 : 1 << (cp.comps[c].prec - 1);
            planes[c] = ((d->reduce & 1) ? d->planeB[c] : d->planeA[c]) + origin;
            dst.base[c] = d->d_out + ((size_t)(ry0 - iy0) * d->outW + (rx0 - ix0)) * d->ncomp * d->S + (size_t)c * d->S;
            dst.colbytes[c] = (int64_t)d->ncomp * d->S;
            dst.rowbytes[c] = (int64_t)d->outW * d->ncomp * d->S;
        }
        launch_mct_inverse(dst, mp, planes, st);
        ++launches;
    }
    CK(cudaEventRecord(d->ev[5], st));
    CK(stream_wait(st, d->ev_wait));
    CK(cudaGetLastError());
    cudaEventElapsedTime(&d->ms[1], d->ev[0], d->ev[1]);
    cudaEventElapsedTime(&d->ms[2], d->ev[2], d->ev[3]);
    cudaEventElapsedTime(&d->ms[3], d->ev[3], d->ev[4]);
    cudaEventElapsedTime(&d->ms[4], d->ev[4], d->ev[5]);
    cudaEventElapsedTime(&d->ms[6], d->ev[6], d->ev[4]);
    d->counters[0] = launches;
    d->ran = true;
    return J2KCUDA_OK;
}

// device image -> caller planes (CopyBuffer semantics, src/common/j2k_codec.cpp:222-427)
static int download(j2kcuda_decoder* d, const j2kcuda_plane* planes, int nplanes) {
    if (!d->ran) return set_error(J2KCUDA_ERR_PARAM, "decoder_run has not completed");
    if (nplanes < 1) return set_error(J2KCUDA_ERR_PARAM, "no planes");
    CK(cudaSetDevice(d->device));
    const CodingParams& cp = d->host.cp;
    const int nc = std::min(nplanes, d->ncomp);
    const size_t out_bytes = (size_t)d->outW * d->outH * d->ncomp * d->S;
    for (int c = 0; c < nc; ++c)
        if (!planes[c].buf || (int)planes[c].width > d->outW || (int)planes[c].height > d->outH)
            return set_error(J2KCUDA_ERR_PARAM, "destination plane missing or larger than the decoded image");
    const int S = d->S;
    auto sb = [](int t) { return t == J2KCUDA_UCHAR ? 1 : (t == J2KCUDA_USHORT ? 2 : 4); };
    bool tight = nplanes >= d->ncomp;
    for (int c = 0; c < nc && tight; ++c) {
        const j2kcuda_plane& p = planes[c];
        const int depth = p.depth ? p.depth : cp.comps[c].prec;
        tight = sb(p.sample_type) == S && (uint8_t*)p.buf == (uint8_t*)planes[0].buf + (size_t)c * S &&
                p.colbytes == (int64_t)d->ncomp * S && p.rowbytes == (int64_t)d->outW * d->ncomp * S &&
                (int)p.width == d->outW && (int)p.height == d->outH && depth == cp.comps[c].prec && !p.sgnd == !cp.comps[c].sgnd;
    }
    // decode window: the planes receive the window's samples (its top-left sample first), cut out of the full image on the device
    size_t win_off = 0;
    int win_w = d->outW, win_h = d->outH;
    if (d->win_x1) {
        const int ix0 = (int)ceildivpow2(cp.x0, d->reduce), iy0 = (int)ceildivpow2(cp.y0, d->reduce);
        const int wx0 = std::max((int)ceildivpow2(d->win_x0, d->reduce) - ix0, 0), wy0 = std::max((int)ceildivpow2(d->win_y0, d->reduce) - iy0, 0);
        const int wx1 = std::min((int)ceildivpow2(d->win_x1, d->reduce) - ix0, d->outW), wy1 = std::min((int)ceildivpow2(d->win_y1, d->reduce) - iy0, d->outH);
        if (wx1 <= wx0 || wy1 <= wy0) return set_error(J2KCUDA_ERR_PARAM, "decode window lies outside the image");
        win_off = ((size_t)wy0 * d->outW + wx0) * d->ncomp * S;
        win_w = wx1 - wx0; win_h = wy1 - wy0;
        for (int c = 0; c < nc; ++c)
            if ((int)planes[c].width > win_w || (int)planes[c].height > win_h) return set_error(J2KCUDA_ERR_PARAM, "destination plane larger than the decode window");
        tight = false;
    }
    CK(cudaEventRecord(d->ev[7], d->st));
    if (d->host.tile_count) {
        // this context decoded a range of the frame's tiles (the others are on other GPUs): their rectangles go straight
        // into the caller's image
        if (!tight) return set_error(J2KCUDA_ERR_UNSUPPORTED, "tile-range decode needs a destination in the codestream's own layout");
        const size_t row = (size_t)d->outW * d->ncomp * S;
        const int ix0 = (int)ceildivpow2(cp.x0, d->reduce), iy0 = (int)ceildivpow2(cp.y0, d->reduce);
        for (auto& t : d->host.tiles) {
            if (t.geo.comps.empty()) continue;
            const Rect& r = t.geo.comps[0].r;
            const int rx0 = (int)ceildivpow2((uint32_t)r.x0, d->reduce) - ix0, ry0 = (int)ceildivpow2((uint32_t)r.y0, d->reduce) - iy0;
            const int rx1 = (int)ceildivpow2((uint32_t)r.x1, d->reduce) - ix0, ry1 = (int)ceildivpow2((uint32_t)r.y1, d->reduce) - iy0;
            if (rx1 <= rx0 || ry1 <= ry0) continue;
            const size_t off = (size_t)ry0 * row + (size_t)rx0 * d->ncomp * S;
            CK(cudaMemcpy2DAsync((uint8_t*)planes[0].buf + off, row, d->d_out + off, row, (size_t)(rx1 - rx0) * d->ncomp * S, (size_t)(ry1 - ry0),
                                 cudaMemcpyDeviceToHost, d->st));
        }
        CK(stream_wait(d->st, d->ev_wait));
    } else if (tight) {
        cudaPointerAttributes attr;
        bool pinned = cudaPointerGetAttributes(&attr, planes[0].buf) == cudaSuccess && attr.type == cudaMemoryTypeHost;
        cudaGetLastError();
        if (pinned) {
            CK(cudaMemcpyAsync(planes[0].buf, d->d_out, out_bytes, cudaMemcpyDeviceToHost, d->st));
            CK(stream_wait(d->st, d->ev_wait));
        } else {
            if (int rc = grow_host(d->h_out, d->h_out_cap, out_bytes)) return rc;
            // pageable destination: D2H into the page-locked buffer in quarters, each quarter copied out by its own
            // thread as soon as its DMA has finished
            const int parts = out_bytes >= (32u << 20) ? 4 : 1;
            cudaEvent_t evp[4] = {nullptr, nullptr, nullptr, nullptr};
            size_t cut[5];
            for (int i = 0; i <= parts; ++i) cut[i] = (out_bytes * (size_t)i / (size_t)parts) & ~(size_t)255;
            cut[parts] = out_bytes;
            bool ok = true;
            for (int i = 0; i < parts; ++i) {
                ok = ok && cudaMemcpyAsync(d->h_out + cut[i], d->d_out + cut[i], cut[i + 1] - cut[i], cudaMemcpyDeviceToHost, d->st) == cudaSuccess;
                ok = ok && cudaEventCreateWithFlags(&evp[i], cudaEventBlockingSync | cudaEventDisableTiming) == cudaSuccess;
                ok = ok && cudaEventRecord(evp[i], d->st) == cudaSuccess;
            }
            std::vector<std::thread> th;
            auto work = [&](int i) {
                if (evp[i] && cudaEventSynchronize(evp[i]) == cudaSuccess) memcpy((uint8_t*)planes[0].buf + cut[i], d->h_out + cut[i], cut[i + 1] - cut[i]);
            };
            for (int i = 1; i < parts && ok; ++i) th.emplace_back(work, i);
            if (ok) work(0);
            for (auto& x : th) x.join();
            for (int i = 0; i < parts; ++i) if (evp[i]) cudaEventDestroy(evp[i]);
            if (!ok) { CK(cudaGetLastError()); return set_error(J2KCUDA_ERR_CUDA, "download failed"); }
            CK(stream_wait(d->st, d->ev_wait));
        }
    } else {
        // Any other destination (other depth, sample type, sign, strides, or fewer channels than components): the
        // CopyBuffer arithmetic (src/common/j2k_codec.cpp:222-427) runs on the device, one tight plane per channel in the
        // caller's sample type; the host only moves bytes to the caller's strides.
        size_t stage = 0;
        size_t offs[4] = {0, 0, 0, 0};
        for (int c = 0; c < nc; ++c) {
            const j2kcuda_plane& p = planes[c];
            if (p.sample_type > J2KCUDA_INT) return set_error(J2KCUDA_ERR_PARAM, "unknown sample type");
            const int depth = p.depth ? p.depth : cp.comps[c].prec;
            if (depth > 8 * sb(p.sample_type)) return set_error(J2KCUDA_ERR_PARAM, "plane depth does not fit its sample type");
            offs[c] = stage;
            stage += ((size_t)p.width * p.height * sb(p.sample_type) + 255) & ~(size_t)255;
        }
        if (int rc = grow_dev(d->d_stage, d->d_stage_cap, stage)) return rc;
        if (int rc = grow_host(d->h_out, d->h_out_cap, stage)) return rc;
        for (int c = 0; c < nc; ++c) {
            const j2kcuda_plane& p = planes[c];
            const int depth = p.depth ? p.depth : cp.comps[c].prec;
            const CopyConv cc = make_copy_conv(J2KCUDA_INT, cp.comps[c].prec, cp.comps[c].sgnd, p.sample_type, depth, p.sgnd);
            // a plane of exactly the component's own (sub-sampled) size takes the component's samples, any other the replicated image
            int sx = 1, sy = 1;
            if (cp.comps[c].dx > 1 || cp.comps[c].dy > 1) {
                const int cw = (int)(ceildivpow2(ceildiv(cp.x1, cp.comps[c].dx), d->reduce) - ceildivpow2(ceildiv(cp.x0, cp.comps[c].dx), d->reduce));
                const int chh = (int)(ceildivpow2(ceildiv(cp.y1, cp.comps[c].dy), d->reduce) - ceildivpow2(ceildiv(cp.y0, cp.comps[c].dy), d->reduce));
                if ((int)p.width == cw && (int)p.height == chh) { sx = cp.comps[c].dx; sy = cp.comps[c].dy; }
            }
            launch_copy_channel(d->d_out + win_off, S, d->ncomp, c, cp.comps[0].sgnd, d->outW, (int)p.width, (int)p.height, sx, sy, d->d_stage + offs[c], cc, d->st);
        }
        CK(cudaMemcpyAsync(d->h_out, d->d_stage, stage, cudaMemcpyDeviceToHost, d->st));
        CK(stream_wait(d->st, d->ev_wait));
        CK(cudaGetLastError());
        for (int c = 0; c < nc; ++c) {
            const j2kcuda_plane& p = planes[c];
            const size_t ds = (size_t)sb(p.sample_type);
            for (uint32_t y = 0; y < p.height; ++y) {
                const uint8_t* s = d->h_out + offs[c] + (size_t)y * p.width * ds;
                uint8_t* q = (uint8_t*)p.buf + (size_t)y * p.rowbytes;
                if (p.colbytes == (int64_t)ds) memcpy(q, s, (size_t)p.width * ds);
                else if (ds == 1) for (uint32_t x = 0; x < p.width; ++x, q += p.colbytes) *q = s[x];
                else if (ds == 2) for (uint32_t x = 0; x < p.width; ++x, q += p.colbytes) memcpy(q, s + 2 * (size_t)x, 2);
                else for (uint32_t x = 0; x < p.width; ++x, q += p.colbytes) memcpy(q, s + 4 * (size_t)x, 4);
            }
        }
    }
    float t = 0;
    cudaEventRecord(d->ev[0], d->st); cudaEventSynchronize(d->ev[0]);
    cudaEventElapsedTime(&t, d->ev[7], d->ev[0]);
    d->ms[5] = t;
    d->counters[5] = out_bytes;
    return J2KCUDA_OK;
}

// ---- decoder contexts kept between one-shot calls (ReadFile after ReadFile); a decoder grows to whatever it is given ----
namespace {
std::mutex g_pool_mu;
std::vector<j2kcuda_decoder*> g_pool;
constexpr size_t POOL_MAX = 2;
j2kcuda_decoder* pool_take(int device) {
    if (!context_cache_enabled()) return nullptr;
    std::lock_guard<std::mutex> g(g_pool_mu);
    for (size_t i = 0; i < g_pool.size(); ++i)
        if (g_pool[i]->device == device) { j2kcuda_decoder* d = g_pool[i]; g_pool.erase(g_pool.begin() + (long)i); return d; }
    return nullptr;
}
void pool_give(j2kcuda_decoder* d) {
    j2kcuda_decoder* drop = nullptr;
    if (!context_cache_enabled()) drop = d;
    else {
        std::lock_guard<std::mutex> g(g_pool_mu);
        g_pool.push_back(d);
        if (g_pool.size() > POOL_MAX) { drop = g_pool.front(); g_pool.erase(g_pool.begin()); }
    }
    if (drop) free_decoder(drop);
}
}  // namespace
namespace j2k {
void decoder_pool_clear() {
    std::vector<j2kcuda_decoder*> all;
    { std::lock_guard<std::mutex> g(g_pool_mu); all.swap(g_pool); }
    for (auto* d : all) free_decoder(d);
}
}  // namespace j2k

namespace j2k {
void decoder_set_tile_range(::j2kcuda_decoder* d, uint32_t first, uint32_t count) { if (d) { d->host.tile_first = first; d->host.tile_count = count; } }
}  // namespace j2k

extern "C" {

j2kcuda_decoder* j2kcuda_decoder_create(int device) { try {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); set_error(J2KCUDA_ERR_NO_DEVICE, "no CUDA device: libj2kcuda has no CPU path"); return nullptr; }
    j2kcuda_decoder* d = new j2kcuda_decoder();
    for (auto& e : d->ev) e = nullptr;
    if (device >= 0) d->device = device; else cudaGetDevice(&d->device);
    bool ok = cudaSetDevice(d->device) == cudaSuccess && device_tables_init() == nullptr &&
              cudaStreamCreateWithFlags(&d->st, cudaStreamNonBlocking) == cudaSuccess;
    for (auto& e : d->ev) ok = ok && cudaEventCreate(&e) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&d->ev_wait, cudaEventBlockingSync | cudaEventDisableTiming) == cudaSuccess;
    if (!ok) { set_error(J2KCUDA_ERR_CUDA, "decoder_create: CUDA initialisation failed"); free_decoder(d); return nullptr; }
    return d;
} catch (...) { set_error(J2KCUDA_ERR_NOMEM, "j2kcuda_decoder_create: out of memory or internal error"); return nullptr; } }
void j2kcuda_decoder_destroy(j2kcuda_decoder* d) { free_decoder(d); }
int j2kcuda_decoder_prepare(j2kcuda_decoder* d, const uint8_t* bytes, size_t n, unsigned reduce) { return j2k::guarded([&]() -> int {
    if (!d || !bytes) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    d->prepared = false;
    return prepare(d, bytes, n, reduce);
}); }
// opj_set_decode_area (openjpeg.h:1287-1302): only the rectangle [x0, x1) x [y0, y1) of the reference grid (full-resolution
// coordinates) is wanted; code-blocks that cannot influence it are not decoded.  x1 == 0 && y1 == 0 resets to the whole image.
int j2kcuda_decoder_set_window(j2kcuda_decoder* d, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1) { return j2k::guarded([&]() -> int {
    if (!d) return set_error(J2KCUDA_ERR_PARAM, "null decoder");
    if ((x1 || y1) && (x1 <= x0 || y1 <= y0)) return set_error(J2KCUDA_ERR_PARAM, "empty decode window");
    d->win_x0 = x0; d->win_y0 = y0; d->win_x1 = x1; d->win_y1 = y1;
    return J2KCUDA_OK;
}); }

int j2kcuda_decoder_run(j2kcuda_decoder* d) { return j2k::guarded([&]() -> int { return d ? run(d) : set_error(J2KCUDA_ERR_PARAM, "null decoder"); }); }
int j2kcuda_decoder_download(j2kcuda_decoder* d, const j2kcuda_plane* planes, int nplanes) { return j2k::guarded([&]() -> int {
    if (!d || !planes) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    return download(d, planes, nplanes);
}); }
int j2kcuda_decoder_get_device_output(j2kcuda_decoder* d, void** d_pixels, int64_t* pitch_bytes, int* sample_type) { return j2k::guarded([&]() -> int {
    if (!d || !d->ran) return set_error(J2KCUDA_ERR_PARAM, "decoder_run has not completed");
    if (d_pixels) *d_pixels = d->d_out;
    if (pitch_bytes) *pitch_bytes = (int64_t)d->outW * d->ncomp * d->S;
    if (sample_type) *sample_type = d->S == 1 ? J2KCUDA_UCHAR : J2KCUDA_USHORT;
    return J2KCUDA_OK;
}); }
int j2kcuda_decoder_timings(j2kcuda_decoder* d, float* ms, int n) { return j2k::guarded([&]() -> int {
    if (!d || !ms) return 0;
    int k = n < 7 ? n : 7;
    for (int i = 0; i < k; ++i) ms[i] = d->ms[i];
    return k;
}); }
int j2kcuda_decoder_counters(j2kcuda_decoder* d, uint64_t* v, int n) { return j2k::guarded([&]() -> int {
    if (!d || !v) return 0;
    int k = n < 6 ? n : 6;
    for (int i = 0; i < k; ++i) v[i] = d->counters[i];
    return k;
}); }

int j2kcuda_probe(const uint8_t* bytes, size_t n, j2kcuda_info* out) { return j2k::guarded([&]() -> int {
    if (!bytes || !out) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    size_t off = 0, len = 0; bool jp2 = false;
    if (!find_codestream(bytes, n, off, len, jp2)) return set_error(J2KCUDA_ERR_STREAM, "not a JPEG 2000 codestream or JP2 file");
    DecoderHost h;
    std::string err;
    size_t pos = 0;
    if (!h.parse_main_header(bytes + off, len, pos, err)) return set_error(J2KCUDA_ERR_STREAM, err);
    fill_info(h.cp, out);
    out->is_jp2 = jp2; out->codestream_offset = off; out->codestream_length = len;
    return J2KCUDA_OK;
}); }

int j2kcuda_decode(const uint8_t* bytes, size_t n, unsigned reduce, const j2kcuda_plane* planes, int nplanes,
                   j2kcuda_abort_fn abort_fn, void* user) { return j2k::guarded([&]() -> int {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return set_error(J2KCUDA_ERR_NO_DEVICE, "no CUDA device: libj2kcuda has no CPU path"); }
    j2kcuda_decoder* d = pool_take(dev);
    if (!d) d = j2kcuda_decoder_create(-1);
    if (!d) return J2KCUDA_ERR_NO_DEVICE;
    int rc = j2kcuda_decoder_prepare(d, bytes, n, reduce);
    if (!rc && abort_fn && abort_fn(user)) rc = set_error(J2KCUDA_ERR_ABORTED, "aborted by the caller");
    if (!rc) rc = j2kcuda_decoder_run(d);
    if (!rc) rc = j2kcuda_decoder_download(d, planes, nplanes);
    if (rc == J2KCUDA_ERR_CUDA) j2kcuda_decoder_destroy(d); else pool_give(d);
    return rc;
}); }

int j2kcuda_decode_region(const uint8_t* bytes, size_t n, unsigned reduce, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1,
                          const j2kcuda_plane* planes, int nplanes) { return j2k::guarded([&]() -> int {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return set_error(J2KCUDA_ERR_NO_DEVICE, "no CUDA device: libj2kcuda has no CPU path"); }
    j2kcuda_decoder* d = pool_take(dev);
    if (!d) d = j2kcuda_decoder_create(-1);
    if (!d) return J2KCUDA_ERR_NO_DEVICE;
    int rc = j2kcuda_decoder_set_window(d, x0, y0, x1, y1);
    if (!rc) rc = j2kcuda_decoder_prepare(d, bytes, n, reduce);
    if (!rc) rc = j2kcuda_decoder_run(d);
    if (!rc) rc = j2kcuda_decoder_download(d, planes, nplanes);
    j2kcuda_decoder_set_window(d, 0, 0, 0, 0);
    if (rc == J2KCUDA_ERR_CUDA) j2kcuda_decoder_destroy(d); else pool_give(d);
    return rc;
}); }

int j2kcuda_decoder_copy_to_device(j2kcuda_decoder* d, void* d_pixels, int64_t pitch_bytes) { return j2k::guarded([&]() -> int {
    if (!d || !d_pixels) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    if (!d->ran) return set_error(J2KCUDA_ERR_PARAM, "decoder_run has not completed");
    const size_t row = (size_t)d->outW * d->ncomp * d->S;
    if (pitch_bytes < (int64_t)row) return set_error(J2KCUDA_ERR_PARAM, "destination pitch smaller than a decoded row");
    CK(cudaSetDevice(d->device));
    CK(cudaMemcpy2DAsync(d_pixels, (size_t)pitch_bytes, d->d_out, row, row, (size_t)d->outH, cudaMemcpyDeviceToDevice, d->st));
    CK(stream_wait(d->st, d->ev_wait));
    return J2KCUDA_OK;
}); }

int j2kcuda_t2_parse(const uint8_t* bytes, size_t n, unsigned reduce, int32_t** blkinfo, int* nblocks, uint8_t** seg, size_t* seg_len) { return j2k::guarded([&]() -> int {
    if (!bytes || !blkinfo || !nblocks || !seg || !seg_len) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    size_t off = 0, len = 0; bool jp2 = false;
    if (!find_codestream(bytes, n, off, len, jp2)) return set_error(J2KCUDA_ERR_STREAM, "not a JPEG 2000 codestream or JP2 file");
    DecoderHost h;
    std::string err;
    if (!h.parse(bytes + off, len, err)) return set_error(J2KCUDA_ERR_STREAM, err);
    (void)reduce;
    size_t total = 0; int nb = 0;
    for (auto& t : h.tiles) for (auto& b : t.blocks) { total += (size_t)b.len; ++nb; }
    int32_t* info = (int32_t*)out_alloc(sizeof(int32_t) * 4 * (size_t)(nb ? nb : 1));
    uint8_t* data = (uint8_t*)out_alloc(total);
    if (!info || !data) { out_free(info); out_free(data); return set_error(J2KCUDA_ERR_NOMEM, "malloc"); }
    size_t pos = 0; int k = 0;
    for (auto& t : h.tiles)
        for (auto& b : t.blocks) {
            info[4 * k] = b.included ? b.numbps : 0; info[4 * k + 1] = b.npasses; info[4 * k + 2] = b.len; info[4 * k + 3] = (int32_t)pos;
            for (auto& ch : b.chunks) { memcpy(data + pos, bytes + off + ch.first, ch.second); pos += ch.second; }
            ++k;
        }
    *blkinfo = info; *nblocks = nb; *seg = data; *seg_len = total;
    return J2KCUDA_OK;
}); }

int j2kcuda_stage_t1_decode(const uint8_t* data, const int32_t* blk, int nblk, int32_t* out) { return j2k::guarded([&]() -> int {
    if (!data || !blk || !out || nblk < 0) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return set_error(J2KCUDA_ERR_NO_DEVICE, "no CUDA device: libj2kcuda has no CPU path"); }
    if (const char* e = device_tables_init()) return set_error(J2KCUDA_ERR_CUDA, e);
    std::vector<T1DecBlock> hb(nblk);
    std::vector<uint8_t> arena;
    size_t in = 0, outn = 0;
    for (int i = 0; i < nblk; ++i) {
        const int32_t* g = blk + 6 * (size_t)i;
        if (g[0] < 0 || g[1] < 0 || g[0] > 64 || g[1] > 64 || g[5] < 0) return set_error(J2KCUDA_ERR_PARAM, "bad block");
        size_t o = arena.size();
        arena.insert(arena.end(), data + in, data + in + g[5]);
        arena.push_back(0xff); arena.push_back(0xff);
        while (arena.size() & 7) arena.push_back(0xff);
        T1DecBlock& b = hb[i];
        memset(&b, 0, sizeof(b));
        b.w = (uint16_t)g[0]; b.h = (uint16_t)g[1]; b.pitch = g[0];
        b.orient = (uint8_t)g[2]; b.qmfbid = 1; b.numbps = (uint8_t)g[3]; b.npasses = (uint8_t)g[4];
        b.data_off = (uint32_t)o; b.data_len = (uint32_t)g[5]; b.stepsize = 1.f;
        b.coeff = (int32_t*)(uintptr_t)outn;              // patched below once the device buffer exists
        in += (size_t)g[5]; outn += (size_t)g[0] * g[1];
    }
    arena.insert(arena.end(), 32, 0xff);
    int32_t* d_out = nullptr; uint8_t* d_seg = nullptr; T1DecBlock* d_blk = nullptr;
    CK(cudaMalloc(&d_out, (outn ? outn : 1) * 4)); CK(cudaMalloc(&d_seg, arena.size())); CK(cudaMalloc(&d_blk, sizeof(T1DecBlock) * (nblk ? nblk : 1)));
    for (auto& b : hb) b.coeff = d_out + (uintptr_t)b.coeff;
    CK(cudaMemcpy(d_seg, arena.data(), arena.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_blk, hb.data(), sizeof(T1DecBlock) * nblk, cudaMemcpyHostToDevice));
    launch_t1_decode(d_blk, nblk, d_seg, nullptr, 0, 1, 0);
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpy(out, d_out, outn * 4, cudaMemcpyDeviceToHost);
    cudaFree(d_out); cudaFree(d_seg); cudaFree(d_blk);
    if (e != cudaSuccess) return set_error(J2KCUDA_ERR_CUDA, cudaGetErrorString(e));
    return J2KCUDA_OK;
}); }

}  // extern "C"
