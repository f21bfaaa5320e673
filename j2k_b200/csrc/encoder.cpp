// Persistent encoder context: device buffers, block tables and the launch sequence
//   upload -> MCT -> DWT levels -> Tier-1 context -> Tier-1 MQ -> compaction -> D2H -> host Tier-2.
// Mirrors the order of opj_tcd_encode_tile (ext/grok/src/lib/openjp2/tcd.cpp:1000-1104); the hand-off
// of per-block results to the host Tier-2 is the one the reference defines for accelerators
// (plugin_bridge.cpp:72-174).  There is no CPU fallback for the pixel path.
#include "../../include/j2kcuda.h"
#include "common.h"
#include "device.h"
#include "t2.h"
#include <algorithm>
#include <atomic>
#include <thread>
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

struct j2kcuda_encoder {
    j2kcuda_params p;
    CodingParams cp;
    int device = 0;
    cudaStream_t st = nullptr;
    int W = 0, H = 0, ncomp = 0, pitch = 0, numres = 0, irreversible = 0;
    int32_t* planeA[4] = {nullptr, nullptr, nullptr, nullptr};
    int32_t* planeB[4] = {nullptr, nullptr, nullptr, nullptr};
    // pixel staging
    uint8_t* d_pix = nullptr; size_t d_pix_cap = 0;
    uint8_t* h_pix = nullptr; size_t h_pix_cap = 0;
    PixelSrc src;
    MctParams mp;
    bool have_input = false;
    // geometry
    std::vector<EncTile> tiles;
    std::vector<std::vector<DwtJob>> level_jobs;     // [level-1]
    std::vector<int> level_items, level_seg;
    DwtJob* d_jobs = nullptr;                        // all levels, concatenated
    DwtPxJob* d_px_jobs = nullptr; int n_px_jobs = 0, px_items = 0;   // level 1 fused with unpack + shift + RCT/ICT (one job per tile)
    std::vector<int> level_off;
    // tier-1
    int nblocks = 0;
    std::vector<T1Block> h_blocks;
    std::vector<int> blk_tile, blk_index;            // global block -> (tile, index in tile)
    T1Block* d_blocks = nullptr;
    uint8_t* d_sym = nullptr; size_t sym_bytes = 0;
    uint8_t* d_out = nullptr; size_t out_bytes = 0;
    uint32_t* d_pass_nsym = nullptr;
    T1Result* d_results = nullptr;
    uint32_t* d_offsets = nullptr;
    uint8_t* d_packed = nullptr;
    T1Result* h_results = nullptr;                   // pinned
    uint32_t* h_offsets = nullptr;                   // pinned
    uint8_t* h_packed = nullptr; size_t h_packed_cap = 0;   // pinned
    size_t last_packed = 0;          // packed bytes of the previous frame of this context: sizes the speculative D2H
    std::vector<std::vector<double>> dist;           // per block cumulative distortion
    int mq_lanes = 4;
    uint32_t* d_mq_order = nullptr; int mq_nlong = 592;  // MQ kernel: blocks by decreasing symbol count, at most mq_nlong of the longest get a warp each
    int row0 = 0, row1 = 0;    // image rows covered by this context's tiles (all rows unless a tile range is set)
    // timing
    cudaEvent_t ev[10];
    cudaEvent_t ev_wait = nullptr;   // blocking-sync event: the host worker sleeps instead of spinning (many workers per core)
    float ms[8] = {0};
    uint64_t counters[6] = {0};
    bool ran = false;
};

static void free_encoder(j2kcuda_encoder* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    for (int c = 0; c < 4; ++c) { cudaFree(e->planeA[c]); cudaFree(e->planeB[c]); }
    cudaFree(e->d_pix); cudaFreeHost(e->h_pix);
    cudaFree(e->d_mq_order); cudaFree(e->d_jobs); cudaFree(e->d_px_jobs); cudaFree(e->d_blocks); cudaFree(e->d_sym); cudaFree(e->d_out);
    cudaFree(e->d_pass_nsym); cudaFree(e->d_results); cudaFree(e->d_offsets); cudaFree(e->d_packed);
    cudaFreeHost(e->h_results); cudaFreeHost(e->h_offsets); cudaFreeHost(e->h_packed);
    for (auto& ev : e->ev) if (ev) cudaEventDestroy(ev);
    if (e->ev_wait) cudaEventDestroy(e->ev_wait);
    if (e->st) cudaStreamDestroy(e->st);
    delete e;
}

static int setup_encoder(j2kcuda_encoder* e) {
    const j2kcuda_params& p = e->p;
    if (const char* why = check_params(p)) return set_error(J2KCUDA_ERR_PARAM, why);
    setup_coding_params(p, e->cp);
    const CodingParams& cp = e->cp;
    e->W = (int)p.width; e->H = (int)p.height; e->ncomp = (int)p.ncomp;
    e->pitch = (e->W + 3) & ~3;
    e->numres = cp.comps[0].numres;
    e->irreversible = cp.comps[0].qmfbid == 0;
    for (int c = 0; c < e->ncomp; ++c) {
        if (cp.comps[c].numres != e->numres || cp.comps[c].dx != 1 || cp.comps[c].dy != 1)
            return set_error(J2KCUDA_ERR_UNSUPPORTED, "sub-sampled components are not supported by the encoder");
        if ((1 << cp.comps[c].cblkw) > 64 || (1 << cp.comps[c].cblkh) > 64)
            return set_error(J2KCUDA_ERR_UNSUPPORTED, "code-blocks larger than 64x64 samples");
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return set_error(J2KCUDA_ERR_NO_DEVICE, "no CUDA device: libj2kcuda has no CPU path");
    if (p.device >= 0) e->device = p.device; else cudaGetDevice(&e->device);
    CK(cudaSetDevice(e->device));
    if (const char* err = device_tables_init()) return set_error(J2KCUDA_ERR_CUDA, err);
    CK(cudaStreamCreateWithFlags(&e->st, cudaStreamNonBlocking));
    for (auto& ev : e->ev) CK(cudaEventCreate(&ev));
    CK(cudaEventCreateWithFlags(&e->ev_wait, cudaEventBlockingSync | cudaEventDisableTiming));
    const size_t plane_bytes = (size_t)e->pitch * e->H * sizeof(int32_t);
    for (int c = 0; c < e->ncomp; ++c) {
        CK(cudaMalloc(&e->planeA[c], plane_bytes));
        CK(cudaMalloc(&e->planeB[c], plane_bytes));
    }
    // ---- geometry: tiles, DWT jobs per level, block table ----
    const int nt = (int)(cp.tw * cp.th);
    e->tiles.resize(nt);
    const int levels = e->numres - 1;
    e->level_jobs.assign(levels, {});
    size_t sym_units = 0, out_off = 0;
    e->row0 = e->H; e->row1 = 0;
    for (int t = 0; t < nt; ++t) {
        if (!cp.tile_in_range(t)) continue;           // another rank's tile (tile ranges, config C4)
        init_enc_tile(cp, t, e->tiles[t]);
        const TileGeo& g = e->tiles[t].geo;
        e->row0 = std::min(e->row0, g.comps[0].r.y0); e->row1 = std::max(e->row1, g.comps[0].r.y1);
        for (int c = 0; c < e->ncomp; ++c) {
            const CompGeo& cg = g.comps[c];
            const size_t origin = (size_t)cg.r.y0 * e->pitch + cg.r.x0;
            for (int l = 1; l <= levels; ++l) {
                const ResGeo& in = cg.res[e->numres - l];          // region this level transforms
                if (in.r.w() <= 0 || in.r.h() <= 0) continue;
                DwtJob j;
                int32_t* from = ((l - 1) & 1) ? e->planeB[c] : e->planeA[c];
                int32_t* to = (l & 1) ? e->planeB[c] : e->planeA[c];
                j.src = from + origin; j.dst = to + origin;
                j.x0 = in.r.x0; j.y0 = in.r.y0; j.x1 = in.r.x1; j.y1 = in.r.y1;
                j.pitch = e->pitch; j.strip0 = 0; j.nstrips = 0; j.nsegs = 0;
                e->level_jobs[l - 1].push_back(j);
            }
            (void)origin;
        }
        for (size_t k = 0; k < g.cblks.size(); ++k) {
            const CblkGeo& cb = g.cblks[k];
            const CompGeo& cg = g.comps[cb.comp];
            const BandGeo& band = cg.res[cb.res].bands[cb.band];
            T1Block b;
            memset(&b, 0, sizeof(b));
            const int level = cb.res == 0 ? levels : e->numres - cb.res;     // decomposition level holding the band
            const int32_t* plane = levels == 0 ? e->planeA[cb.comp] : ((level & 1) ? e->planeB[cb.comp] : e->planeA[cb.comp]);
            b.coeff = plane + (size_t)(cg.r.y0 + cb.py) * e->pitch + (cg.r.x0 + cb.px);
            b.pitch = e->pitch;
            b.w = (uint16_t)cb.r.w(); b.h = (uint16_t)cb.r.h();
            b.orient = (uint8_t)cb.orient;
            b.qmfbid = (uint8_t)cp.comps[cb.comp].qmfbid;
            b.max_planes = (uint16_t)std::max(band.numbps, 1);
            b.bandconst = band_const(band.stepsize);
            // worst case per plane: one decision per sample plus two extra per run-length column,
            // and one sign per sample overall
            const size_t area = (size_t)b.w * b.h;
            size_t cap = area * ((size_t)b.max_planes + 2) * 3 / 2 + area + 64;
            cap = (cap + 15) & ~(size_t)15;
            b.sym_off = (uint32_t)sym_units; b.sym_cap = (uint32_t)cap;
            sym_units += cap / 16;
            size_t ocap = 64 + area * ((size_t)b.max_planes + 2) / 4;
            if (cp.comps[cb.comp].cblksty & 4) ocap += 3 * 3 * ((size_t)b.max_planes + 1);      // TERMALL: a flush (up to 3 bytes) after every pass
            ocap = (ocap + 15) & ~(size_t)15;
            b.out_off = (uint32_t)out_off; b.out_cap = (uint32_t)ocap;
            out_off += ocap;
            e->h_blocks.push_back(b);
            e->blk_tile.push_back(t); e->blk_index.push_back((int)k);
        }
    }
    if (sym_units * 16 > 0xffffffffull * 16 || out_off > 0xffffffffull)
        return set_error(J2KCUDA_ERR_UNSUPPORTED, "frame too large for 32-bit arena offsets");
    e->nblocks = (int)e->h_blocks.size();
    e->sym_bytes = sym_units * 16; e->out_bytes = out_off;
    // DWT job tables
    std::vector<DwtJob> all;
    e->level_off.assign(levels, 0); e->level_items.assign(levels, 0); e->level_seg.assign(levels, 0);
    for (int l = 0; l < levels; ++l) {
        auto& v = e->level_jobs[l];
        e->level_off[l] = (int)all.size();
        if (v.empty()) continue;
        e->level_items[l] = dwt_plan_jobs(v.data(), (int)v.size());
        e->level_seg[l] = v[0].strip0;
        all.insert(all.end(), v.begin(), v.end());
    }
    if (!all.empty()) {
        CK(cudaMalloc(&e->d_jobs, all.size() * sizeof(DwtJob)));
        CK(cudaMemcpy(e->d_jobs, all.data(), all.size() * sizeof(DwtJob), cudaMemcpyHostToDevice));
    }
    // level 1 fused with the sample unpacking and the component transform: one job per tile over all components
    if (levels >= 1 && e->ncomp >= 3 && dwt_px_fusion_available() && !e->level_jobs[0].empty() && e->level_jobs[0].size() % (size_t)e->ncomp == 0) {
        const auto& v = e->level_jobs[0];                    // tile-major, then component
        std::vector<DwtPxJob> pj;
        bool ok = true;
        for (size_t i = 0; i < v.size() && ok; i += (size_t)e->ncomp) {
            DwtPxJob j;
            memset(&j, 0, sizeof(j));
            for (int c = 0; c < e->ncomp; ++c) {
                const DwtJob& d = v[i + (size_t)c];
                ok = ok && d.x0 == v[i].x0 && d.y0 == v[i].y0 && d.x1 == v[i].x1 && d.y1 == v[i].y1 && d.y1 - d.y0 >= 2;
                j.dst[c] = d.dst;
                j.dc[c] = cp.comps[c].sgnd ? 0 : 1 << (cp.comps[c].prec - 1);
            }
            j.x0 = v[i].x0; j.y0 = v[i].y0; j.x1 = v[i].x1; j.y1 = v[i].y1; j.pitch = v[i].pitch;
            j.strip0 = v[i].strip0; j.nstrips = v[i].nstrips; j.nsegs = v[i].nsegs; j.strip_cols = v[i].strip_cols;
            j.mct = cp.mct ? 1 : 0; j.irreversible = e->irreversible;
            pj.push_back(j);
        }
        if (ok) {
            CK(cudaMalloc(&e->d_px_jobs, pj.size() * sizeof(DwtPxJob)));
            CK(cudaMemcpy(e->d_px_jobs, pj.data(), pj.size() * sizeof(DwtPxJob), cudaMemcpyHostToDevice));
            e->n_px_jobs = (int)pj.size(); e->px_items = e->level_items[0];
        }
    }
    if (e->nblocks) {
        CK(cudaMalloc(&e->d_blocks, e->nblocks * sizeof(T1Block)));
        CK(cudaMemcpy(e->d_blocks, e->h_blocks.data(), e->nblocks * sizeof(T1Block), cudaMemcpyHostToDevice));
        CK(cudaMalloc(&e->d_sym, e->sym_bytes + 64));      // the MQ kernel reads one 8-byte word ahead
        CK(cudaMalloc(&e->d_out, e->out_bytes));
        CK(cudaMalloc(&e->d_packed, e->out_bytes));
        CK(cudaMalloc(&e->d_pass_nsym, (size_t)e->nblocks * T1_MAX_PASSES * sizeof(uint32_t)));
        CK(cudaMalloc(&e->d_results, (size_t)e->nblocks * sizeof(T1Result)));
        CK(cudaMalloc(&e->d_offsets, ((size_t)e->nblocks + 1) * sizeof(uint32_t)));
        CK(cudaMalloc(&e->d_mq_order, ((size_t)e->nblocks + 2) * sizeof(uint32_t)));
        CK(cudaMallocHost(&e->h_results, (size_t)e->nblocks * sizeof(T1Result)));
        CK(cudaMallocHost(&e->h_offsets, ((size_t)e->nblocks + 1) * sizeof(uint32_t)));
    }
    e->dist.assign(e->nblocks, {});
    e->mq_lanes = e->nblocks >= 4096 ? 8 : 4;
    if (const char* s = getenv("J2KCUDA_MQ_LANES")) { int v = atoi(s); if (v >= 1 && v <= 32) e->mq_lanes = v; }
    if (const char* s = getenv("J2KCUDA_MQ_NLONG")) e->mq_nlong = atoi(s);
    // sample conversion parameters
    memset(&e->mp, 0, sizeof(e->mp));
    e->mp.width = e->W; e->mp.height = e->H; e->mp.pitch = e->pitch; e->mp.ncomp = e->ncomp;
    e->mp.mct = cp.mct ? 1 : 0; e->mp.irreversible = e->irreversible;
    for (int c = 0; c < e->ncomp; ++c) {
        e->mp.prec[c] = cp.comps[c].prec;
        e->mp.dc_shift[c] = cp.comps[c].sgnd ? 0 : 1 << (cp.comps[c].prec - 1);
    }
    return J2KCUDA_OK;
}

static int sample_bytes_of(int sample_type) {
    return sample_type == J2KCUDA_UCHAR ? 1 : (sample_type == J2KCUDA_USHORT ? 2 : 4);
}

// Host planes -> device.  One DMA per contiguous host range: directly from the caller's memory when
// it is page-locked, through the context's pinned staging buffer otherwise.
static int upload_planes(j2kcuda_encoder* e, const j2kcuda_plane* planes, int nplanes) {
    if (nplanes < e->ncomp) return set_error(J2KCUDA_ERR_PARAM, "fewer planes than components");
    const int S = sample_bytes_of(planes[0].sample_type);
    struct Range { const uint8_t* lo; const uint8_t* hi; size_t dev; };
    std::vector<Range> ranges;
    for (int c = 0; c < e->ncomp; ++c) {
        const j2kcuda_plane& pl = planes[c];
        if (!pl.buf || (int)pl.width != e->W || (int)pl.height != e->H)
            return set_error(J2KCUDA_ERR_PARAM, "plane size differs from the image size");
        if (sample_bytes_of(pl.sample_type) != S || pl.colbytes < S || pl.rowbytes < 0)
            return set_error(J2KCUDA_ERR_PARAM, "planes must share one sample size and have positive strides");
        if (pl.sample_type > J2KCUDA_INT || pl.depth > 8 * S)
            return set_error(J2KCUDA_ERR_PARAM, "plane depth does not fit its sample type");
        const uint8_t* lo = (const uint8_t*)pl.buf + (size_t)e->row0 * pl.rowbytes;     // only the rows of this context's tiles
        const uint8_t* hi = (const uint8_t*)pl.buf + (size_t)(e->row1 - 1) * pl.rowbytes + (size_t)(e->W - 1) * pl.colbytes + S;
        ranges.push_back({lo, hi, 0});
    }
    // merge overlapping / adjacent ranges
    std::vector<Range> merged;
    {
        std::vector<Range> s = ranges;
        std::sort(s.begin(), s.end(), [](const Range& a, const Range& b) { return a.lo < b.lo; });
        for (auto& r : s) {
            if (!merged.empty() && r.lo <= merged.back().hi) { if (r.hi > merged.back().hi) merged.back().hi = r.hi; }
            else merged.push_back(r);
        }
    }
    size_t total = 0;
    for (auto& m : merged) {
        total = (total + 255) & ~(size_t)255;
        total += (uintptr_t)m.lo & 15;                 // keep the caller's alignment modulo 16
        m.dev = total;
        total += (size_t)(m.hi - m.lo);
    }
    if (total > e->d_pix_cap) {
        cudaFree(e->d_pix); e->d_pix = nullptr; e->d_pix_cap = 0;
        CK(cudaMalloc(&e->d_pix, total + 256));
        e->d_pix_cap = total;
    }
    uint64_t h2d = 0;
    for (auto& m : merged) {
        const size_t n = (size_t)(m.hi - m.lo);
        cudaPointerAttributes attr;
        bool pinned = cudaPointerGetAttributes(&attr, m.lo) == cudaSuccess && attr.type == cudaMemoryTypeHost;
        cudaGetLastError();
        if (pinned) {
            CK(cudaMemcpyAsync(e->d_pix + m.dev, m.lo, n, cudaMemcpyHostToDevice, e->st));
        } else {
            if (total > e->h_pix_cap) {
                CK(stream_wait(e->st, e->ev_wait));
                cudaFreeHost(e->h_pix); e->h_pix = nullptr; e->h_pix_cap = 0;
                CK(cudaMallocHost(&e->h_pix, total + 256));
                e->h_pix_cap = total;
            }
            // pageable caller memory (the plug-in's case): staged through the page-locked buffer in 4 MB chunks, the DMA of
            // one chunk under the memcpy of the next; one thread copies ~9 GB/s, so large frames use a few
            const size_t chunk = 4u << 20;
            const size_t nchunks = (n + chunk - 1) / chunk;
            const int nthreads = nchunks >= 8 ? 4 : 1;
            std::atomic<int> failed{0};
            auto work = [&](int t) {
                cudaSetDevice(e->device);
                for (size_t ci = (size_t)t; ci < nchunks; ci += (size_t)nthreads) {
                    const size_t off = ci * chunk, k = std::min(chunk, n - off);
                    memcpy(e->h_pix + m.dev + off, m.lo + off, k);
                    if (cudaMemcpyAsync(e->d_pix + m.dev + off, e->h_pix + m.dev + off, k, cudaMemcpyHostToDevice, e->st) != cudaSuccess) failed = 1;
                }
            };
            if (nthreads == 1) work(0);
            else {
                std::vector<std::thread> th;
                for (int t = 1; t < nthreads; ++t) th.emplace_back(work, t);
                work(0);
                for (auto& x : th) x.join();
            }
            if (failed) { CK(cudaGetLastError()); return set_error(J2KCUDA_ERR_CUDA, "upload failed"); }
        }
        h2d += n;
    }
    memset(&e->src, 0, sizeof(e->src));
    e->src.sample_bytes = S;
    e->src.sgnd = 0;                                   // sign handling lives in mp.conv (copybuffer.cuh)
    for (int c = 0; c < e->ncomp; ++c) {
        const uint8_t* lo = (const uint8_t*)planes[c].buf + (size_t)e->row0 * planes[c].rowbytes;
        for (auto& m : merged)          // base = where row 0 would be (rows above row0 are never touched)
            if (lo >= m.lo && lo < m.hi) { e->src.base[c] = e->d_pix + m.dev + (lo - m.lo) - (size_t)e->row0 * planes[c].rowbytes; break; }
        e->src.colbytes[c] = planes[c].colbytes;
        e->src.rowbytes[c] = planes[c].rowbytes;
        const int depth = planes[c].depth ? planes[c].depth : e->mp.prec[c];
        e->mp.conv[c] = make_copy_conv(planes[c].sample_type, depth, planes[c].sgnd, J2KCUDA_INT, e->mp.prec[c], (int)e->p.sgnd);
    }
    e->counters[4] = h2d;
    e->have_input = true;
    return J2KCUDA_OK;
}

// Per code-block results of the last run, in Tier-2 order: the cumulative weighted distortion of every pass
// (t1_opt.cpp:711-722 with t1.cpp:1114-1143, in double and in pass order like the reference) and the block's bytes / rates.
static void fill_block_results(j2kcuda_encoder* e) {
    const CodingParams& cp = e->cp;
    for (int i = 0; i < e->nblocks; ++i) {
        const T1Result& r = e->h_results[i];
        EncTile& tile = e->tiles[e->blk_tile[i]];
        const CblkGeo& cb = tile.geo.cblks[e->blk_index[i]];
        const BandGeo& band = tile.geo.comps[cb.comp].res[cb.res].bands[cb.band];
        EncBlock& b = tile.blocks[e->blk_index[i]];
        const bool real = cp.comps[cb.comp].qmfbid == 0;
        const int level = cp.comps[cb.comp].numres - 1 - cb.res;
        const double w1 = (cp.mct && cb.comp < 3 && cp.ncomp >= 3) ? mct_norm(cb.comp, real) : 1.0;
        const double w2 = dwt_norm(level, cb.orient, real);
        std::vector<double>& d = e->dist[i];
        d.resize(r.npasses + 1);
        double cum = 0;
        int bpno = r.numbps - 1, passtype = 2;
        for (int k = 0; k < r.npasses; ++k) {
            double w = w1 * w2 * (double)band.stepsize * (double)(1 << bpno);
            w *= w * r.nmsedec[k] / 8192.0;
            cum += w;
            d[k] = cum;
            if (++passtype == 3) { passtype = 0; --bpno; }
        }
        b.data = e->h_packed + e->h_offsets[i];
        b.numbps = r.numbps; b.totalpasses = r.npasses;
        b.rates = r.rate; b.dist = d.data();
    }
}

// ---- contexts kept between one-shot calls (WriteFile after WriteFile with the same settings) ----
namespace {
std::mutex g_pool_mu;
std::vector<j2kcuda_encoder*> g_pool;
constexpr size_t POOL_MAX = 2;
j2kcuda_encoder* pool_take(const j2kcuda_params& p) {
    if (!context_cache_enabled()) return nullptr;
    int cur = 0;
    if (p.device < 0 && cudaGetDevice(&cur) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    std::lock_guard<std::mutex> g(g_pool_mu);
    for (size_t i = 0; i < g_pool.size(); ++i)
        if (memcmp(&g_pool[i]->p, &p, sizeof(p)) == 0 && (p.device >= 0 || g_pool[i]->device == cur)) { j2kcuda_encoder* e = g_pool[i]; g_pool.erase(g_pool.begin() + (long)i); return e; }
    return nullptr;
}
void pool_give(j2kcuda_encoder* e) {
    j2kcuda_encoder* drop = nullptr;
    if (!context_cache_enabled()) drop = e;
    else {
        std::lock_guard<std::mutex> g(g_pool_mu);
        g_pool.push_back(e);
        if (g_pool.size() > POOL_MAX) { drop = g_pool.front(); g_pool.erase(g_pool.begin()); }
    }
    if (drop) free_encoder(drop);
}
}  // namespace
namespace j2k {
void encoder_pool_clear() {
    std::vector<j2kcuda_encoder*> all;
    { std::lock_guard<std::mutex> g(g_pool_mu); all.swap(g_pool); }
    for (auto* e : all) free_encoder(e);
}
}  // namespace j2k

extern "C" {

j2kcuda_encoder* j2kcuda_encoder_create(const j2kcuda_params* p) { try {
    if (!p) { set_error(J2KCUDA_ERR_PARAM, "null params"); return nullptr; }
    j2kcuda_encoder* e = new j2kcuda_encoder();
    for (auto& ev : e->ev) ev = nullptr;
    e->p = *p;
    if (setup_encoder(e) != J2KCUDA_OK) { free_encoder(e); return nullptr; }
    return e;
} catch (...) { set_error(J2KCUDA_ERR_NOMEM, "j2kcuda_encoder_create: out of memory or internal error"); return nullptr; } }

void j2kcuda_encoder_destroy(j2kcuda_encoder* e) { free_encoder(e); }

int j2kcuda_encoder_upload(j2kcuda_encoder* e, const j2kcuda_plane* planes, int nplanes) { return j2k::guarded([&]() -> int {
    if (!e || !planes) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    CK(cudaSetDevice(e->device));
    CK(cudaEventRecord(e->ev[8], e->st));
    int rc = upload_planes(e, planes, nplanes);
    if (rc) return rc;
    CK(cudaEventRecord(e->ev[9], e->st));
    return J2KCUDA_OK;
}); }

int j2kcuda_encoder_set_device_input(j2kcuda_encoder* e, const void* d_pixels, int64_t pitch_bytes, int sample_type) { return j2k::guarded([&]() -> int {
    if (!e || !d_pixels) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    const int S = sample_bytes_of(sample_type);
    memset(&e->src, 0, sizeof(e->src));
    e->src.sample_bytes = S;
    e->src.sgnd = 0;
    for (int c = 0; c < e->ncomp; ++c) {
        e->src.base[c] = (const uint8_t*)d_pixels + (size_t)c * S;
        e->src.colbytes[c] = (int64_t)e->ncomp * S;
        e->src.rowbytes[c] = pitch_bytes;
        e->mp.conv[c] = make_copy_conv(sample_type, e->mp.prec[c], (int)e->p.sgnd, J2KCUDA_INT, e->mp.prec[c], (int)e->p.sgnd);
    }
    CK(cudaSetDevice(e->device));
    CK(cudaEventRecord(e->ev[8], e->st));
    CK(cudaEventRecord(e->ev[9], e->st));
    e->counters[4] = 0;
    e->have_input = true;
    return J2KCUDA_OK;
}); }

int j2kcuda_encoder_run(j2kcuda_encoder* e) { return j2k::guarded([&]() -> int {
    if (!e) return set_error(J2KCUDA_ERR_PARAM, "null encoder");
    if (!e->have_input) return set_error(J2KCUDA_ERR_PARAM, "no input uploaded");
    CK(cudaSetDevice(e->device));
    cudaStream_t st = e->st;
    uint64_t launches = 0;
    CK(cudaEventRecord(e->ev[0], st));
    // Fused first level: straight from the caller's interleaved pixels when they need no depth / sign conversion
    bool fused = false;
    if (e->n_px_jobs && (e->src.sample_bytes == 1 || e->src.sample_bytes == 2)) {
        DwtPxSrc ps;
        memset(&ps, 0, sizeof(ps));
        const int nch = detect_pixel_interleave(e->src, e->ncomp, ps.chan);
        bool plain = nch == 3 || nch == 4;
        for (int c = 0; c < e->ncomp && plain; ++c) plain = e->mp.conv[c].plain != 0;
        if (plain) {
            const uint8_t* lo = e->src.base[0];
            for (int c = 1; c < e->ncomp; ++c) if (e->src.base[c] < lo) lo = e->src.base[c];
            ps.px = lo; ps.rowbytes = e->src.rowbytes[0];
            CK(cudaEventRecord(e->ev[1], st));
            fused = launch_dwt_forward_px(e->d_px_jobs, ps, e->n_px_jobs, e->px_items, e->irreversible ? DWT_97_FIX : DWT_53, e->ncomp, nch,
                                          e->src.sample_bytes, st);
            if (fused) ++launches;
        }
    }
    if (!fused) {
        // only the image rows this context's tiles cover (all of them unless a tile range is set)
        PixelSrc src = e->src;
        MctParams mp = e->mp;
        int32_t* planes[4] = {nullptr, nullptr, nullptr, nullptr};
        for (int c = 0; c < e->ncomp; ++c) {
            src.base[c] = e->src.base[c] + (int64_t)e->row0 * e->src.rowbytes[c];
            planes[c] = e->planeA[c] + (size_t)e->row0 * e->pitch;
        }
        mp.height = e->row1 - e->row0;
        if (mp.height > 0) { launch_mct_forward(src, mp, planes, st); ++launches; }
    }
    if (!fused) CK(cudaEventRecord(e->ev[1], st));
    const int levels = e->numres - 1;
    for (int l = 0; l < levels; ++l) {
        if (e->level_jobs[l].empty()) continue;
        if (l == 1) CK(cudaEventRecord(e->ev[6], st));
        if (l == 0 && fused) continue;
        launch_dwt_forward(e->d_jobs + e->level_off[l], (int)e->level_jobs[l].size(), e->level_items[l], e->level_seg[l],
                           e->irreversible ? DWT_97_FIX : DWT_53, st);
        ++launches;
    }
    if (levels < 2) CK(cudaEventRecord(e->ev[6], st));
    CK(cudaEventRecord(e->ev[2], st));
    launch_t1_context(e->d_blocks, e->nblocks, e->d_sym, e->d_pass_nsym, e->d_results, st); ++launches;
    CK(cudaEventRecord(e->ev[3], st));
    launch_t1_mq(e->d_blocks, e->nblocks, e->d_sym, e->d_pass_nsym, e->d_out, e->d_results, e->mq_lanes, e->d_mq_order, e->mq_nlong, st, e->cp.comps[0].cblksty); launches += e->cp.comps[0].cblksty ? 1 : 2;
    CK(cudaEventRecord(e->ev[4], st));
    launch_t1_compact(e->d_blocks, e->nblocks, e->d_out, e->d_results, e->d_offsets, e->d_packed, st); launches += 2;
    if (e->nblocks) {
        CK(cudaMemcpyAsync(e->h_results, e->d_results, (size_t)e->nblocks * sizeof(T1Result), cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(e->h_offsets, e->d_offsets, ((size_t)e->nblocks + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    }
    // The packed code bytes follow in the same wait: their size is only known on the device at this point, so the copy is sized
    // from the previous frame of this context (+6 %, frames of a sequence differ by a few percent); the rare remainder is
    // fetched afterwards.  One blocking wait per frame instead of two: on a busy host every wake-up costs scheduler latency.
    size_t guess = 0;
    static const bool speculate = [] { const char* v = getenv("J2KCUDA_SPEC_D2H"); return !(v && v[0] == '0'); }();
    if (speculate && e->nblocks && e->last_packed && e->h_packed) {
        guess = e->last_packed + e->last_packed / 16 + 65536;
        if (guess > e->h_packed_cap) guess = e->h_packed_cap;
        if (guess > e->out_bytes) guess = e->out_bytes;
        if (guess) CK(cudaMemcpyAsync(e->h_packed, e->d_packed, guess, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaEventRecord(e->ev[5], st));
    CK(stream_wait(st, e->ev_wait));
    CK(cudaGetLastError());
    size_t packed = e->nblocks ? e->h_offsets[e->nblocks] : 0;
    if (packed > e->h_packed_cap) {
        cudaFreeHost(e->h_packed); e->h_packed = nullptr; e->h_packed_cap = 0;
        size_t cap = packed + packed / 4 + 4096;
        CK(cudaMallocHost(&e->h_packed, cap));
        e->h_packed_cap = cap;
        guess = 0;
    }
    if (packed > guess) {
        CK(cudaMemcpyAsync(e->h_packed + guess, e->d_packed + guess, packed - guess, cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(e->ev[5], st));
        CK(stream_wait(st, e->ev_wait));
    }
    e->last_packed = packed;
    uint64_t nsym = 0;
    for (int i = 0; i < e->nblocks; ++i) {
        if (e->h_results[i].status != 0)
            return set_error(J2KCUDA_ERR_CUDA, e->h_results[i].status == 1 ? "tier-1 symbol slot overflow" : "tier-1 byte slot overflow");
        nsym += e->h_results[i].nsym;
    }
    cudaEventElapsedTime(&e->ms[0], e->ev[8], e->ev[9]);
    cudaEventElapsedTime(&e->ms[1], e->ev[0], e->ev[1]);
    cudaEventElapsedTime(&e->ms[2], e->ev[1], e->ev[2]);
    cudaEventElapsedTime(&e->ms[3], e->ev[2], e->ev[3]);
    cudaEventElapsedTime(&e->ms[4], e->ev[3], e->ev[4]);
    cudaEventElapsedTime(&e->ms[5], e->ev[4], e->ev[5]);
    cudaEventElapsedTime(&e->ms[7], e->ev[1], e->ev[6]);
    e->counters[0] = launches; e->counters[1] = (uint64_t)e->nblocks; e->counters[2] = nsym;
    e->counters[3] = packed;
    e->counters[5] = (uint64_t)e->nblocks * sizeof(T1Result) + ((uint64_t)e->nblocks + 1) * 4 + packed;
    e->ran = true;
    return J2KCUDA_OK;
}); }

int j2kcuda_encoder_finish(j2kcuda_encoder* e, uint8_t** out, size_t* out_len) { return j2k::guarded([&]() -> int {
    if (!e || !out || !out_len) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    if (!e->ran) return set_error(J2KCUDA_ERR_PARAM, "encoder_run has not completed");
    auto t0 = std::chrono::steady_clock::now();
    const CodingParams& cp = e->cp;
    for (size_t t = 0; t < e->tiles.size(); ++t) {
        if (!cp.tile_in_range((int)t)) continue;
        EncTile& tile = e->tiles[t];
        if (tile.geo.cblks.empty()) { init_enc_tile(cp, (int)t, tile); continue; }
        // the geometry and the tag trees of a context never change: only the per-frame block state is reset
        for (auto& b : tile.blocks) {
            b.numpasses = 0; b.numlenbits = 0; b.included_other = 0;
            for (auto& l : b.layers) l = EncLayer();
        }
        tile.distotile = 0; tile.packno = 0;
    }
    fill_block_results(e);
    EncoderHost host;
    host.cp = cp;
    host.ntiles = (int)e->tiles.size();
    ByteBuf cs;                                      // built in place in the buffer the caller receives (recycled by j2kcuda_free)
    std::string err;
    if (!host.assemble(e->tiles, cs, err)) return set_error(J2KCUDA_ERR_PARAM, err);
    *out_len = cs.size();
    *out = cs.release();
    e->ms[6] = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return J2KCUDA_OK;
}); }

// What grok's accelerator hand-off needs (plugin_bridge.cpp:72-174, openjpeg.h:1308-1360): the device results of every
// code-block after j2kcuda_encoder_run, in Tier-2 order.  The pointers stay valid until the next upload / run on this context.
int j2kcuda_encoder_block_results(j2kcuda_encoder* e, j2kcuda_block_result* out, int max_blocks) { return j2k::guarded([&]() -> int {
    if (!e || !out) { set_error(J2KCUDA_ERR_PARAM, "null argument"); return -1; }
    if (!e->ran) { set_error(J2KCUDA_ERR_PARAM, "encoder_run has not completed"); return -1; }
    if (max_blocks < e->nblocks) { set_error(J2KCUDA_ERR_PARAM, "block array too small"); return -1; }
    fill_block_results(e);
    for (int i = 0; i < e->nblocks; ++i) {
        const EncBlock& b = e->tiles[e->blk_tile[i]].blocks[e->blk_index[i]];
        out[i].data = b.data; out[i].numbps = b.numbps; out[i].npasses = b.totalpasses; out[i].rates = b.rates; out[i].dist = b.dist;
    }
    return e->nblocks;
}); }

// The DWT coefficients of one component after j2kcuda_encoder_run, assembled into the reference's single in-place tile buffer
// (Mallat layout, tile_buf.cpp:147-157): grok's OPJ_PLUGIN_STATE_DEBUG_ENCODE cross-check wants them as the image data
// (tcd.cpp:1042-1079).  out = width x height int32, row-major.  Single-tile frames only.
int j2kcuda_encoder_download_coefficients(j2kcuda_encoder* e, int comp, int32_t* out) { return j2k::guarded([&]() -> int {
    if (!e || !out || comp < 0 || comp >= e->ncomp) return set_error(J2KCUDA_ERR_PARAM, "bad argument");
    if (!e->ran) return set_error(J2KCUDA_ERR_PARAM, "encoder_run has not completed");
    if (e->tiles.size() != 1) return set_error(J2KCUDA_ERR_UNSUPPORTED, "coefficient download is for single-tile frames");
    CK(cudaSetDevice(e->device));
    const CompGeo& cg = e->tiles[0].geo.comps[comp];
    const int levels = e->numres - 1;
    for (int r = 0; r < e->numres; ++r) {
        const ResGeo& rg = cg.res[r];
        for (int b = 0; b < rg.numbands; ++b) {
            const BandGeo& bg = rg.bands[b];
            if (bg.r.w() <= 0 || bg.r.h() <= 0) continue;
            const int level = r == 0 ? levels : e->numres - r;              // decomposition level whose output holds the band
            const int32_t* plane = levels == 0 ? e->planeA[comp] : ((level & 1) ? e->planeB[comp] : e->planeA[comp]);
            const int32_t* src = plane + (size_t)(cg.r.y0 + bg.offy) * e->pitch + (cg.r.x0 + bg.offx);
            CK(cudaMemcpy2DAsync(out + (size_t)bg.offy * e->W + bg.offx, (size_t)e->W * 4, src, (size_t)e->pitch * 4, (size_t)bg.r.w() * 4, (size_t)bg.r.h(),
                                 cudaMemcpyDeviceToHost, e->st));
        }
    }
    CK(stream_wait(e->st, e->ev_wait));
    return J2KCUDA_OK;
}); }

int j2kcuda_encoder_timings(j2kcuda_encoder* e, float* ms, int n) { return j2k::guarded([&]() -> int {
    if (!e || !ms) return 0;
    int k = n < 8 ? n : 8;
    for (int i = 0; i < k; ++i) ms[i] = e->ms[i];
    return k;
}); }

int j2kcuda_encoder_counters(j2kcuda_encoder* e, uint64_t* v, int n) { return j2k::guarded([&]() -> int {
    if (!e || !v) return 0;
    int k = n < 6 ? n : 6;
    for (int i = 0; i < k; ++i) v[i] = e->counters[i];
    return k;
}); }

int j2kcuda_encode(const j2kcuda_params* p, const j2kcuda_plane* planes, int nplanes, uint8_t** out, size_t* out_len) { return j2k::guarded([&]() -> int {
    if (!p) return set_error(J2KCUDA_ERR_PARAM, "null params");
    j2kcuda_encoder* e = pool_take(*p);
    if (!e) e = j2kcuda_encoder_create(p);
    if (!e) return J2KCUDA_ERR_PARAM;
    int rc = j2kcuda_encoder_upload(e, planes, nplanes);
    if (!rc) rc = j2kcuda_encoder_run(e);
    if (!rc) rc = j2kcuda_encoder_finish(e, out, out_len);
    if (rc) j2kcuda_encoder_destroy(e); else pool_give(e);
    return rc;
}); }

}  // extern "C"

namespace j2k {
void encoder_set_throughput_mode(::j2kcuda_encoder* e, bool many) {
    if (!e || getenv("J2KCUDA_MQ_LANES")) return;
    e->mq_lanes = many ? 16 : (e->nblocks >= 4096 ? 8 : 4);
}
}  // namespace j2k
