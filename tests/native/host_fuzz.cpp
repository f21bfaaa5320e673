// TEST INFRASTRUCTURE.  Host-only mutation fuzzer for the stream-facing host code of libj2kcuda: the codestream parser
// (t2_dec.cpp, geometry.cpp), what j2kcuda_probe / j2kcuda_t2_parse do around it, and the JP2 box reader (jp2.cpp).  Built with
// AddressSanitizer + UndefinedBehaviorSanitizer by tests/native/Makefile (target host_fuzz); tests/test_host_fuzz.py runs it
// for a few seconds on the committed golden codestreams, tools or a developer run it for as long as they like:
//     tests/native/host_fuzz <seed directory> <seconds> <rng seed>
// Every mutated stream must end in a status, never in a sanitizer report (the process aborts on the first one).
#include "t2.h"
#include "../../include/j2kcuda.h"
#include <chrono>
#include <cstdio>
#include <cstring>
#include <dirent.h>
#include <random>
#include <string>
#include <vector>
using namespace j2k;
static std::vector<std::vector<uint8_t>> seeds;
// ---- second mode: host Tier-2 writer -> host parser over random geometry ------------------------------------------------
// Random (valid or refused) parameters, random code-block results; whatever j2kcuda_t2_assemble writes, DecoderHost::parse must
// read back: per code-block the same bit-plane count, a pass count the writer could have chosen, and exactly the bytes of
// those passes.  tests/native/host_fuzz --roundtrip <seconds> <rng seed>
static int roundtrip(double seconds, unsigned seed0) {
    std::mt19937 rng(seed0);
    auto t0 = std::chrono::steady_clock::now();
    long iters = 0, written = 0, refused = 0;
    while (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() < seconds) {
        ++iters;
        j2kcuda_params p; j2kcuda_default_params(&p);
        p.width = 1 + rng() % 300; p.height = 1 + rng() % 300; p.ncomp = 1 + rng() % 4; p.prec = 1 + rng() % 16;
        p.irreversible = rng() & 1; p.numresolution = 1 + rng() % 7; p.mct = (p.ncomp >= 3) ? (rng() & 1) : 0;
        const int cw = 2 + rng() % 5, chmax = 12 - cw < 6 ? 12 - cw : 6, ch = 2 + rng() % (chmax - 1);
        p.cblk_w = 1u << cw; p.cblk_h = 1u << ch;
        p.prog_order = rng() % 5;
        if (rng() % 3 == 0) { p.tile_w = 16 + rng() % 200; p.tile_h = 16 + rng() % 200; }
        p.numlayers = 1 + rng() % 4;
        const bool lossy = rng() % 2;
        for (unsigned k = 0; k < p.numlayers; ++k) p.rates[k] = lossy ? (k + 1 == p.numlayers && (rng() & 1) ? 0.0 : 40.0 / (k + 1) + (rng() % 10)) : 0.0;
        p.compat = rng() % 3;
        { static const uint32_t stys[6] = {0, 0, 0x02, 0x04, 0x20, 0x26}; p.cblksty = stys[rng() % 6]; }   // RESET, TERMALL, SEGSYM
        const int nb = j2kcuda_t2_block_count(&p);
        if (nb < 0) { ++refused; continue; }
        if (nb == 0 || nb > 20000) continue;
        std::vector<j2kcuda_block_result> res((size_t)nb);
        std::vector<std::vector<int32_t>> rates((size_t)nb);
        std::vector<std::vector<double>> dist((size_t)nb);
        std::vector<std::vector<uint8_t>> data((size_t)nb);
        std::vector<int32_t> geo(12 * (size_t)nb);
        if (j2kcuda_t2_block_geometry(&p, geo.data(), nb) != nb) { printf("block geometry disagrees with block count\n"); return 2; }
        for (int i = 0; i < nb; ++i) {
            const int band_bps = geo[12 * (size_t)i + 8];             // a block never has more bit-planes than its band (t1_opt.cpp:680)
            int numbps = (rng() % 5 == 0) ? 0 : 1 + rng() % 12;
            if (numbps > band_bps) numbps = band_bps > 0 ? band_bps : 0;
            const int np = numbps ? 1 + rng() % (3 * numbps - 2) : 0;
            int r = 0; double dd = 0;
            rates[i].resize(np + 1); dist[i].resize(np + 1);
            for (int k = 0; k < np; ++k) { r += (rng() % 4 == 0) ? 0 : rng() % 40; dd += (rng() % 4 == 0) ? 0.0 : (double)(rng() % 100000); rates[i][k] = r; dist[i][k] = dd; }
            data[i].resize((size_t)r + 4);
            for (auto& b : data[i]) b = (uint8_t)(rng() % 255);           // no 0xFF: bodies never look like markers
            res[i].data = data[i].data(); res[i].numbps = numbps; res[i].npasses = np; res[i].rates = rates[i].data(); res[i].dist = dist[i].data();
        }
        uint8_t* out = nullptr; size_t n = 0;
        if (j2kcuda_t2_assemble(&p, res.data(), nb, &out, &n) != 0) { ++refused; continue; }
        ++written;
        DecoderHost h; std::string err;
        if (!h.parse(out, n, err)) { printf("parser refused the writer's stream: %s\n", err.c_str()); return 2; }
        int k = 0;
        for (auto& t : h.tiles)
            for (auto& b : t.blocks) {
                if (k >= nb) { printf("more blocks read than written\n"); return 2; }
                if (b.included) {
                    if (b.numbps != res[k].numbps || b.npasses < 1 || b.npasses > res[k].npasses || b.len != rates[k][b.npasses - 1]) {
                        printf("block %d: numbps %d/%d passes %d/%d len %d\n", k, b.numbps, res[k].numbps, b.npasses, res[k].npasses, b.len); return 2;
                    }
                    if (p.cblksty & 4) {                                      // TERMALL: one codeword segment per pass, each with the pass's length
                        if ((int)b.segs.size() != b.npasses) { printf("block %d: %zu segments for %d terminated passes\n", k, b.segs.size(), b.npasses); return 2; }
                        for (int q = 0; q < b.npasses; ++q)
                            if (b.segs[q].numpasses != 1 || b.segs[q].len != rates[k][q] - (q ? rates[k][q - 1] : 0)) { printf("block %d: segment %d differs\n", k, q); return 2; }
                    } else if (b.segs.size() != 1) { printf("block %d: %zu segments without TERMALL\n", k, b.segs.size()); return 2; }
                    size_t pos = 0;
                    for (auto& c : b.chunks) { if (memcmp(out + c.first, data[k].data() + pos, c.second)) { printf("block %d: bytes differ\n", k); return 2; } pos += c.second; }
                    if ((int)pos != b.len) { printf("block %d: chunks do not add up\n", k); return 2; }
                } else if (b.len != 0) { printf("block %d: bytes without inclusion\n", k); return 2; }
                ++k;
            }
        if (k != nb) { printf("%d blocks read, %d written\n", k, nb); return 2; }
        // the same stream inside JP2 boxes (j2kcuda_wrap_jp2) must be found again, with what the boxes say (j2kcuda_probe_jp2)
        if (p.ncomp <= 4) {
            j2kcuda_jp2_params jp; memset(&jp, 0, sizeof(jp));
            jp.width = p.width; jp.height = p.height; jp.ncomp = p.ncomp;
            const bool mixed = rng() % 4 == 0;
            const unsigned sg = rng() & 1;        // one sign for all: like the reference, the writer falls back to bpcc only when DEPTHS differ (jp2.cpp:2115-2124)
            for (unsigned c = 0; c < p.ncomp; ++c) { jp.prec[c] = mixed ? 1 + rng() % 16 : p.prec; jp.sgnd[c] = sg; }
            jp.enumcs = p.ncomp >= 3 ? 16 : 17;
            std::vector<uint8_t> icc;
            if (rng() % 3 == 0) { icc.resize(128 + rng() % 600); for (auto& b : icc) b = (uint8_t)rng(); jp.icc = icc.data(); jp.icc_len = icc.size(); }
            jp.alpha_comp = (p.ncomp == 2 || p.ncomp == 4) && (rng() & 1) ? (int)p.ncomp - 1 : -1;
            uint8_t* boxed = nullptr; size_t bn = 0;
            if (j2kcuda_wrap_jp2(out, n, &jp, &boxed, &bn) != 0) { printf("wrap_jp2 refused: %s\n", j2kcuda_last_error()); return 2; }
            j2kcuda_jp2_info ji;
            if (j2kcuda_probe_jp2(boxed, bn, &ji) != 0 || !ji.is_jp2) { printf("probe_jp2 refused the wrapper's file: %s\n", j2kcuda_last_error()); return 2; }
            if (ji.width != p.width || ji.height != p.height || ji.ncomp != p.ncomp || ji.codestream_length != n ||
                ji.codestream_offset + n > bn || memcmp(boxed + ji.codestream_offset, out, n)) { printf("jp2: header or codestream differs\n"); return 2; }
            for (unsigned c = 0; c < p.ncomp; ++c)
                if (ji.bpcc[c] != ((jp.prec[c] - 1) | (jp.sgnd[c] << 7))) { printf("jp2: bpcc of component %u differs\n", c); return 2; }
            if (jp.icc) { if (ji.colr_method != 2 || ji.icc_length != jp.icc_len || memcmp(boxed + ji.icc_offset, jp.icc, jp.icc_len)) { printf("jp2: ICC profile differs\n"); return 2; } }
            else if (ji.colr_method != 1 || ji.enumcs != jp.enumcs) { printf("jp2: colour specification differs\n"); return 2; }
            if (jp.alpha_comp >= 0 && !jp.icc) {      // (with an ICC profile the reference writes no cdef either: jp2.cpp:2173-2181)
                bool found = false;
                for (unsigned i = 0; i < ji.cdef_n; ++i) if (ji.cdef_cn[i] == (unsigned)jp.alpha_comp && ji.cdef_typ[i] == 1) found = true;
                if (!found) { printf("jp2: opacity channel not declared\n"); return 2; }
            }
            // and through the decoder's own locator
            size_t off = 0, len = 0; bool isjp2 = false;
            if (!find_codestream(boxed, bn, off, len, isjp2) || !isjp2 || off != ji.codestream_offset || len != n) { printf("find_codestream disagrees with probe_jp2\n"); return 2; }
            j2kcuda_free(boxed);
        }
        j2kcuda_free(out);
    }
    printf("%ld iterations, %ld written and read back, %ld refused\n", iters, written, refused);
    fflush(stdout);
    j2k::out_release_all();          // the recycled output buffers (what j2kcuda_shutdown does)
    return written > 0 ? 0 : 3;
}

int main(int argc, char** argv) {
    if (argc > 1 && !strcmp(argv[1], "--roundtrip")) return roundtrip(argc > 2 ? atof(argv[2]) : 10, argc > 3 ? atoi(argv[3]) : 1);
    const char* dir = argc > 1 ? argv[1] : ".";
    double seconds = argc > 2 ? atof(argv[2]) : 60;
    unsigned seed0 = argc > 3 ? atoi(argv[3]) : 1;
    DIR* d = opendir(dir); dirent* e;
    while ((e = readdir(d))) { if (e->d_name[0] == '.') continue; std::string p = std::string(dir) + "/" + e->d_name; FILE* f = fopen(p.c_str(), "rb"); std::vector<uint8_t> b(1 << 20); b.resize(fread(b.data(), 1, b.size(), f)); fclose(f); seeds.push_back(b); }
    closedir(d);
    std::mt19937 rng(seed0);
    auto t0 = std::chrono::steady_clock::now();
    long iters = 0, ok = 0;
    DecoderHost reuse;                                           // the geometry-reuse path across frames
    while (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() < seconds) {
        std::vector<uint8_t> b = seeds[rng() % seeds.size()];
        const size_t sod = [&] { for (size_t i = 0; i + 1 < b.size(); ++i) if (b[i] == 0xff && b[i + 1] == 0x93) return i; return b.size() / 2; }();
        int nm = 1 + rng() % 6;
        for (int m = 0; m < nm; ++m) {
            const int kind = rng() % 8;
            // bias towards the headers: they drive every allocation
            size_t pos = (rng() % 3) ? rng() % (sod + 64 < b.size() ? sod + 64 : b.size()) : rng() % b.size();
            if (kind == 0) b[pos] ^= 1u << (rng() % 8);
            else if (kind == 1) b[pos] = (uint8_t)rng();
            else if (kind == 2) b[pos] = (rng() & 1) ? 0xff : 0x00;
            else if (kind == 3) b.resize(pos + 1);
            else if (kind == 4 && pos + 4 < b.size()) { uint32_t v = rng(); memcpy(&b[pos], &v, 4); }
            else if (kind == 5 && pos + 2 < b.size()) { b[pos] = 0xff; b[pos + 1] = (uint8_t)(0x4f + rng() % 0x50); }
            else if (kind == 6) b.insert(b.begin() + pos, (size_t)(rng() % 16), (uint8_t)rng());
            else if (pos + 8 < b.size()) b.erase(b.begin() + pos, b.begin() + pos + 1 + rng() % 8);
        }
        std::string err;
        { DecoderHost h; if (h.parse(b.data(), b.size(), err)) ++ok; }
        reuse.parse(b.data(), b.size(), err);
        { size_t off = 0, len = 0; bool jp2 = false;          // what j2kcuda_probe / j2kcuda_t2_parse do (decoder.cpp), minus the CUDA file
          if (find_codestream(b.data(), b.size(), off, len, jp2)) {
              DecoderHost h; size_t pos = 0; h.parse_main_header(b.data() + off, len, pos, err);
              DecoderHost g;
              if (g.parse(b.data() + off, len, err)) {
                  size_t total = 0; for (auto& t : g.tiles) for (auto& bl : t.blocks) total += (size_t)bl.len;
                  std::vector<uint8_t> data(total + 1); size_t p = 0;
                  for (auto& t : g.tiles) for (auto& bl : t.blocks) for (auto& ch : bl.chunks) { memcpy(data.data() + p, b.data() + off + ch.first, ch.second); p += ch.second; }
              }
          } }
        j2kcuda_jp2_info ji; j2kcuda_probe_jp2(b.data(), b.size(), &ji);
        uint32_t pal[1024]; j2kcuda_jp2_palette(b.data(), b.size(), pal, 1024);
        ++iters;
    }
    printf("%ld iterations, %ld parsed ok\n", iters, ok);
    fflush(stdout);
}
