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
int main(int argc, char** argv) {
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
}
