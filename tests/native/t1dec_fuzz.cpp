// TEST INFRASTRUCTURE.  The Tier-1 decoder's block routine (j2k_b200/csrc/t1_dec_core.cuh, the code k_t1_decode runs) on the host
// under AddressSanitizer, fed random and damaged code-block bytes in an arena laid out EXACTLY as decoder.cpp lays it out on the
// device: the segment 8-byte aligned, two 0xFF bytes behind it, the next segment (or the arena's slack) after that.  The MQ
// reader fetches 8-byte words ahead of the byte it decodes (t1_dec_core.cuh, MqIo); the contract in that header is "two 0xFF
// and at least 24 more readable bytes behind every segment".  Here every segment sits in a heap block of exactly that size, so
// a read beyond the contract is an ASan report.  Whatever the bytes, the routine must also terminate and stay inside `out`.
//     tests/native/t1dec_fuzz <seconds> <rng seed>
#include "t1dec_host.cpp"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
namespace j2k { const T1Tables& t1_tables_host(); }

int main(int argc, char** argv) {
    const double seconds = argc > 1 ? atof(argv[1]) : 10;
    std::mt19937 rng(argc > 2 ? atoi(argv[2]) : 1);
    const j2k::T1Tables& T = j2k::t1_tables_host();
    uint32_t tab[96];
    for (int i = 0; i < 94; ++i) tab[i] = mq_state_word(T.mq, i);
    uint8_t sc2[256];
    for (int i = 0; i < 256; ++i) sc2[i] = sc_entry(T.sc, i);
    auto t0 = std::chrono::steady_clock::now();
    long iters = 0;
    static DRows R; static ColBuf cb; static MqIo io;
    while (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() < seconds) {
        const int w = 1 + rng() % 64, h = 1 + rng() % 64, orient = rng() % 4;
        const int numbps = 1 + rng() % 30, maxp = 3 * numbps - 2;
        const int npasses = 1 + rng() % maxp;
        const bool generic = rng() % 3 == 0;
        const int sty = generic ? (int)(rng() % 64) : 0, roishift = generic && (rng() % 4 == 0) ? (int)(rng() % (31 - numbps)) : 0;
        // segments: one, or several with pass counts that add up
        int nsegs = 1;
        std::vector<int> seg_passes{npasses};
        if (generic && npasses > 1 && (sty & 5)) {
            seg_passes.clear();
            int left = npasses;
            while (left > 0) { int take = (sty & 4) ? 1 : 1 + (int)(rng() % (left < 10 ? left : 10)); if (take > left) take = left; seg_passes.push_back(take); left -= take; }
            nsegs = (int)seg_passes.size();
        }
        std::vector<T1DecSeg> segs;
        size_t total = 0;
        std::vector<int> lens(nsegs);
        for (int i = 0; i < nsegs; ++i) { lens[i] = (rng() % 8 == 0) ? 0 : (int)(rng() % (rng() % 4 == 0 ? 6000 : 200)); total += (((size_t)lens[i] + 2) + 7) & ~(size_t)7; }
        // exactly what the contract promises behind the LAST segment: 24 readable bytes (the others are followed by their successors)
        uint8_t* arena = (uint8_t*)aligned_alloc(8, total + 24);
        const int fill = rng() % 4;
        for (size_t i = 0; i < total + 24; ++i) arena[i] = fill == 0 ? 0xff : fill == 1 ? 0x00 : (uint8_t)rng();
        size_t pos = 0;
        for (int i = 0; i < nsegs; ++i) {
            arena[pos + lens[i]] = 0xff; arena[pos + lens[i] + 1] = 0xff;
            segs.push_back(T1DecSeg{(uint32_t)pos, (uint32_t)lens[i], (uint32_t)seg_passes[i], 0});
            pos += (((size_t)lens[i] + 2) + 7) & ~(size_t)7;
        }
        int32_t* out = (int32_t*)malloc(sizeof(int32_t) * (size_t)w * h);
        memset(out, 0, sizeof(int32_t) * (size_t)w * h);
        T1DecBlock b;
        memset(&b, 0, sizeof(b));
        b.coeff = out; b.pitch = w; b.w = (uint16_t)w; b.h = (uint16_t)h; b.orient = (uint8_t)orient; b.qmfbid = (uint8_t)(rng() & 1);
        b.numbps = (uint8_t)numbps; b.npasses = (uint8_t)npasses; b.stepsize = 0.5f;
        b.data_off = segs[0].off; b.data_len = segs[0].len; b.nsegs = (uint16_t)nsegs; b.seg_index = 0;
        b.cblksty = (uint8_t)sty; b.roishift = (uint8_t)roishift;
        memset(&R, 0, sizeof(R));
        uint32_t st[32];
        for (int i = 0; i < 32; ++i) st[i] = tab[0];
        st[CTX_UNI] = tab[46 * 2]; st[CTX_AGG] = tab[3 * 2]; st[CTX_ZC] = tab[4 * 2];
        if (generic) decode_block<true>(b, R, cb, &io, st, T.zc9, sc2, tab, arena, segs.data(), rng() & 1, 0, 1);
        else decode_block<false>(b, R, cb, &io, st, T.zc9, sc2, tab, arena, nullptr, rng() & 1, 0, 1);
        free(out); free(arena);
        ++iters;
    }
    printf("%ld blocks decoded\n", iters);
    return 0;
}
