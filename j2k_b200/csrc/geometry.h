// Tile geometry and coding parameters shared by the encoder, the decoder and the kernels' block
// tables.  Restates what opj_tcd_init_tile computes (ext/grok/src/lib/openjp2/tcd.cpp:458-872)
// as flat vectors instead of nested allocations.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>
#include <string>

namespace j2k {

constexpr int MAX_COMPS = 4;
constexpr int MAX_RES = 33;

struct StepSize { int expn = 0, mant = 0; };

struct CompParams {           // opj_tccp_t + opj_image_comp_t, j2k.h:173-203
    int prec = 8, sgnd = 0, dx = 1, dy = 1;
    int numres = 6;
    int cblkw = 6, cblkh = 6;          // log2
    int cblksty = 0;
    int qmfbid = 1;                    // 1 = 5/3, 0 = 9/7
    int qntsty = 0;                    // 0 none, 1 derived, 2 expounded
    int numgbits = 2;
    int csty = 0;                      // bit0: user precincts
    int prcw[MAX_RES], prch[MAX_RES];  // log2 precinct sizes
    int roishift = 0;
    StepSize steps[3 * MAX_RES];
    int dc_shift = 0;
};

struct PocEntry { int resno0, compno0, layno1, resno1, compno1, prg; };

struct CodingParams {         // opj_cp_t + opj_tcp_t subset
    uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0;       // image area on the canvas
    uint32_t tx0 = 0, ty0 = 0, tdx = 0, tdy = 0;   // tile grid
    uint32_t tw = 1, th = 1;
    int ncomp = 0;
    int rsiz = 0;
    int csty = 0;              // Scod: bit0 precincts, bit1 SOP, bit2 EPH
    int prog = 0;
    int numlayers = 1;
    int mct = 0;
    CompParams comps[MAX_COMPS];
    std::vector<PocEntry> pocs;
    // encoder only
    double rates[16] = {0};    // per layer; converted to bytes by update_rates
    uint64_t max_cs_size = 0;
    uint32_t max_comp_size = 0;
    int compat = 0;
    uint32_t tile_first = 0, tile_count = 0;   // encode only these tiles, emit tile-parts alone (0 = all, whole codestream)
    bool tile_in_range(int t) const { return tile_count == 0 || ((uint32_t)t >= tile_first && (uint32_t)t < tile_first + tile_count); }
    bool tp_on = false;        // tile parts split by component (cinema)
    int tp_flag = 0;
    std::string comment;
};

struct Rect { int x0 = 0, y0 = 0, x1 = 0, y1 = 0; int w() const { return x1 - x0; } int h() const { return y1 - y0; } };

struct CblkGeo {
    Rect r;                 // in band coordinates
    int comp, res, band;    // band = index within resolution (0..2)
    int orient;             // bandno: 0 LL, 1 HL, 2 LH, 3 HH
    int prec;               // precinct index within resolution
    int px, py;             // position inside the tile-component plane (Mallat layout)
};

struct PrecGeo {
    Rect r;
    int cw = 0, ch = 0;
    int first_cblk = 0;     // index into TileGeo::cblks
};

struct BandGeo {
    Rect r;
    int orient = 0;
    float stepsize = 1.f;
    int numbps = 0;
    int offx = 0, offy = 0; // Mallat offset of the band inside the plane
    std::vector<PrecGeo> precs;
};

struct ResGeo {
    Rect r;
    int pw = 0, ph = 0, pdx = 15, pdy = 15;
    int numbands = 1;
    BandGeo bands[3];
};

struct CompGeo {
    Rect r;                 // tile-component bounds
    int numres = 0;
    std::vector<ResGeo> res;
};

struct TileGeo {
    int tileno = 0;
    Rect r;
    std::vector<CompGeo> comps;
    std::vector<CblkGeo> cblks;   // comp -> res -> band -> precinct -> cblk (raster)
};

inline uint32_t ceildiv(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a + b - 1) / b); }
inline uint32_t ceildivpow2(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a + ((uint64_t)1 << b) - 1) >> b); }
inline int64_t ceildivpow2_64(int64_t a, int b) { return (a + ((int64_t)1 << b) - 1) >> b; }
inline uint32_t floordivpow2(uint32_t a, uint32_t b) { return a >> b; }
inline int floorlog2(uint32_t a) { int l = 0; while (a > 1) { a >>= 1; ++l; } return l; }

// fraction = 1.0 for the encoder, 0.5 for the decoder (tcd.cpp:876,888)
void build_tile(const CodingParams& cp, int tileno, float fraction, TileGeo& out);
bool geometry_within_limits(const CodingParams& cp, std::string& err);   // call before build_tile on untrusted parameters

// dwt.cpp:193-208, 531-593; mct.cpp:73,78
double dwt_norm(int level, int orient, bool real);
double mct_norm(int compno, bool real);
int dwt_gain(int orient, bool real);
void calc_explicit_stepsizes(CompParams& c);
// t1.cpp:1482: Q13 reciprocal of the band step used by the 9/7 quantiser
inline int32_t band_const(float stepsize) {
    int32_t d = (int32_t)std::floor(stepsize * 8192);
    return d ? 8192 * 8192 / d : 0;
}

}  // namespace j2k
