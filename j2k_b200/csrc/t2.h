// Host Tier-2: bit I/O, tag trees, progression order, packet headers, layer formation, markers.
// Consumes the per-code-block results copied back from the device (encoder) and produces the
// per-code-block segments uploaded to the device (decoder).
// Reference behaviour: ext/grok/src/lib/openjp2/{bio,tgt,pi,t2,tcd,j2k}.cpp (cited per function).
#pragma once
#include "geometry.h"
#include "common.h"
#include "../../include/j2kcuda.h"
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

namespace j2k {

// bio.cpp:99-235.  A byte that follows 0xFF carries 7 bits.
struct BitWriter {
    ByteBuf* out = nullptr;                // null: count only (simulate)
    size_t n = 0;
    uint32_t buf = 0;
    int ct = 8;
    void byteout() {
        ct = buf == 0xff ? 7 : 8;
        if (out) out->push_back((uint8_t)buf);
        ++n;
        buf = 0;
    }
    void putbit(uint32_t b) {
        if (ct == 0) byteout();
        --ct;
        buf |= b << ct;
    }
    void write(uint32_t v, int nb) { for (int i = nb - 1; i >= 0; --i) putbit((v >> i) & 1); }
    void flush() { byteout(); if (ct == 7) byteout(); }
};

struct BitReader {
    const uint8_t* bp = nullptr; const uint8_t* end = nullptr;
    uint32_t buf = 0; int ct = 0; bool overrun = false;
    void init(const uint8_t* p, size_t len) { bp = p; end = p + len; buf = 0; ct = 0; overrun = false; }
    void bytein() {
        ct = buf == 0xff ? 7 : 8;
        if (bp >= end) { overrun = true; buf = 0; return; }
        buf = *bp++;
    }
    uint32_t getbit() { if (ct == 0) bytein(); --ct; return (buf >> ct) & 1; }
    uint32_t read(int nb) { uint32_t v = 0; for (int i = nb - 1; i >= 0; --i) v |= getbit() << i; return v; }
    void inalign() { if ((buf & 0xff) == 0xff) bytein(); ct = 0; }
};

// tgt.cpp:66-352
struct TagTree {
    struct Node { int parent; int value; int low; int known; };
    std::vector<Node> nodes;
    int w = 0, h = 0;
    void init(int numleafsh, int numleafsv);
    void reset();
    void setvalue(int leaf, int value);
    void encode(BitWriter& bio, int leaf, int threshold);
    int decode(BitReader& bio, int leaf, int threshold);
};

struct PacketId { int layno, resno, compno, precno; };

// pi.cpp:276-567: packets of one progression-order change, in order.
void progression_packets(const CodingParams& cp, const TileGeo& t, int prg, int resno0, int resno1,
                         int compno0, int compno1, int layno0, int layno1,
                         std::vector<uint8_t>& include, std::vector<PacketId>& out);

// ---- encoder side ---------------------------------------------------------------------------
struct EncLayer { int numpasses = 0; int len = 0; int data_off = 0; double disto = 0; };
struct EncBlock {
    const uint8_t* data = nullptr;
    int numbps = 0, totalpasses = 0;
    const int32_t* rates = nullptr;      // cumulative, clamped + trimmed
    const double* dist = nullptr;        // cumulative
    int numpasses = 0;                   // included so far while writing packets
    int numlenbits = 0;
    int included_other = 0;              // num_passes_included_in_other_layers
    std::vector<EncLayer> layers;
};

struct EncTile {
    TileGeo geo;
    std::vector<EncBlock> blocks;        // same order as geo.cblks
    std::vector<TagTree> incl, imsb;     // per (comp,res,band,prec), indexed by prec_index()
    std::vector<int> prec_base;          // [comp][res] -> first tree index
    double distotile = 0;
    int packno = 0;
};

struct EncoderHost {
    CodingParams cp;
    int ntiles = 0;
    // codestream assembly: tile results -> bytes
    bool assemble(std::vector<EncTile>& tiles, ByteBuf& out, std::string& err);
};

const char* check_params(const ::j2kcuda_params& p);                      // nullptr = acceptable (j2k.cpp:6640-6680)
void setup_coding_params(const ::j2kcuda_params& p, CodingParams& cp);   // j2k.cpp:5842-6216
void init_enc_tile(const CodingParams& cp, int tileno, EncTile& t);
bool join_tile_streams(const CodingParams& cp, const uint8_t* const* frags, const size_t* lens, int nfrags,
                       ByteBuf& out, std::string& err);

// ---- decoder side ---------------------------------------------------------------------------
struct DecSeg { int numpasses = 0, maxpasses = 109, len = 0; };   // opj_tcd_seg_t, t2.cpp:1525-1563
struct DecBlock {
    std::vector<DecSeg> segs;  // codeword segments (one unless the code-block style terminates passes)
    int numbps = 0;
    int npasses = 0;          // total passes over all layers read
    int numlenbits = 3;
    bool included = false;
    std::vector<std::pair<uint32_t, uint32_t>> chunks;  // (offset in codestream, length) per layer
    int len = 0;
};
struct DecTile {
    TileGeo geo;
    std::vector<DecBlock> blocks;
    std::vector<TagTree> incl, imsb;
    std::vector<int> prec_base;
    std::vector<std::pair<size_t, size_t>> parts;   // (offset, length) of tile-part bodies
};
struct DecoderHost {
    CodingParams cp;
    std::vector<DecTile> tiles;
    uint32_t tile_first = 0, tile_count = 0;      // read packets of these tiles only (0 = all): tiles of one frame over several GPUs
    bool tile_in_range(int t) const { return tile_count == 0 || ((uint32_t)t >= tile_first && (uint32_t)t < tile_first + tile_count); }
    // geometry of the previous codestream parsed by this object: frames of a sequence share it, so the tile / precinct /
    // code-block tables and the tag trees are built once and only reset per frame
    CodingParams prev_cp; bool have_prev = false; uint32_t prev_first = 0, prev_count = 0;
    bool parse_main_header(const uint8_t* d, size_t n, size_t& pos, std::string& err);
    bool parse(const uint8_t* d, size_t n, std::string& err);   // headers + packets for all tiles
};
bool find_codestream(const uint8_t* d, size_t n, size_t& off, size_t& len, bool& is_jp2);

}  // namespace j2k
