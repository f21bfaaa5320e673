// Device-side data layout and kernel launch wrappers of libj2kcuda (sm_100a only).
//
// HBM layout.  Every component of a frame owns two full-image int32 planes "A" and "B" with a
// common row pitch (multiple of 4 samples, rows 16-byte aligned).  Tiles are sub-rectangles of
// those planes.  The forward DWT ping-pongs: level 1 reads A and writes all four sub-bands in
// Mallat position into B; level 2 reads B's LL corner and writes into A's corner, and so on.
// The sub-bands of decomposition level l therefore live in B for odd l and in A for even l (the
// final LL lives in the buffer the last level wrote), always at the coordinates the reference's
// single in-place tile buffer would use (tile_buf.cpp:147-157, t1.cpp:1492-1501), so every
// level moves exactly 4 B in + 4 B out per sample and no copy-back is needed: the Tier-1 block
// table simply points each code-block at the plane that holds its band.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "copybuffer.cuh"

namespace j2k {

// ---- multi-component transform + sample (un)packing ------------------------------------------
struct PixelSrc {              // caller samples (host layout mirrored on the device)
    const uint8_t* base[4];    // device pointer of sample (0,0) of each component
    int64_t colbytes[4];
    int64_t rowbytes[4];
    int sample_bytes;          // 1, 2 or 4
    int sgnd;
};
struct MctParams {
    int width, height;         // image size
    int pitch;                 // plane pitch in samples
    int ncomp;
    int mct;                   // apply RCT/ICT to components 0..2
    int irreversible;
    int dc_shift[4];           // 1 << (prec-1) for unsigned components
    int prec[4];
    CopyConv conv[4];          // caller sample -> component sample (CopyBuffer, j2k_codec.cpp:222-427), forward only
};
void launch_mct_forward(const PixelSrc& src, const MctParams& mp, int32_t* const planes[4], cudaStream_t st);
// inverse: planes (int32, or float bit patterns when irreversible) -> caller samples
struct PixelDst {
    uint8_t* base[4];
    int64_t colbytes[4];
    int64_t rowbytes[4];
    int sample_bytes;
    int sgnd;
};
void launch_mct_inverse(const PixelDst& dst, const MctParams& mp, const int32_t* const planes[4], cudaStream_t st);
// decode side of CopyBuffer for a destination that is not the tight image the inverse MCT kernel writes: component c of
// the tight interleaved image `img` (S-byte samples, signed when sgnd) -> a tight w x h plane of the caller's sample type
// (sx, sy: read every sx-th / sy-th sample of the image: a destination at a sub-sampled component's own resolution)
void launch_copy_channel(const uint8_t* img, int S, int ncomp, int c, int sgnd, int img_w, int w, int h, int sx, int sy, uint8_t* dst,
                         const CopyConv& conv, cudaStream_t st);
// Output stage when some components are sub-sampled (no MCT then): level shift, clamp, and every component replicated to the
// reference grid -- output sample (x, y) of component c = its plane at (x / dx, y / dy), the rule of Codec::CopyBuffer
// (src/common/j2k_codec.cpp:262-279: the source advances once every relativeSub destination samples).
struct SubOut {
    const int32_t* plane[4];
    int dx[4], dy[4], cw[4], ch[4], prec[4], dc_shift[4];
    int width, height, pitch, ncomp, sample_bytes, irreversible, sgnd;
};
void launch_output_subsampled(const SubOut& so, uint8_t* d_out, cudaStream_t st);
// bare three-plane transforms used by the stage entry points (mct.cpp:94-471)
void launch_mct_planes(int32_t* c0, int32_t* c1, int32_t* c2, size_t n, int irreversible, int inverse, cudaStream_t st);

// ---- DWT -------------------------------------------------------------------------------------
// One (tile-component, level) job.  Coordinates are canvas coordinates of the level's INPUT
// region for the forward transform / OUTPUT region for the inverse transform.
struct DwtJob {
    int32_t* src;     // forward: interleaved region origin; inverse: Mallat region origin
    int32_t* dst;     // forward: Mallat region origin;      inverse: interleaved region origin
    int x0, y0, x1, y1;
    int pitch;        // samples, same for src and dst
    int strip0;       // row pairs per segment chosen by dwt_plan_jobs
    int nstrips, nsegs;
    int strip_cols;   // columns per warp strip (multiple of the columns one lane holds), set by dwt_plan_jobs
    int pad;
};
enum { DWT_53 = 0, DWT_97_FIX = 1, DWT_97_FLOAT = 2 };
// Level 1 of the forward transform fused with sample unpacking, DC level shift and RCT / ICT: one job per tile, all components
// at once, straight from the caller's interleaved pixels (dwt.cu).  Used when the source is an interleaved 1- or 2-byte image
// with 3 or 4 channels per pixel whose samples need no depth / sign conversion.
struct DwtPxSrc {             // per frame: where the pixels are
    const uint8_t* px;        // sample (0, 0) of channel 0 of the interleaved image on the device
    long long rowbytes;
    int chan[4];              // channel of component c inside a pixel
};
struct DwtPxJob {             // per context: fixed geometry
    int32_t* dst[4];          // Mallat region origin of every component
    int x0, y0, x1, y1, pitch;
    int strip0, nstrips, nsegs, strip_cols;
    int dc[4];
    int mct, irreversible;
};
bool dwt_px_fusion_available();
bool launch_dwt_forward_px(const DwtPxJob* d_jobs, const DwtPxSrc& src, int njobs, int max_items, int kind, int ncomp, int nch, int sample_bytes,
                           cudaStream_t st);
// interleave detection shared with the MCT kernels: channels per pixel (0 = not one interleaved array), chan[c] = channel of component c
int detect_pixel_interleave(const PixelSrc& s, int ncomp, int chan[4]);
// host-side helper: fills strip/segment counts (jobs[i].strip0 = row pairs per segment, the same
// for all jobs of the launch); returns the largest number of warp work items of any job
int dwt_plan_jobs(DwtJob* jobs, int njobs);
void launch_dwt_forward(const DwtJob* d_jobs, int njobs, int max_items, int seg_pairs, int kind, cudaStream_t st);
void launch_dwt_inverse(const DwtJob* d_jobs, int njobs, int max_items, int seg_pairs, int kind, cudaStream_t st);

// ---- Tier-1 ----------------------------------------------------------------------------------
struct T1Block {               // one code-block, 48 bytes
    const int32_t* coeff;      // top-left coefficient of the block inside its plane
    int pitch;
    uint16_t w, h;
    uint8_t orient;            // 0 LL, 1 HL, 2 LH, 3 HH
    uint8_t qmfbid;            // 1: 5/3 (v << 6), 0: 9/7 (fix_mul_t1 by bandconst)
    uint16_t max_planes;       // band numbps: upper bound on coded planes (sizes the stream slot)
    int32_t bandconst;
    uint32_t sym_off;          // 16-byte units into the symbol arena
    uint32_t sym_cap;          // bytes
    uint32_t out_off;          // bytes into the code-byte arena
    uint32_t out_cap;
    uint32_t pad;
};
constexpr int T1_MAX_PASSES = 100;         // tcd.h:86-91 keeps 100 passes per block
struct T1Result {              // per block, copied back to the host (plugin_bridge.cpp:72-174)
    int32_t numbps;
    int32_t npasses;
    int32_t total;             // bytes kept (rate of the last pass)
    int32_t status;            // 0 ok, 1 symbol slot overflow, 2 byte slot overflow
    uint32_t nsym;             // MQ decisions
    uint32_t pad[3];
    int32_t rate[T1_MAX_PASSES];      // cumulative, clamped and 0xFF-trimmed (t1_opt.cpp:722-736)
    int32_t nmsedec[T1_MAX_PASSES];   // per pass (t1.cpp:298-314)
};
void launch_t1_context(const T1Block* d_blocks, int nblocks, uint8_t* sym_arena, uint32_t* pass_nsym,
                       T1Result* results, cudaStream_t st);
void launch_t1_mq(const T1Block* d_blocks, int nblocks, const uint8_t* sym_arena,
                  const uint32_t* pass_nsym, uint8_t* out_arena, T1Result* results,
                  int lanes_per_warp, uint32_t* d_order, int max_long, cudaStream_t st, int cblksty = 0);
// (d_order: scratch of nblocks + 2 words, or null: filled with the block indices by decreasing symbol count;
//  the blocks at least half as long as the longest, at most max_long of them, get a warp each)
// gather the kept bytes of every block into one contiguous buffer (offsets = exclusive scan of
// T1Result::total, written to d_offsets[nblocks+1])
void launch_t1_compact(const T1Block* d_blocks, int nblocks, const uint8_t* out_arena,
                       const T1Result* results, uint32_t* d_offsets, uint8_t* packed, cudaStream_t st);

struct T1DecBlock {            // decode job, one code-block
    int32_t* coeff;            // destination inside the plane (int32, or float when 9/7)
    int pitch;
    uint16_t w, h;
    uint8_t orient, qmfbid;
    uint8_t numbps;            // band numbps - zero bit-planes  (t2.cpp:1230-1240)
    uint8_t npasses;
    uint32_t data_off, data_len;   // into the segment arena
    float stepsize;            // 9/7 dequantiser step, already halved (tcd.cpp:888)
    uint32_t seg_index;        // first entry of the block's codeword segments in the segment table (when nsegs > 1)
    uint16_t nsegs;            // codeword segments (1 unless the code-block style terminates passes: TERMALL, LAZY)
    uint8_t cblksty;           // code-block style bits of the component (j2k.h:77-82)
    uint8_t roishift;          // region-of-interest up-shift of the component (RGN marker)
    uint32_t pad;
};
struct T1DecSeg { uint32_t off, len, passes, pad; };   // one codeword segment: bytes at off in the segment arena (8-byte aligned, 0xFF 0xFF after)
// generic = 1: some block uses a code-block style, several segments or an ROI shift (slower kernel variant)
void launch_t1_decode(const T1DecBlock* d_blocks, int nblocks, const uint8_t* seg_arena, const T1DecSeg* d_segs, int generic,
                      int raw_output, cudaStream_t st);

const char* device_tables_init();   // uploads LUTs once per device; returns error text or null

}  // namespace j2k
