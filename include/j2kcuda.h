/*
 * j2kcuda.h -- C ABI of libj2kcuda, the B200-native JPEG 2000 codec core.
 *
 * Drop-in boundary: everything the reference's plug-in layer reaches through the abstract
 * j2k::Codec (src/common/j2k_codec.h:279-326) and that its only registered implementation,
 * OpenJPEGCodec (src/common/j2k_openjpeg_codec.cpp:213-758), forwards to the opj_* C API
 * (ext/grok/src/lib/openjp2/openjpeg.h).  Plain C types only: no exceptions, no torch types,
 * caller-owned pixel memory, library-owned output freed by j2kcuda_free().
 *
 * There is NO CPU fallback: every pixel-path entry point returns J2KCUDA_ERR_NO_DEVICE when no
 * sm_100 device is usable.  Tier-2 (packets, PCRD, markers) runs on the host by design, as in
 * the reference's accelerator hand-off (ext/grok/src/lib/openjp2/plugin_bridge.cpp:72-174).
 */
#ifndef J2KCUDA_H
#define J2KCUDA_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define J2KCUDA_VERSION 0x000100

enum {
    J2KCUDA_OK = 0,
    J2KCUDA_ERR_PARAM = 1,      /* bad argument (what OpenJPEGCodec would throw Exception for) */
    J2KCUDA_ERR_NO_DEVICE = 2,  /* no CUDA device / kernels unavailable: never falls back */
    J2KCUDA_ERR_CUDA = 3,       /* a CUDA call failed; see j2kcuda_last_error() */
    J2KCUDA_ERR_STREAM = 4,     /* malformed / unsupported codestream */
    J2KCUDA_ERR_NOMEM = 5,
    J2KCUDA_ERR_ABORTED = 6,    /* abort callback returned non-zero */
    J2KCUDA_ERR_UNSUPPORTED = 7
};

/* j2k::SampleType, src/common/j2k_codec.h:212-219 */
enum { J2KCUDA_UCHAR = 0, J2KCUDA_USHORT = 1, J2KCUDA_UINT = 2, J2KCUDA_INT = 3 };

/* One channel of pixels in caller memory; same fields as j2k::Channel
 * (src/common/j2k_codec.h:221-247).  From After Effects the four channels alias one interleaved
 * ARGB world (buf = base + i*sample, colbytes = 4*sample; src/aftereffects/j2k.cpp:324-362). */
typedef struct j2kcuda_plane {
    void*    buf;
    int64_t  colbytes;     /* bytes between horizontally adjacent samples */
    int64_t  rowbytes;     /* bytes between rows */
    uint32_t width, height;
    uint8_t  sample_type;  /* J2KCUDA_UCHAR ... */
    uint8_t  depth;        /* significant bits in the caller's samples */
    uint8_t  sgnd;
    uint8_t  reserved;
} j2kcuda_plane;

/* OPJ_PROG_ORDER, ext/grok/src/lib/openjp2/openjpeg.h:303-310 */
enum { J2KCUDA_LRCP = 0, J2KCUDA_RLCP = 1, J2KCUDA_RPCL = 2, J2KCUDA_PCRL = 3, J2KCUDA_CPRL = 4 };
/* rsiz profiles, openjpeg.h:225-226 */
enum { J2KCUDA_PROFILE_NONE = 0, J2KCUDA_PROFILE_CINEMA_2K = 3, J2KCUDA_PROFILE_CINEMA_4K = 4 };
/* Which encoder's Tier-2 policy to reproduce byte for byte. */
enum { J2KCUDA_COMPAT_GROK_2_1_1 = 0, J2KCUDA_COMPAT_OPENJPEG_2_5 = 1 /* COM text of 2.5.4 (Pillow's) */,
       J2KCUDA_COMPAT_OPENJPEG_2_5_3 = 2 /* same policy, COM text of 2.5.3 (OpenCV's) */ };

/* Encoder parameters: the subset of opj_cparameters_t (openjpeg.h:356-486) the plug-in and the
 * BASELINE configs use; zero means "library default" (openjpeg.cpp:706-728). */
typedef struct j2kcuda_params {
    uint32_t width, height;      /* image size (x0 = y0 = 0 as the plug-in always sets) */
    uint32_t ncomp;              /* 1..4 */
    uint32_t prec;               /* bits per sample in the codestream, 1..16 */
    uint32_t sgnd;
    uint32_t irreversible;       /* 0: 5/3 + RCT, 1: 9/7 + ICT */
    uint32_t numresolution;      /* 0 -> 6 */
    uint32_t cblk_w, cblk_h;     /* 0 -> 64 */
    uint32_t mct;                /* tcp_mct: 0 none, 1 RCT/ICT on components 0..2 */
    uint32_t prog_order;         /* J2KCUDA_LRCP ... */
    uint32_t tile_w, tile_h;     /* 0 -> single tile */
    uint32_t numlayers;          /* 0 -> 1 */
    uint32_t profile;            /* J2KCUDA_PROFILE_* */
    uint32_t max_comp_size;      /* cinema: bytes per component per frame; 0 -> profile default */
    uint64_t max_cs_size;        /* cinema: bytes per frame; 0 -> profile default */
    double   rates[16];          /* tcp_rates: compression ratio per layer, 0 = lossless */
    uint32_t compat;             /* J2KCUDA_COMPAT_* */
    int32_t  device;             /* CUDA ordinal, -1 = current */
    /* Tiles of one very large frame sharded across GPUs (config C4): this context encodes only tiles
     * [tile_first, tile_first + tile_count) of the frame and returns their tile-parts alone (no main
     * header, no EOC); j2kcuda_join_tile_streams() puts the fragments of all ranks together.
     * tile_count == 0: the whole frame, a complete codestream. */
    uint32_t tile_first, tile_count;
    /* Code-block style (Scod / SPcod byte, j2k.h:77-82; opj_cparameters_t::mode).  The encoder supports RESET (0x02), TERMALL
     * (0x04), SEGSYM (0x20) and their combinations: styles the arithmetic-coder stage can do alone (t1.cpp:1537-1673, row
     * a21); LAZY, VSC and PTERM are refused.  The decoder reads all six.  0 = none (what the plug-in writes). */
    uint32_t cblksty;
    uint32_t reserved[3];
} j2kcuda_params;

/* What GetFileInfo needs from the main header (j2k_openjpeg_codec.cpp:222-429). */
typedef struct j2kcuda_info {
    uint32_t width, height, ncomp;
    uint32_t prec[4], sgnd[4], dx[4], dy[4];
    uint32_t reversible, numresolution, numlayers, mct, prog_order;
    uint32_t tile_w, tile_h, tiles_x, tiles_y;
    uint32_t cblk_w, cblk_h, profile;
    uint32_t is_jp2;             /* input was a JP2 box file; codestream located inside */
    uint64_t codestream_offset, codestream_length;
} j2kcuda_info;

typedef int  (*j2kcuda_abort_fn)(void* user);                 /* non-zero = stop */
typedef void (*j2kcuda_log_fn)(void* user, int level, const char* msg);

/* ---- lifetime ------------------------------------------------------------------------------ */
int  j2kcuda_init(int device, j2kcuda_log_fn log, void* user);     /* replaces opj_initialize */
void j2kcuda_shutdown(void);
int  j2kcuda_version(void);
int  j2kcuda_device_count(void);
const char* j2kcuda_last_error(void);                               /* thread-local */
void j2kcuda_free(void* p);
void j2kcuda_default_params(j2kcuda_params* p);     /* opj_set_default_encoder_parameters */

/* ---- the reference-facing calls ------------------------------------------------------------ */
/* replaces opj_read_header(_ex) as used by OpenJPEGCodec::GetFileInfo */
int j2kcuda_probe(const uint8_t* bytes, size_t n, j2kcuda_info* out);
/* replaces opj_setup_encoder + opj_start_compress + opj_encode + opj_end_compress
 * (j2k_openjpeg_codec.cpp:722-736).  planes[c] live in HOST memory.  *out is malloc'd. */
int j2kcuda_encode(const j2kcuda_params* p, const j2kcuda_plane* planes, int nplanes,
                   uint8_t** out, size_t* out_len);
/* replaces opj_read_header + opj_decode (j2k_openjpeg_codec.cpp:489-512) and the CopyBuffer that
 * follows (:528-562).  reduce = log2(subsample).  planes[c] are HOST destinations sized
 * ceil(w / 2^reduce) x ceil(h / 2^reduce). */
int j2kcuda_decode(const uint8_t* bytes, size_t n, unsigned reduce, const j2kcuda_plane* planes,
                   int nplanes, j2kcuda_abort_fn abort_fn, void* user);

/* j2kcuda_decode restricted to a window (see j2kcuda_decoder_set_window) */
int j2kcuda_decode_region(const uint8_t* bytes, size_t n, unsigned reduce, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1,
                          const j2kcuda_plane* planes, int nplanes);

/* ---- persistent contexts (image sequences: configs C4/C5; also what bench.py times) -------- */
typedef struct j2kcuda_encoder j2kcuda_encoder;
typedef struct j2kcuda_decoder j2kcuda_decoder;

j2kcuda_encoder* j2kcuda_encoder_create(const j2kcuda_params* p);
void j2kcuda_encoder_destroy(j2kcuda_encoder* e);
/* host planes -> device staging (async on the context's stream) */
int j2kcuda_encoder_upload(j2kcuda_encoder* e, const j2kcuda_plane* planes, int nplanes);
/* pixels already resident on the device: interleaved samples, `ncomp` per pixel,
 * sample_type UCHAR/USHORT, pitch in bytes */
int j2kcuda_encoder_set_device_input(j2kcuda_encoder* e, const void* d_pixels, int64_t pitch_bytes,
                                     int sample_type);
/* device stages (MCT, DWT, T1) for all tiles + D2H of code-block bytes / pass tables */
int j2kcuda_encoder_run(j2kcuda_encoder* e);
/* host Tier-2: PCRD, packets, markers -> codestream (malloc'd) */
int j2kcuda_encoder_finish(j2kcuda_encoder* e, uint8_t** out, size_t* out_len);
/* per-stage device times of the last run, milliseconds (CUDA events on the context's stream):
 * [0] upload  [1] mct  [2] dwt  [3] t1 context  [4] t1 mq  [5] compact + D2H  [6] host tier-2 (wall)
 * [7] dwt level-1 launch only.  Returns number of entries written. */
int j2kcuda_encoder_timings(j2kcuda_encoder* e, float* ms, int n);
/* counters of the last run: [0] kernels launched [1] code-blocks [2] MQ symbols
 * [3] compressed bytes [4] h2d bytes [5] d2h bytes */
int j2kcuda_encoder_counters(j2kcuda_encoder* e, uint64_t* v, int n);

j2kcuda_decoder* j2kcuda_decoder_create(int device);
void j2kcuda_decoder_destroy(j2kcuda_decoder* d);
/* host Tier-2 parse + H2D of code-block segments */
int j2kcuda_decoder_prepare(j2kcuda_decoder* d, const uint8_t* bytes, size_t n, unsigned reduce);
/* decode window, opj_set_decode_area (openjpeg.h:1287-1302; dwt_region.cpp, tile_buf.cpp:233-340): only [x0, x1) x [y0, y1)
 * of the reference grid (full-resolution coordinates) is wanted by the next prepare / run / download: code-blocks that cannot
 * reach it through the synthesis filters are skipped, the planes receive the window (sized ceil of it at `reduce`).
 * x1 == y1 == 0: whole image again. */
int j2kcuda_decoder_set_window(j2kcuda_decoder* d, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1);
/* device stages: T1 decode, dequantise, inverse DWT, inverse MCT, level shift, clamp */
int j2kcuda_decoder_run(j2kcuda_decoder* d);
/* device -> host planes */
int j2kcuda_decoder_download(j2kcuda_decoder* d, const j2kcuda_plane* planes, int nplanes);
/* keep result on device: interleaved u8/u16, `ncomp` samples per pixel */
int j2kcuda_decoder_get_device_output(j2kcuda_decoder* d, void** d_pixels, int64_t* pitch_bytes,
                                      int* sample_type);
/* copy the decoded interleaved image into caller-owned DEVICE memory (rows `pitch_bytes` apart) */
int j2kcuda_decoder_copy_to_device(j2kcuda_decoder* d, void* d_pixels, int64_t pitch_bytes);
/* [0] host tier-2 parse (wall) [1] H2D [2] t1 [3] idwt [4] imct+pack [5] D2H [6] idwt finest level */
int j2kcuda_decoder_timings(j2kcuda_decoder* d, float* ms, int n);
/* [0] kernels launched [1] code-blocks [2] code-blocks skipped by the decode window [3] codestream bytes [4] h2d bytes [5] output bytes */
int j2kcuda_decoder_counters(j2kcuda_decoder* d, uint64_t* v, int n);

/* ---- sequences: frames of an image sequence on this process's device (one process per GPU;
 *      configs C4/C5).  Frames are independent, so a sequence keeps `nworkers` contexts (own CUDA
 *      stream + buffers + host thread each): host Tier-2 of one frame overlaps the kernels of the
 *      others.  What FrameSeq's per-frame loop (src/aftereffects/FrameSeq.cpp) would call once per
 *      render batch instead of once per frame. ------------------------------------------------ */
typedef struct j2kcuda_sequence j2kcuda_sequence;
enum { J2KCUDA_SEQ_DEVICE_PIXELS = 1 };  /* frames[i][0].buf is a DEVICE pointer to interleaved pixels,
                                            rowbytes = pitch; ncomp samples per pixel */
/* p may be NULL for a decode-only sequence; device < 0 = p->device / current; nworkers 0 -> 4 */
j2kcuda_sequence* j2kcuda_sequence_create(const j2kcuda_params* p, int device, int nworkers);
void j2kcuda_sequence_destroy(j2kcuda_sequence* s);
int j2kcuda_sequence_encode(j2kcuda_sequence* s, const j2kcuda_plane* const* frames, int nframes,
                            int nplanes, unsigned flags, uint8_t** outs, size_t* out_lens);
int j2kcuda_sequence_decode(j2kcuda_sequence* s, const uint8_t* const* streams, const size_t* lens,
                            int nframes, unsigned reduce, const j2kcuda_plane* const* frames,
                            int nplanes, unsigned flags);
/* totals over all frames since the last reset: the per-frame arrays of j2kcuda_encoder_timings /
 * _counters and j2kcuda_decoder_timings / _counters summed (8 doubles + 6 counters each way),
 * frames[0] = frames encoded, frames[1] = frames decoded.  Any pointer may be NULL. */
int j2kcuda_sequence_stats(j2kcuda_sequence* s, double* enc_ms, uint64_t* enc_cnt, double* dec_ms,
                           uint64_t* dec_cnt, uint64_t* frames, int reset);
/* one-shot forms (create a sequence, run it, destroy it) */
int j2kcuda_encode_batch(const j2kcuda_params* p, const j2kcuda_plane* const* frames, int nframes,
                         int nplanes, uint8_t** outs, size_t* out_lens);
int j2kcuda_decode_batch(const uint8_t* const* streams, const size_t* lens, int nframes,
                         unsigned reduce, const j2kcuda_plane* const* frames, int nplanes);

/* ---- several GPUs behind one call (device assignment inside the library; 8e: the path shards with no
 *      exchange step, so no collective is involved).  For a C++ host such as the plug-in; bench.py uses one process
 *      per GPU instead. ---------------------------------------------------------------------------------------- */
/* Frames of a sequence over `ndevices` GPUs: `workers_per_device` contexts on each, all pulling frames from one shared
 * counter.  Use with j2kcuda_sequence_encode / _decode / _stats / _destroy (host frames; the DEVICE_PIXELS flag is
 * for single-device sequences). */
j2kcuda_sequence* j2kcuda_sequence_create_multi(const j2kcuda_params* p, const int* devices, int ndevices,
                                                int workers_per_device);
/* Tiles of ONE very large frame over several GPUs (config C4): device k encodes / decodes a contiguous run of tiles
 * with its own context and host thread; encode concatenates the tile-part fragments behind one main header (the
 * result is byte-identical to a one-device encode), decode routes tile-parts by their SOT tile index
 * (j2k.cpp:9247-9333) and every device writes its tiles' rectangles into the caller's image (which must have the
 * codestream's own sample layout: interleaved, same sample size and depth).  p may be NULL for decode only. */
typedef struct j2kcuda_tiles j2kcuda_tiles;
j2kcuda_tiles* j2kcuda_tiles_create(const j2kcuda_params* p, const int* devices, int ndevices);
void j2kcuda_tiles_destroy(j2kcuda_tiles* t);
int j2kcuda_tiles_encode(j2kcuda_tiles* t, const j2kcuda_plane* planes, int nplanes, uint8_t** out, size_t* out_len);
int j2kcuda_tiles_decode(j2kcuda_tiles* t, const uint8_t* bytes, size_t n, unsigned reduce,
                         const j2kcuda_plane* planes, int nplanes);

/* ---- one frame whose tiles were encoded on several GPUs ----------------------------------- */
/* p describes the whole frame (tile_count ignored).  frags[i] are the fragments returned by encoders
 * created with tile ranges that together cover every tile exactly once, in any order; the result
 * is byte-identical to the codestream one context would have written for the whole frame
 * (tile-parts are self-delimiting: Psot, j2k.cpp:3826-3862).  Cinema profiles (TLM) are not
 * joinable.  *out is malloc'd. */
int j2kcuda_join_tile_streams(const j2kcuda_params* p, const uint8_t* const* frags, const size_t* lens,
                              int nfrags, uint8_t** out, size_t* out_len);

/* ---- JP2 file wrapper (host only; ext/grok/src/lib/openjp2/jp2.cpp) ------------------------- */
/* What OpenJPEGCodec::GetFileInfo takes from the JP2 header boxes through opj_read_header_ex
 * (j2k_openjpeg_codec.cpp:303-408): enumerated colour space or ICC profile, channel definitions,
 * palette + component mapping, resolution.  Replaces opj_jp2_read_header_procedure / read_ihdr /
 * read_bpcc / read_colr / read_cdef / read_pclr / read_cmap / read_res (jp2.cpp:560-1816, 2383-2500). */
#define J2KCUDA_JP2_MAX_CHANNELS 16
typedef struct j2kcuda_jp2_info {
    uint32_t is_jp2;                     /* 0: raw codestream, nothing else is set */
    uint32_t brand, minversion;          /* ftyp: 'jp2 ' = 0x6a703220, 'jpx ' = 0x6a707820 */
    uint32_t width, height, ncomp;       /* ihdr */
    uint32_t bpc;                        /* ihdr: (depth - 1) | sign << 7; 255 = per component, see bpcc */
    uint32_t unk_c, ipr;
    uint32_t bpcc[J2KCUDA_JP2_MAX_CHANNELS];   /* (depth - 1) | sign << 7 per component (from bpcc, else bpc) */
    uint32_t colr_method;                /* 0 no colr box, 1 enumerated, 2 restricted ICC, 3 any ICC, 4 vendor */
    uint32_t enumcs;                     /* 12 CMYK, 16 sRGB, 17 grey, 18 sYCC, 19 esYCC ... */
    uint32_t precedence, approx;
    uint64_t icc_offset, icc_length;     /* ICC profile bytes inside the probed buffer */
    uint32_t cdef_n;                     /* channel definitions: component cn has type typ (0 colour, 1 opacity, */
    uint16_t cdef_cn[J2KCUDA_JP2_MAX_CHANNELS];     /* 2 premultiplied opacity, 65535 unspecified) and is */
    uint16_t cdef_typ[J2KCUDA_JP2_MAX_CHANNELS];    /* associated with colour asoc (0 whole image, 1..n)  */
    uint16_t cdef_asoc[J2KCUDA_JP2_MAX_CHANNELS];
    uint32_t pclr_entries, pclr_channels;            /* palette: entries x channels */
    uint8_t  pclr_depth[J2KCUDA_JP2_MAX_CHANNELS];   /* (depth - 1) | sign << 7 per palette column */
    uint64_t pclr_offset;                /* first entry inside the probed buffer (big-endian, ceil(depth / 8) bytes each) */
    uint32_t cmap_n;                     /* component mapping: output channel i = component cmp, direct (mtyp 0) */
    uint16_t cmap_cmp[J2KCUDA_JP2_MAX_CHANNELS];    /* or through palette column pcol (mtyp 1) */
    uint8_t  cmap_mtyp[J2KCUDA_JP2_MAX_CHANNELS], cmap_pcol[J2KCUDA_JP2_MAX_CHANNELS];
    uint32_t has_capture_res, has_display_res;
    double   capture_res[2], display_res[2];         /* vertical, horizontal, grid points per metre */
    uint64_t codestream_offset, codestream_length;   /* contents of the jp2c box */
} j2kcuda_jp2_info;
int j2kcuda_probe_jp2(const uint8_t* bytes, size_t n, j2kcuda_jp2_info* out);
/* palette entries as uint32 [entry][channel] (what opj_jp2_pclr_t::entries holds, jp2.cpp:1317-1419);
 * returns the number of values written (entries x channels) or a negative error */
int j2kcuda_jp2_palette(const uint8_t* bytes, size_t n, uint32_t* entries, size_t cap);

/* Wrap a codestream into a JP2 file: jp, ftyp, jp2h{ihdr, [bpcc], colr, [cdef]}, jp2c -- the boxes
 * and defaults of opj_jp2_setup_encoder + opj_jp2_write_* (jp2.cpp:646-1110, 1817-2060, 2058-2262);
 * byte-identical to what the reference writes around the same codestream. */
typedef struct j2kcuda_jp2_params {
    uint32_t width, height, ncomp;
    uint32_t prec[J2KCUDA_JP2_MAX_CHANNELS], sgnd[J2KCUDA_JP2_MAX_CHANNELS];
    uint32_t enumcs;                     /* used when icc == NULL: 16 sRGB, 17 grey, 18 sYCC */
    const uint8_t* icc; size_t icc_len;  /* restricted ICC profile -> colr method 2 */
    int32_t  alpha_comp;                 /* component holding opacity (-> cdef box), -1 = none */
} j2kcuda_jp2_params;
int j2kcuda_wrap_jp2(const uint8_t* codestream, size_t len, const j2kcuda_jp2_params* p,
                     uint8_t** out, size_t* out_len);

/* ---- stage entry points (host arrays in/out; what the parity tests drive, one stage at a
 *      time, the way ext/grok's OPJ_PLUGIN_STATE_DEBUG_ENCODE cross-check does) -------------- */
/* RCT / ICT on three int32 planes (inverse ICT: float32 bit patterns). mct.cpp:94-471 */
int j2kcuda_stage_mct(int32_t* c0, int32_t* c1, int32_t* c2, size_t n, int irreversible, int inverse);
/* multi-level DWT of one tile-component, Mallat layout in and out. dwt.cpp:420-506, 615-704,
 * 927-1062.  (x0,y0,x1,y1) = canvas bounds of the tile-component. */
int j2kcuda_stage_dwt(int32_t* data, int x0, int y0, int x1, int y1, int numres, int irreversible,
                      int inverse);
/* Quantise + tier-1 encode `nblk` code-blocks cut from one plane.  blk[i] = {x, y, w, h, orient,
 * bandconst}; 5/3 when irreversible == 0.  Outputs per block: numbps, npasses, cumulative rates
 * (stride 128), per-pass nmsedec (stride 128), bytes at data + offsets[i]. t1_opt.cpp:659-739 */
int j2kcuda_stage_t1_encode(const int32_t* plane, int stride, int height, const int32_t* blk,
                            int nblk, int irreversible, int32_t* numbps, int32_t* npasses,
                            int32_t* rates, int32_t* nmsedec, uint8_t* data, size_t data_cap,
                            uint64_t* offsets);
/* Tier-1 decode: blk[i] = {w, h, orient, numbps, npasses, len}; bytes concatenated in `data`.
 * out = concatenated w*h int32 per block in the reference decoder's representation
 * (t1.cpp:1337-1448: magnitude with one extra fractional bit). */
int j2kcuda_stage_t1_decode(const uint8_t* data, const int32_t* blk, int nblk, int32_t* out);

/* ---- allocation-free forward DWT on DEVICE planes -------------------------------------------
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same transform as j2kcuda_stage_dwt(..., inverse = 0) -- 5/3, or the 9/7 in the
 * reference's Q13 fixed point when `irreversible` -- but input and output are the CALLER'S device
 * planes (int32, `pitch` samples per row, pitch >= x1 - x0 and a multiple of 4, both 16-byte
 * aligned, not aliased; the input is not modified) and the ping-pong plane + job table every
 * other path allocates live in one file-scope device array of fixed size inside the library
 * (dwt.cu; J2KCUDA_NOALLOC_MAX_SCRATCH_SAMPLES * 4 + 4096 bytes, held by every GPU that loads the
 * library until the process exits).  The call makes no cudaMalloc / cudaFree / allocating library
 * call, enqueues its kernels and device-to-device copies on `stream` (a cudaStream_t, NULL = the
 * default stream) of the CURRENT device and returns without synchronising.
 * Limit: pitch * ceil((y1 - y0) / 2) <= J2KCUDA_NOALLOC_MAX_SCRATCH_SAMPLES when numres >= 3
 * (e.g. 2048 x 2048 or 4096 x 1024), else J2KCUDA_ERR_UNSUPPORTED with the sizes in the message.
 * All calls on one device share the scratch array, which keeps the previous call's contents
 * (nothing in it has to start at zero: each level reads only what the level before wrote): calls
 * whose work overlaps on the GPU race -- order them on one stream or with events. */
#define J2KCUDA_NOALLOC_MAX_SCRATCH_SAMPLES (2u * 1024u * 1024u)       /* 8 MiB of int32 */
int j2kcuda_dwt_forward_noalloc(const int32_t* d_in, int32_t* d_out, int pitch, int x0, int y0,
                                int x1, int y1, int numres, int irreversible, void* stream);

/* Host Tier-2 alone, from externally supplied code-block results (the grok plug-in hand-off,
 * plugin_bridge.cpp:72-174): blocks in tile -> component -> resolution -> band -> precinct ->
 * code-block order.  rates are raw cumulative `mqc_numbytes + 3` values already clamped and
 * 0xFF-trimmed (t1_opt.cpp:722-736); dist are cumulative distortion decreases. */
typedef struct j2kcuda_block_result {
    const uint8_t* data;
    int32_t numbps, npasses;
    const int32_t* rates;
    const double* dist;
} j2kcuda_block_result;
int j2kcuda_t2_block_count(const j2kcuda_params* p);
/* The other direction of the same hand-off: the DEVICE results of every code-block of the frame a context has just run
 * (after j2kcuda_encoder_run), in the same order; returns the number of blocks or -1.  rates as above, dist cumulative.
 * Valid until the next upload / run on the context.  What integration/grok_plugin.cpp fills opj_plugin_tile_t from
 * (openjpeg.h:1308-1360). */
int j2kcuda_encoder_block_results(j2kcuda_encoder* e, j2kcuda_block_result* out, int max_blocks);
/* DWT coefficients of one component after j2kcuda_encoder_run, in the reference's in-place Mallat tile buffer layout
 * (width x height int32): the image data grok's OPJ_PLUGIN_STATE_DEBUG_ENCODE cross-check expects.  Single-tile frames. */
int j2kcuda_encoder_download_coefficients(j2kcuda_encoder* e, int comp, int32_t* out);
/* Geometry of every code-block in that order, 12 int32 per block: tile, comp, x, y (position in the
 * tile-component's Mallat plane), w, h, orient, level (numres-1-resno), band numbps, bandconst
 * (t1.cpp:1482), band stepsize (float32 bits), resno.  Returns blocks written.  (tcd.cpp:592-842) */
int j2kcuda_t2_block_geometry(const j2kcuda_params* p, int32_t* geo, int max_blocks);
/* 8 int32 per block, same order: band index within its resolution, precinct index, index within the precinct, precincts of the
 * resolution (empty ones included), bounds in band coordinates x0, y0, x1, y1. */
int j2kcuda_t2_block_structure(const j2kcuda_params* p, int32_t* out, int max_blocks);
int j2kcuda_t2_assemble(const j2kcuda_params* p, const j2kcuda_block_result* blocks, int nblocks,
                        uint8_t** out, size_t* out_len);
/* Host Tier-2 parse alone: returns, per code-block in the same order, numbps, npasses, len and
 * offset of its (concatenated) bytes inside `bytes`-derived scratch returned in *seg (malloc'd). */
int j2kcuda_t2_parse(const uint8_t* bytes, size_t n, unsigned reduce, int32_t** blkinfo,
                     int* nblocks, uint8_t** seg, size_t* seg_len);

#ifdef __cplusplus
}
#endif
#endif
