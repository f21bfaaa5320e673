/*
 * j2k_oracle.h -- TEST INFRASTRUCTURE ONLY (never linked into libj2kcuda).
 *
 * Plain-C, single-threaded restatement of the reference's hot-path arithmetic
 * (ext/grok/src/lib/openjp2: mct.cpp, dwt.cpp, T1Encoder.cpp, T1Decoder.cpp, t1_opt.cpp, t1.cpp,
 * mqc.cpp).  Written from the algorithm, not from the reference's code layout: per-sample state
 * arrays instead of packed flag words, absolute-coordinate lifting instead of interleave buffers.
 * Parity pinned differentially against oracle/_ref/libgrokref.so (the reference compiled here)
 * and against the fixtures in tests/golden/ (written by tests/golden/make_golden.py from the compiled reference and
 * OpenJPEG 2.5.x; tests/test_oracle_vs_ref.py, tests/test_golden.py).
 */
#ifndef J2K_ORACLE_H
#define J2K_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

/* mct.cpp:94-157 / 163-227 (RCT), 241-401 (ICT fwd, Q13), 406-471 (ICT inv, float) */
void orc_rct_fwd(int32_t* c0, int32_t* c1, int32_t* c2, size_t n);
void orc_rct_inv(int32_t* c0, int32_t* c1, int32_t* c2, size_t n);
void orc_ict_fwd(int32_t* c0, int32_t* c1, int32_t* c2, size_t n);
void orc_ict_inv(float* c0, float* c1, float* c2, size_t n);

/* dwt.cpp:420-506 (fwd), 615-704 (inv 5/3), 927-1062 (inv 9/7 float).  `a` is the tile-component
 * buffer [y1-y0][x1-x0]; (x0,y0,x1,y1) are its canvas bounds; numres = decomposition levels+1.
 * Output / input is the Mallat layout. */
void orc_dwt53_fwd(int32_t* a, int x0, int y0, int x1, int y1, int numres);
void orc_dwt53_inv(int32_t* a, int x0, int y0, int x1, int y1, int numres);
void orc_dwt97_fwd(int32_t* a, int x0, int y0, int x1, int y1, int numres);   /* Q13 fixed point */
void orc_dwt97_inv(float* a, int x0, int y0, int x1, int y1, int numres);

/* T1Encoder.cpp:103-174: quantise one block into sign-magnitude (bit 31 sign, 6 fractional bits).
 * Returns max magnitude. */
uint32_t orc_quantize_block(const int32_t* coeff, int stride, int w, int h, int qmfbid,
                            int32_t bandconst, uint32_t* out);

/* t1_opt.cpp:659-739.  smag = w*h sign-magnitude samples.  rates[] = cumulative, already clamped
 * and 0xFF-trimmed like the reference; nmsedec[] = per-pass integer distortion (t1.cpp:298-314).
 * Returns total bytes written to out, or -1 if cap too small. */
int orc_t1_encode_cblk(const uint32_t* smag, int w, int h, int orient, uint32_t maxmag,
                       uint8_t* out, int cap, int* numbps, int* npasses, int* rates, int* nmsedec);

/* t1.cpp:1114-1143 */
double orc_wmsedec(int nmsedec, double mct_norm, double dwt_norm, double stepsize, int bpno);
double orc_dwt_norm(int level, int orient, int real);
double orc_mct_norm(int compno, int real);

/* t1.cpp:1337-1448 for cblksty == 0, one segment.  out = w*h values in the reference decoder's
 * representation (magnitude*2 + half bit, signed).  Returns 0. */
int orc_t1_decode_cblk(const uint8_t* data, int len, int w, int h, int orient, int numbps,
                       int npasses, int32_t* out);

/* LUT accessors so tests can compare our tables with what the kernels embed */
int orc_zc_context(int orient, int h, int v, int d);
void orc_mq_table(uint16_t* qe, uint8_t* nmps, uint8_t* nlps, uint8_t* sw); /* 47 entries each */

#ifdef __cplusplus
}
#endif
#endif
