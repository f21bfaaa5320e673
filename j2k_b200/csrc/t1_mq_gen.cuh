// MQ encoder of one code-block for the code-block styles that only touch the arithmetic coder: RESET (contexts back to their
// initial states after every pass), TERMALL (the coder is flushed and restarted after every pass, so every pass is a
// codeword segment of its own) and SEGSYM (the symbols 1, 0, 1, 0 in the UNI context after every clean-up pass: they are coded
// here, the modelling kernel does not know about them).  PTERM is left out on purpose: the reference's predictable termination
// (mqc.cpp:534-548 with the byte count of t1.cpp:1667-1670) writes small blocks that neither its own decoder nor OpenJPEG reads
// back (64 x 64 image, six resolutions), and there is nothing to pin a corrected version against.  This is the reference's GENERIC loop (ext/grok/src/lib/openjp2/t1.cpp:1537-1673), not the
// optimised one k_t1_mq follows (t1_opt.cpp:659-739): a pass's rate is numbytes + 3, or + 1 when the pass was terminated, and
// only when numbytes > 0; the restart steps back over the last byte (mqc.cpp:521-532) and drops a trailing 0xFF; the final
// flush comes on top of the last pass's own.  The context modelling is unchanged, so the input is the same (context, decision)
// symbol stream k_t1_context writes.
//
// Shared between the CUDA kernel (t1_enc.cu: k_t1_mq_gen, one block per thread) and a host-compiled harness
// (tests/native/t1dec_host.cpp: mqgen_host) that is compared with the compiled reference encoder block by block.
// Written after the reference's byte-pointer form (mqc.cpp:223-306, 441-452) rather than k_t1_mq's branch-free register
// form: streams with these styles are rare and what matters here is that every quirk is reproduced.
#pragma once
#include <cstdint>
#include "t1_tables.cuh"
#ifdef __CUDACC__
#define J2K_MQG_HD __host__ __device__
#else
#define J2K_MQG_HD
#endif

namespace j2k {
namespace mqgen {

enum { STY_RESET = 0x02, STY_TERMALL = 0x04, STY_SEGSYM = 0x20, STY_SUPPORTED = STY_RESET | STY_TERMALL | STY_SEGSYM };

struct Enc {
    uint32_t a, c, ct;
    int bp;                 // index of the last byte written; -1 = the reference's byte in front of the buffer (holds 0)
    uint8_t* out;
    int cap;
    uint8_t pre;            // that byte: a carry out of the very first byte lands here and is lost, as in the reference
    bool overflow;
    J2K_MQG_HD uint8_t get(int i) const { return i < 0 ? pre : out[i]; }
    J2K_MQG_HD void set(int i, uint8_t v) { if (i < 0) pre = v; else if (i < cap) out[i] = v; else overflow = true; }
    J2K_MQG_HD void byteout() {                                  // mqc.cpp:223-257
        if (get(bp) == 0xff) { ++bp; set(bp, (uint8_t)(c >> 20)); c &= 0xfffff; ct = 7; }
        else if ((c & 0x8000000u) == 0) { ++bp; set(bp, (uint8_t)(c >> 19)); c &= 0x7ffff; ct = 8; }
        else {
            set(bp, (uint8_t)(get(bp) + 1));
            if (get(bp) == 0xff) { c &= 0x7ffffff; ++bp; set(bp, (uint8_t)(c >> 20)); c &= 0xfffff; ct = 7; }
            else { ++bp; set(bp, (uint8_t)(c >> 19)); c &= 0x7ffff; ct = 8; }
        }
    }
    J2K_MQG_HD void renorm() { do { a <<= 1; c <<= 1; --ct; if (ct == 0) byteout(); } while ((a & 0x8000u) == 0); }   // mqc.cpp:259-269
    J2K_MQG_HD void flush() {                                    // mqc.cpp:299-306, 441-452
        const uint32_t tempc = c + a;
        c |= 0xffffu;
        if (c >= tempc) c -= 0x8000u;
        c <<= ct; byteout();
        c <<= ct; byteout();
        if (get(bp) != 0xff) ++bp;
    }
    J2K_MQG_HD int numbytes() const { return bp < 0 ? 0 : bp; }  // mqc.cpp:404-410 (bp - start, where start = index 0)
};

J2K_MQG_HD inline void reset_states(uint8_t* st) {               // t1.cpp:1578-1581, mqc.cpp:496-502: index = state * 2 + mps
    for (int i = 0; i < 32; ++i) st[i] = 0;
    st[CTX_UNI] = 46 * 2; st[CTX_AGG] = 3 * 2; st[CTX_ZC] = 4 * 2;
}

J2K_MQG_HD inline void code(Enc& q, uint8_t* st, const uint32_t* mq, uint32_t ctx, uint32_t d) {
    const uint32_t e = mq[st[ctx]], qe = e & 0xffffu;
    q.a -= qe;
    if ((st[ctx] & 1u) == d) {                                   // mqc.cpp:271-283 (codemps)
        if ((q.a & 0x8000u) == 0) {
            if (q.a < qe) q.a = qe; else q.c += qe;
            st[ctx] = (uint8_t)((e >> 16) & 0xffu);
            q.renorm();
        } else q.c += qe;
    } else {                                                     // mqc.cpp:285-297 (codelps)
        if (q.a < qe) q.c += qe; else q.a = qe;
        st[ctx] = (uint8_t)(e >> 24);
        q.renorm();
    }
}

// sym: (ctx << 1) | d per decision; pass_nsym[p] decisions belong to pass p.  mq: T1Tables::mq (qe | next-if-MPS << 16 |
// next-if-LPS << 24, indices state * 2 + mps).  rate[p] = cumulative bytes after pass p as the reference reports them.
// Returns 0, or 2 when `cap` bytes do not hold the block.
J2K_MQG_HD inline int encode_block(const uint8_t* sym, const uint32_t* pass_nsym, int npasses, int cblksty, const uint32_t* mq,
                                   uint8_t* out, int cap, int32_t* rate, int32_t* total) {
    Enc q;
    q.a = 0x8000u; q.c = 0; q.ct = 12; q.bp = -1; q.out = out; q.cap = cap; q.pre = 0; q.overflow = false;   // mqc.cpp:412-419
    uint8_t st[32];
    reset_states(st);
    uint32_t pos = 0;
    for (int p = 0; p < npasses; ++p) {
        const uint32_t end = pos + pass_nsym[p];
        for (; pos < end; ++pos) { const uint32_t sb = sym[pos]; code(q, st, mq, (sb >> 1) & 31u, sb & 1u); }
        if ((cblksty & STY_SEGSYM) && p % 3 == 0)                // the first pass and every third after it are clean-up passes
            for (uint32_t i = 1; i < 5; ++i) code(q, st, mq, CTX_UNI, i & 1u);   // t1.cpp:1598-1600, mqc.cpp:550-558: 1, 0, 1, 0
        const int correction = 3;
        if (cblksty & STY_TERMALL) {                             // t1.cpp:1618-1647
            q.flush();
            q.a = 0x8000u; q.c = 0; q.ct = 12; --q.bp;           // opj_mqc_restart_init_enc
            if (q.get(q.bp) == 0xff) q.ct = 13;
        }
        // The reference adds the correction only when numbytes > 0 (t1.cpp:1649-1651).  After a termination that loses a segment
        // that is exactly one byte long (last byte at index 0: "no bytes" and "one byte" both give numbytes 0): the pass gets
        // length 0 and neither the reference's decoder nor OpenJPEG can read the block (tests/test_t1_mq_generic.py).  A
        // terminated pass always owns the bytes up to and including the one the restart stepped back to.
        rate[p] = (cblksty & STY_TERMALL) ? q.bp + 1 : q.numbytes();
        if (!(cblksty & STY_TERMALL) && rate[p] > 0) rate[p] += correction;
        if (cblksty & STY_RESET) reset_states(st);               // t1.cpp:1654-1656
    }
    q.flush();                                                   // t1.cpp:1659-1663 (neither PTERM nor LAZY)
    const int nbytes = q.numbytes();
    if (q.overflow || nbytes > cap) { *total = 0; return 2; }
    int last = 0;
    for (int p = 0; p < npasses; ++p) {                          // t1.cpp:1667-1677
        int r = rate[p];
        if (r > nbytes) r = nbytes;
        if (r > 1 && out[r - 1] == 0xff) --r;
        rate[p] = r;
        last = r;
    }
    *total = last;
    return 0;
}

}  // namespace mqgen
}  // namespace j2k
