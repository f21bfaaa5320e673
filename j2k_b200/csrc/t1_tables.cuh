// Tier-1 lookup tables shared by the encoder and decoder kernels.  They are generated on the host
// from the coding rules (ISO 15444-1 Tables D.1-D.3, C.2; the reference tabulates the same rules in
// ext/grok/src/lib/openjp2/t1_luts.h and mqc.cpp:120-215) and uploaded once per device.
#pragma once
#include <cstdint>

namespace j2k {

enum { CTX_ZC = 0, CTX_SC = 9, CTX_MAG = 14, CTX_AGG = 17, CTX_UNI = 18, T1_NCTX = 19 };

struct T1Tables {
    // zero-coding context: index = orient*256 + nb, nb bits: 0 NW, 1 N, 2 NE, 3 W, 4 E, 5 SW, 6 S, 7 SE
    uint8_t zc[4 * 256];
    // sign coding: index bits 0 W sig, 1 E sig, 2 N sig, 3 S sig, 4 W neg, 5 E neg, 6 N neg, 7 S neg
    // value = (ctx - CTX_SC) | xorbit << 3
    uint8_t sc[256];
    // distortion estimates (t1_generate_luts.cpp:434-456)
    int16_t nm_sig[128], nm_sig0[128], nm_ref[128], nm_ref0[128];
    // MQ state machine, index = state*2 + mps: qe | next-if-MPS << 16 | next-if-LPS << 24
    // (next values are again state*2 + mps, the LPS one with the switch applied)
    uint32_t mq[94];
    uint32_t pad[2];
    // zero-coding context from the full 3x3 neighbourhood: index = orient*512 + n9, n9 bits: 0 NW, 1 N, 2 NE,
    // 3 W, 4 centre (ignored), 5 E, 6 SW, 7 S, 8 SE (same idea as lut_ctxno_zc_opt, t1_luts.h)
    uint8_t zc9[4 * 512];
};

const T1Tables* t1_tables_device();     // device pointer (valid after device_tables_init)
const T1Tables& t1_tables_host();

}  // namespace j2k
