// Sample conversion between the caller's channels and the codestream's components: bit-depth change by shift and bit
// replication, sign offset -- the arithmetic of Codec::CopyBuffer / CopyChannel<DESTTYPE, SRCTYPE>
// (src/common/j2k_codec.cpp:222-427), one sample at a time, every branch of it:
//   no shift (:266-279), up-shift with one fill (:285-301) and two fills (:302-325) for depths >= 8, the repeated
//   doubling for depths < 8 (:326-355), down-shift (:357-375), signConverter (:252-254), and the truncations the
//   reference's typed temporaries make (const SRCTYPE v, const DESTTYPE t, DESTTYPE *d).
// Used by the MCT kernels (encode: caller type -> int component; decode: int component -> caller type).
#pragma once
#include <cstdint>
#ifdef __CUDACC__
#define J2K_CB_HD __host__ __device__
#else
#define J2K_CB_HD
#endif

namespace j2k {

struct CopyConv {              // one channel
    int8_t src_type, dst_type; // J2KCUDA_UCHAR / USHORT / UINT / INT (0..3)
    int8_t src_depth, dst_depth;
    int32_t sign_conv;         // +-2^(src_depth-1) when exactly one side is signed
    int32_t plain;             // no shift, no sign offset: the sample passes through
    int32_t sext_bits;         // signed samples in an 8/16-bit container (beyond the reference, which asserts): sign-extend first
};

// the value a variable of the given type holds after assignment, widened to 64 bits
J2K_CB_HD inline long long cb_as(int type, long long v) {
    switch (type) {
    case 0: return v & 0xff;
    case 1: return v & 0xffff;
    case 2: return v & 0xffffffffll;
    default: return (long long)(int32_t)(uint32_t)(v & 0xffffffffll);
    }
}

// s = the source sample as loaded (zero-extended for unsigned types, sign-extended for INT)
J2K_CB_HD inline long long copy_convert(long long s, const CopyConv& c) {
    if (c.sext_bits) s = (long long)((int32_t)((uint32_t)s << (32 - c.sext_bits)) >> (32 - c.sext_bits));
    const long long v = cb_as(c.src_type, s + c.sign_conv);                // const SRCTYPE v = (*s + signConverter)
    const int shift = (int)c.dst_depth - (int)c.src_depth;
    long long r;
    if (shift == 0) r = v;
    else if (shift > 0) {
        if (c.src_depth >= 8) {
            if (shift <= c.src_depth) r = (cb_as(c.dst_type, v) << shift) | (v >> (c.src_depth - shift));
            else {
                const int first = c.src_depth, second = shift - first, fill = 2 * c.src_depth - second;
                const long long t = cb_as(c.dst_type, (cb_as(c.dst_type, v) << first) | v);   // const DESTTYPE t
                r = (t << second) | (t >> fill);
            }
        } else {
            int pix = c.src_depth;
            long long t = cb_as(c.dst_type, s + c.sign_conv);              // DESTTYPE v = (*s + signConverter)
            while (pix * 2 < c.dst_depth) { t = cb_as(c.dst_type, (t << pix) | t); pix *= 2; }
            const int second = c.dst_depth - pix, fill = pix - second;
            r = (t << second) | (t >> fill);
        }
    } else r = v >> -shift;
    return cb_as(c.dst_type, r);
}

inline CopyConv make_copy_conv(int src_type, int src_depth, int src_sgnd, int dst_type, int dst_depth, int dst_sgnd) {
    CopyConv c;
    c.sext_bits = (src_sgnd && src_type < 2) ? (src_type == 0 ? 8 : 16) : 0;
    if (c.sext_bits) src_type = 3;
    c.src_type = (int8_t)src_type; c.dst_type = (int8_t)dst_type;
    c.src_depth = (int8_t)src_depth; c.dst_depth = (int8_t)dst_depth;
    c.sign_conv = (!dst_sgnd == !src_sgnd) ? 0 : (src_sgnd ? (int32_t)(1ll << (src_depth - 1)) : -(int32_t)(1ll << (src_depth - 1)));
    c.plain = (src_depth == dst_depth && c.sign_conv == 0 && !c.sext_bits) ? 1 : 0;
    return c;
}

}  // namespace j2k
