// Host-side generation of the Tier-1 tables (see t1_tables.cuh) and their upload.
#include "t1_tables.cuh"
#include "device.h"
#include <cmath>
#include <cstring>
#include <mutex>
#include <map>

namespace j2k {
namespace {

// ISO 15444-1 Table D.1 (t1_luts.h lut_ctxno_zc tabulates the same rule)
int zc_rule(int orient, int h, int v, int d) {
    if (orient == 1) { int t = h; h = v; v = t; }
    if (orient == 3) {
        int hv = h + v;
        if (d >= 3) return 8;
        if (d == 2) return hv >= 1 ? 7 : 6;
        if (d == 1) return hv >= 2 ? 5 : (hv == 1 ? 4 : 3);
        return hv >= 2 ? 2 : (hv == 1 ? 1 : 0);
    }
    if (h == 2) return 8;
    if (h == 1) return v >= 1 ? 7 : (d >= 1 ? 6 : 5);
    if (v == 2) return 4;
    if (v == 1) return 3;
    return d >= 2 ? 2 : (d == 1 ? 1 : 0);
}

// ISO 15444-1 Tables D.2/D.3
void sc_rule(int hc, int vc, int& ctx, int& xorbit) {
    hc = hc > 1 ? 1 : (hc < -1 ? -1 : hc);
    vc = vc > 1 ? 1 : (vc < -1 ? -1 : vc);
    xorbit = 0;
    if (hc < 0 || (hc == 0 && vc < 0)) { xorbit = 1; hc = -hc; vc = -vc; }
    if (hc == 1) ctx = vc == 1 ? 4 : (vc == 0 ? 3 : 2);
    else ctx = vc == 1 ? 1 : 0;
}

const uint16_t kQe[47] = {
    0x5601, 0x3401, 0x1801, 0x0AC1, 0x0521, 0x0221, 0x5601, 0x5401, 0x4801, 0x3801, 0x3001, 0x2401,
    0x1C01, 0x1601, 0x5601, 0x5401, 0x5101, 0x4801, 0x3801, 0x3401, 0x3001, 0x2801, 0x2401, 0x2201,
    0x1C01, 0x1801, 0x1601, 0x1401, 0x1201, 0x1101, 0x0AC1, 0x09C1, 0x08A1, 0x0521, 0x0441, 0x02A1,
    0x0221, 0x0141, 0x0111, 0x0085, 0x0049, 0x0025, 0x0015, 0x0009, 0x0005, 0x0001, 0x5601};
const uint8_t kNmps[47] = {
    1, 2, 3, 4, 5, 38, 7, 8, 9, 10, 11, 12, 13, 29, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27,
    28, 29, 30, 31, 32, 33, 34, 35, 36, 37, 38, 39, 40, 41, 42, 43, 44, 45, 45, 46};
const uint8_t kNlps[47] = {
    1, 6, 9, 12, 29, 33, 6, 14, 14, 14, 17, 18, 20, 21, 14, 14, 15, 16, 17, 18, 19, 19, 20, 21, 22, 23,
    24, 25, 26, 27, 28, 29, 30, 31, 32, 33, 34, 35, 36, 37, 38, 39, 40, 41, 42, 43, 46};
const uint8_t kSwitch[47] = {1, 0, 0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 1};

T1Tables build() {
    T1Tables t;
    memset(&t, 0, sizeof(t));
    for (int orient = 0; orient < 4; ++orient)
        for (int nb = 0; nb < 256; ++nb) {
            int h = ((nb >> 3) & 1) + ((nb >> 4) & 1);
            int v = ((nb >> 1) & 1) + ((nb >> 6) & 1);
            int d = (nb & 1) + ((nb >> 2) & 1) + ((nb >> 5) & 1) + ((nb >> 7) & 1);
            t.zc[orient * 256 + nb] = (uint8_t)(CTX_ZC + zc_rule(orient, h, v, d));
        }
    for (int orient = 0; orient < 4; ++orient)
        for (int n = 0; n < 512; ++n) {
            int h = ((n >> 3) & 1) + ((n >> 5) & 1);
            int v = ((n >> 1) & 1) + ((n >> 7) & 1);
            int d = (n & 1) + ((n >> 2) & 1) + ((n >> 6) & 1) + ((n >> 8) & 1);
            t.zc9[orient * 512 + n] = (uint8_t)(CTX_ZC + zc_rule(orient, h, v, d));
        }
    for (int i = 0; i < 256; ++i) {
        auto contrib = [&](int sigbit, int negbit) { return ((i >> sigbit) & 1) ? (((i >> negbit) & 1) ? -1 : 1) : 0; };
        int hc = contrib(0, 4) + contrib(1, 5), vc = contrib(2, 6) + contrib(3, 7);
        int ctx, xb;
        sc_rule(hc, vc, ctx, xb);
        t.sc[i] = (uint8_t)(ctx | (xb << 3));
    }
    for (int i = 0; i < 128; ++i) {          // t1_generate_luts.cpp:434-456
        double x = i / 64.0, u = x, v = x - 1.5;
        auto q = [](double e) { int r = (int)(std::floor(e * 64.0 + 0.5) / 64.0 * 8192.0); return (int16_t)(r < 0 ? 0 : r); };
        t.nm_sig[i] = q(u * u - v * v);
        t.nm_sig0[i] = q(u * u);
        u = x - 1.0;
        v = (i & 64) ? x - 1.5 : x - 0.5;
        t.nm_ref[i] = q(u * u - v * v);
        t.nm_ref0[i] = q(u * u);
    }
    for (int s = 0; s < 47; ++s)
        for (int mps = 0; mps < 2; ++mps) {
            uint32_t nm = (uint32_t)kNmps[s] * 2 + mps;
            uint32_t nl = (uint32_t)kNlps[s] * 2 + (kSwitch[s] ? 1 - mps : mps);
            t.mq[s * 2 + mps] = kQe[s] | (nm << 16) | (nl << 24);
        }
    return t;
}

std::mutex g_mu;
std::map<int, T1Tables*> g_dev;     // per device ordinal

}  // namespace

const T1Tables& t1_tables_host() {
    static const T1Tables t = build();
    return t;
}

const char* device_tables_init() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return "cudaGetDevice failed";
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_dev.count(dev)) return nullptr;
    T1Tables* d = nullptr;
    if (cudaMalloc(&d, sizeof(T1Tables)) != cudaSuccess) return "cudaMalloc(tables) failed";
    if (cudaMemcpy(d, &t1_tables_host(), sizeof(T1Tables), cudaMemcpyHostToDevice) != cudaSuccess) return "table upload failed";
    g_dev[dev] = d;
    return nullptr;
}

const T1Tables* t1_tables_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_dev.find(dev);
    return it == g_dev.end() ? nullptr : it->second;
}

}  // namespace j2k
