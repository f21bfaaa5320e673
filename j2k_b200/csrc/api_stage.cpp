// C-ABI lifetime calls and the single-stage entry points the parity tests drive (host arrays in and
// out, one device stage each), in the spirit of the reference's OPJ_PLUGIN_STATE_DEBUG_ENCODE
// cross-check (ext/grok/src/lib/openjp2/plugin_bridge.cpp:86-170).  See include/j2kcuda.h.
#include "../../include/j2kcuda.h"
#include "common.h"
#include "device.h"
#include "geometry.h"
#include <cstring>
#include <vector>

using namespace j2k;

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return set_error(J2KCUDA_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

namespace {
struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
    template <class T> T* as() { return (T*)p; }
};
int need_device() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        return set_error(J2KCUDA_ERR_NO_DEVICE, "no CUDA device: libj2kcuda has no CPU path");
    }
    if (const char* err = device_tables_init()) return set_error(J2KCUDA_ERR_CUDA, err);
    return J2KCUDA_OK;
}
}  // namespace

extern "C" {

int j2kcuda_device_count(void) { return j2k::guarded([&]() -> int {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}); }

int j2kcuda_init(int device, j2kcuda_log_fn log, void* user) { return j2k::guarded([&]() -> int {
    j2k::set_log(log, user);
    int n = j2kcuda_device_count();
    if (n == 0) return set_error(J2KCUDA_ERR_NO_DEVICE, "no CUDA device: libj2kcuda has no CPU path");
    if (device >= n) return set_error(J2KCUDA_ERR_PARAM, "device ordinal out of range");
    if (device >= 0) CK(cudaSetDevice(device));
    if (const char* err = device_tables_init()) return set_error(J2KCUDA_ERR_CUDA, err);
    return J2KCUDA_OK;
}); }

void j2kcuda_shutdown(void) { j2k::encoder_pool_clear(); j2k::decoder_pool_clear(); j2k::out_release_all(); j2k::set_log(nullptr, nullptr); }

int j2kcuda_stage_mct(int32_t* c0, int32_t* c1, int32_t* c2, size_t n, int irreversible, int inverse) { return j2k::guarded([&]() -> int {
    if (!c0 || !c1 || !c2) return set_error(J2KCUDA_ERR_PARAM, "null plane");
    if (int rc = need_device()) return rc;
    DevBuf d[3];
    int32_t* h[3] = {c0, c1, c2};
    for (int i = 0; i < 3; ++i) { CK(d[i].alloc(n * 4)); CK(cudaMemcpy(d[i].p, h[i], n * 4, cudaMemcpyHostToDevice)); }
    launch_mct_planes(d[0].as<int32_t>(), d[1].as<int32_t>(), d[2].as<int32_t>(), n, irreversible, inverse, 0);
    CK(cudaGetLastError());
    for (int i = 0; i < 3; ++i) CK(cudaMemcpy(h[i], d[i].p, n * 4, cudaMemcpyDeviceToHost));
    return J2KCUDA_OK;
}); }

int j2kcuda_stage_dwt(int32_t* data, int x0, int y0, int x1, int y1, int numres, int irreversible, int inverse) { return j2k::guarded([&]() -> int {
    if (!data || x1 <= x0 || y1 <= y0 || numres < 1 || numres > 33) return set_error(J2KCUDA_ERR_PARAM, "bad region");
    if (int rc = need_device()) return rc;
    const int w = x1 - x0, h = y1 - y0, pitch = (w + 3) & ~3, levels = numres - 1;
    DevBuf A, B, J;
    const size_t bytes = (size_t)pitch * h * 4;
    CK(A.alloc(bytes)); CK(B.alloc(bytes));
    CK(cudaMemcpy2D(A.p, (size_t)pitch * 4, data, (size_t)w * 4, (size_t)w * 4, h, cudaMemcpyHostToDevice));
    CK(cudaMemcpy2D(B.p, (size_t)pitch * 4, data, (size_t)w * 4, (size_t)w * 4, h, cudaMemcpyHostToDevice));
    CK(J.alloc(sizeof(DwtJob)));
    auto plane = [&](int l) { return (l & 1) ? B.as<int32_t>() : A.as<int32_t>(); };
    const int kind = !irreversible ? DWT_53 : (inverse ? DWT_97_FLOAT : DWT_97_FIX);
    auto region = [&](int l, DwtJob& j) {        // input region of forward level l (1-based)
        j.x0 = (int)ceildivpow2((uint32_t)x0, l - 1); j.x1 = (int)ceildivpow2((uint32_t)x1, l - 1);
        j.y0 = (int)ceildivpow2((uint32_t)y0, l - 1); j.y1 = (int)ceildivpow2((uint32_t)y1, l - 1);
        j.pitch = pitch; j.strip0 = j.nstrips = j.nsegs = 0;
    };
    if (!inverse) {
        for (int l = 1; l <= levels; ++l) {
            DwtJob j; region(l, j);
            if (j.x1 <= j.x0 || j.y1 <= j.y0) continue;
            j.src = plane(l - 1); j.dst = plane(l);
            int items = dwt_plan_jobs(&j, 1);
            CK(cudaMemcpy(J.p, &j, sizeof(j), cudaMemcpyHostToDevice));
            launch_dwt_forward(J.as<DwtJob>(), 1, items, j.strip0, kind, 0);
            CK(cudaDeviceSynchronize());
        }
        std::vector<int32_t> ha((size_t)pitch * h), hb((size_t)pitch * h);
        CK(cudaMemcpy(ha.data(), A.p, bytes, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(hb.data(), B.p, bytes, cudaMemcpyDeviceToHost));
        // level l's detail bands live in plane(l), the last LL in plane(levels)
        for (int l = 1; l <= levels; ++l) {
            DwtJob j; region(l, j);
            const int rw = j.x1 - j.x0, rh = j.y1 - j.y0;
            if (rw <= 0 || rh <= 0) continue;
            const int snx = (int)ceildivpow2((uint32_t)j.x1, 1) - (int)ceildivpow2((uint32_t)j.x0, 1);
            const int sny = (int)ceildivpow2((uint32_t)j.y1, 1) - (int)ceildivpow2((uint32_t)j.y0, 1);
            const std::vector<int32_t>& src = (l & 1) ? hb : ha;
            for (int y = 0; y < rh; ++y)
                for (int x = 0; x < rw; ++x)
                    if (l == levels || x >= snx || y >= sny) data[(size_t)y * w + x] = src[(size_t)y * pitch + x];
        }
    } else {
        for (int l = levels; l >= 1; --l) {
            DwtJob j; region(l, j);
            if (j.x1 <= j.x0 || j.y1 <= j.y0) continue;
            j.src = plane(l); j.dst = plane(l - 1);
            int items = dwt_plan_jobs(&j, 1);
            CK(cudaMemcpy(J.p, &j, sizeof(j), cudaMemcpyHostToDevice));
            launch_dwt_inverse(J.as<DwtJob>(), 1, items, j.strip0, kind, 0);
            CK(cudaDeviceSynchronize());
        }
        CK(cudaMemcpy2D(data, (size_t)w * 4, A.p, (size_t)pitch * 4, (size_t)w * 4, h, cudaMemcpyDeviceToHost));
    }
    CK(cudaGetLastError());
    return J2KCUDA_OK;
}); }

int j2kcuda_stage_t1_encode(const int32_t* plane, int stride, int height, const int32_t* blk, int nblk, int irreversible,
                            int32_t* numbps, int32_t* npasses, int32_t* rates, int32_t* nmsedec, uint8_t* data,
                            size_t data_cap, uint64_t* offsets) { return j2k::guarded([&]() -> int {
    if (!plane || !blk || nblk < 0 || !numbps || !npasses || !rates || !nmsedec || !data || !offsets)
        return set_error(J2KCUDA_ERR_PARAM, "null argument");
    if (int rc = need_device()) return rc;
    DevBuf P, BL, SYM, OUT, PN, RES;
    CK(P.alloc((size_t)stride * height * 4));
    CK(cudaMemcpy(P.p, plane, (size_t)stride * height * 4, cudaMemcpyHostToDevice));
    std::vector<T1Block> hb(nblk);
    size_t sym_units = 0, out_off = 0;
    for (int i = 0; i < nblk; ++i) {
        const int32_t* g = blk + 6 * (size_t)i;
        if (g[2] < 0 || g[3] < 0 || g[2] > 64 || g[3] > 64 || g[0] < 0 || g[1] < 0 || g[0] + g[2] > stride || g[1] + g[3] > height)
            return set_error(J2KCUDA_ERR_PARAM, "block outside the plane or larger than 64x64");
        T1Block& b = hb[i];
        memset(&b, 0, sizeof(b));
        b.coeff = P.as<int32_t>() + (size_t)g[1] * stride + g[0];
        b.pitch = stride; b.w = (uint16_t)g[2]; b.h = (uint16_t)g[3];
        b.orient = (uint8_t)g[4]; b.qmfbid = irreversible ? 0 : 1; b.bandconst = g[5]; b.max_planes = 26;
        const size_t area = (size_t)b.w * b.h;
        size_t cap = (area * 28 * 3 / 2 + area + 64 + 15) & ~(size_t)15;
        b.sym_off = (uint32_t)sym_units; b.sym_cap = (uint32_t)cap; sym_units += cap / 16;
        size_t ocap = (64 + area * 28 / 4 + 15) & ~(size_t)15;
        b.out_off = (uint32_t)out_off; b.out_cap = (uint32_t)ocap; out_off += ocap;
    }
    CK(BL.alloc(sizeof(T1Block) * nblk)); CK(SYM.alloc(sym_units * 16 + 64)); CK(OUT.alloc(out_off));
    CK(PN.alloc((size_t)nblk * T1_MAX_PASSES * 4)); CK(RES.alloc((size_t)nblk * sizeof(T1Result)));
    CK(cudaMemcpy(BL.p, hb.data(), sizeof(T1Block) * nblk, cudaMemcpyHostToDevice));
    CK(cudaMemset(RES.p, 0, (size_t)nblk * sizeof(T1Result)));
    launch_t1_context(BL.as<T1Block>(), nblk, SYM.as<uint8_t>(), PN.as<uint32_t>(), RES.as<T1Result>(), 0);
    CK(cudaDeviceSynchronize());
    launch_t1_mq(BL.as<T1Block>(), nblk, SYM.as<uint8_t>(), PN.as<uint32_t>(), OUT.as<uint8_t>(), RES.as<T1Result>(), 32, nullptr, 0, 0);
    CK(cudaDeviceSynchronize());
    std::vector<T1Result> hr(nblk);
    std::vector<uint8_t> ho(out_off);
    CK(cudaMemcpy(hr.data(), RES.p, (size_t)nblk * sizeof(T1Result), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ho.data(), OUT.p, out_off, cudaMemcpyDeviceToHost));
    size_t off = 0;
    for (int i = 0; i < nblk; ++i) {
        if (hr[i].status) return set_error(J2KCUDA_ERR_CUDA, "tier-1 slot overflow");
        numbps[i] = hr[i].numbps; npasses[i] = hr[i].npasses;
        for (int k = 0; k < hr[i].npasses && k < 128; ++k) { rates[(size_t)i * 128 + k] = hr[i].rate[k]; nmsedec[(size_t)i * 128 + k] = hr[i].nmsedec[k]; }
        offsets[i] = off;
        if (off + hr[i].total > data_cap) return set_error(J2KCUDA_ERR_PARAM, "data buffer too small");
        memcpy(data + off, ho.data() + hb[i].out_off, hr[i].total);
        off += hr[i].total;
    }
    offsets[nblk] = off;
    return J2KCUDA_OK;
}); }

}  // extern "C"
