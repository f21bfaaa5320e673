// One very large frame whose TILES are spread over several GPUs of the box, driven from one process
// (include/j2kcuda.h, "tiles of one frame on several GPUs"; BASELINE config C4: 8192x4320, 1024x1024 tiles).
//
// Tiles never read across a tile edge (tcd.cpp:508-526; each tile has its own opj_tcd_encode_tile), so the path shards
// with no exchange step: device k gets a contiguous run of tiles, encodes them with its own context and host
// thread and returns their tile-parts; the fragments are concatenated on the host behind one main header
// (j2kcuda_join_tile_streams) -- every byte of them went through host Tier-2 anyway, so no device-to-device copy or
// collective is involved.  Decode is the mirror: every device parses the main header, reads the packets of its own tiles
// only (routed by the Isot field of the SOT markers, j2k.cpp:9247-9333) and writes their rectangles into the caller's image.
#include "../../include/j2kcuda.h"
#include "common.h"
#include <string>
#include <thread>
#include <vector>
#include <cuda_runtime.h>
using namespace j2k;

struct j2kcuda_tiles {
    j2kcuda_params params;
    bool have_params = false;
    std::vector<int> devs;
    std::vector<j2kcuda_encoder*> enc;      // one per device, created on first use (tile range baked in)
    std::vector<j2kcuda_decoder*> dec;
    int ntiles = 0;
};

namespace {

// contiguous runs, the first (ntiles % n) devices take one tile more: the fragments concatenate in codestream order
void tile_range(int ntiles, int k, int n, uint32_t& first, uint32_t& count) {
    const int base = ntiles / n, extra = ntiles % n;
    first = (uint32_t)(k * base + (k < extra ? k : extra));
    count = (uint32_t)(base + (k < extra ? 1 : 0));
}

// run fn(k) for every device on its own thread; first failure wins, its message goes to the caller's error slot
template <class F>
int on_all_devices(int n, F fn) {
    std::vector<int> rc((size_t)n, 0);
    std::vector<std::string> msg((size_t)n);
    auto body = [&](int k) { rc[(size_t)k] = fn(k); if (rc[(size_t)k]) msg[(size_t)k] = j2kcuda_last_error(); };
    std::vector<std::thread> th;
    for (int k = 1; k < n; ++k) th.emplace_back(body, k);
    body(0);
    for (auto& t : th) t.join();
    for (int k = 0; k < n; ++k) if (rc[(size_t)k]) return set_error(rc[(size_t)k], msg[(size_t)k]);
    return J2KCUDA_OK;
}

}  // namespace

extern "C" {

j2kcuda_tiles* j2kcuda_tiles_create(const j2kcuda_params* p, const int* devices, int ndevices) { try {
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have == 0) { cudaGetLastError(); set_error(J2KCUDA_ERR_NO_DEVICE, "no CUDA device: libj2kcuda has no CPU path"); return nullptr; }
    if (!devices || ndevices < 1 || ndevices > 64) { set_error(J2KCUDA_ERR_PARAM, "bad device list"); return nullptr; }
    for (int i = 0; i < ndevices; ++i)
        if (devices[i] < 0 || devices[i] >= have) { set_error(J2KCUDA_ERR_PARAM, "device ordinal out of range"); return nullptr; }
    j2kcuda_tiles* t = new j2kcuda_tiles();
    t->devs.assign(devices, devices + ndevices);
    if (p) {
        t->params = *p; t->have_params = true;
        const uint32_t tw = p->tile_w ? p->tile_w : p->width, th = p->tile_h ? p->tile_h : p->height;
        if (!tw || !th) { delete t; set_error(J2KCUDA_ERR_PARAM, "bad image or tile size"); return nullptr; }
        t->ntiles = (int)(((p->width + tw - 1) / tw) * ((p->height + th - 1) / th));
    }
    return t;
} catch (...) { set_error(J2KCUDA_ERR_NOMEM, "j2kcuda_tiles_create: out of memory or internal error"); return nullptr; } }

void j2kcuda_tiles_destroy(j2kcuda_tiles* t) {
    if (!t) return;
    for (auto* e : t->enc) j2kcuda_encoder_destroy(e);
    for (auto* d : t->dec) j2kcuda_decoder_destroy(d);
    delete t;
}

int j2kcuda_tiles_encode(j2kcuda_tiles* t, const j2kcuda_plane* planes, int nplanes, uint8_t** out, size_t* out_len) { return j2k::guarded([&]() -> int {
    if (!t || !planes || !out || !out_len) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    if (!t->have_params) return set_error(J2KCUDA_ERR_PARAM, "created without encoder parameters");
    const int n = (int)t->devs.size() < t->ntiles ? (int)t->devs.size() : t->ntiles;   // never more devices than tiles
    if (n <= 1) {                                    // one device: an ordinary whole-frame encode
        if (t->enc.empty()) {
            j2kcuda_params p = t->params; p.device = t->devs[0]; p.tile_first = p.tile_count = 0;
            j2kcuda_encoder* e = j2kcuda_encoder_create(&p);
            if (!e) return J2KCUDA_ERR_PARAM;
            t->enc.push_back(e);
        }
        int rc = j2kcuda_encoder_upload(t->enc[0], planes, nplanes);
        if (!rc) rc = j2kcuda_encoder_run(t->enc[0]);
        if (!rc) rc = j2kcuda_encoder_finish(t->enc[0], out, out_len);
        return rc;
    }
    while ((int)t->enc.size() < n) {
        const int k = (int)t->enc.size();
        j2kcuda_params p = t->params;
        p.device = t->devs[(size_t)k];
        tile_range(t->ntiles, k, n, p.tile_first, p.tile_count);
        j2kcuda_encoder* e = j2kcuda_encoder_create(&p);
        if (!e) return J2KCUDA_ERR_PARAM;
        t->enc.push_back(e);
    }
    std::vector<uint8_t*> frag((size_t)n, nullptr);
    std::vector<size_t> len((size_t)n, 0);
    int rc = on_all_devices(n, [&](int k) -> int {
        j2kcuda_encoder* e = t->enc[(size_t)k];
        int r = j2kcuda_encoder_upload(e, planes, nplanes);          // only the rows of this device's tiles travel
        if (!r) r = j2kcuda_encoder_run(e);
        if (!r) r = j2kcuda_encoder_finish(e, &frag[(size_t)k], &len[(size_t)k]);
        return r;
    });
    if (!rc) rc = j2kcuda_join_tile_streams(&t->params, frag.data(), len.data(), n, out, out_len);
    for (auto* f : frag) j2kcuda_free(f);
    return rc;
}); }

int j2kcuda_tiles_decode(j2kcuda_tiles* t, const uint8_t* bytes, size_t nbytes, unsigned reduce, const j2kcuda_plane* planes, int nplanes) { return j2k::guarded([&]() -> int {
    if (!t || !bytes || !planes) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    j2kcuda_info info;
    if (int rc = j2kcuda_probe(bytes, nbytes, &info)) return rc;
    const int ntiles = (int)(info.tiles_x * info.tiles_y);
    const int n = (int)t->devs.size() < ntiles ? (int)t->devs.size() : ntiles;
    while ((int)t->dec.size() < (n < 1 ? 1 : n)) {
        j2kcuda_decoder* d = j2kcuda_decoder_create(t->devs[t->dec.size()]);
        if (!d) return J2KCUDA_ERR_NO_DEVICE;
        t->dec.push_back(d);
    }
    if (n <= 1) {
        decoder_set_tile_range(t->dec[0], 0, 0);
        int rc = j2kcuda_decoder_prepare(t->dec[0], bytes, nbytes, reduce);
        if (!rc) rc = j2kcuda_decoder_run(t->dec[0]);
        if (!rc) rc = j2kcuda_decoder_download(t->dec[0], planes, nplanes);
        return rc;
    }
    return on_all_devices(n, [&](int k) -> int {
        j2kcuda_decoder* d = t->dec[(size_t)k];
        uint32_t first, count;
        tile_range(ntiles, k, n, first, count);
        decoder_set_tile_range(d, first, count);
        int r = j2kcuda_decoder_prepare(d, bytes, nbytes, reduce);
        if (!r) r = j2kcuda_decoder_run(d);
        if (!r) r = j2kcuda_decoder_download(d, planes, nplanes);     // this device's tile rectangles only
        return r;
    });
}); }

}  // extern "C"
