// Image sequences on this process's device (include/j2kcuda.h, "sequences").
//
// Frames of a sequence are independent, so a sequence object keeps a small pool of
// encoder / decoder contexts, each with its own CUDA stream and buffers, and one host worker per
// context.  Workers pull frame indices from a shared counter and take a frame through all of its
// stages (upload, device stages, result D2H, host Tier-2 -- or the mirror for decode).  While one
// worker sits in host Tier-2, the kernels of the other workers' frames run, and the latency-bound
// Tier-1 kernels of different frames share the SMs.  No frame ever waits for another one.
#include "../../include/j2kcuda.h"
#include "common.h"
#include <atomic>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include <cuda_runtime.h>
using namespace j2k;

struct j2kcuda_sequence {
    int device = -1;
    int nworkers = 0;
    std::vector<int> devs;            // device of worker w (one entry per worker; all equal for a single-GPU sequence)
    bool have_params = false;
    j2kcuda_params params;
    std::vector<j2kcuda_encoder*> enc;
    std::vector<j2kcuda_decoder*> dec;
    // running totals over all frames since the last j2kcuda_sequence_stats(reset)
    std::mutex mu;
    double enc_ms[8] = {0}, dec_ms[8] = {0};
    uint64_t enc_cnt[6] = {0}, dec_cnt[6] = {0};
    uint64_t enc_frames = 0, dec_frames = 0;
};

namespace {

// run fn(worker, frame) for every frame on `nw` threads; returns the first non-zero status and
// hands its message to the calling thread's error slot
template <class F>
int run_workers(int nw, int nframes, F fn) {
    std::atomic<int> next{0};
    std::atomic<int> status{0};
    std::mutex mu;
    std::string msg;
    auto body = [&](int w) {
        for (;;) {
            if (status.load(std::memory_order_relaxed)) return;
            const int i = next.fetch_add(1);
            if (i >= nframes) return;
            const int rc = fn(w, i);
            if (rc) {
                std::lock_guard<std::mutex> g(mu);
                if (!status.load()) { status.store(rc); msg = j2kcuda_last_error(); }
                return;
            }
        }
    };
    if (nw > nframes) nw = nframes;
    if (nw <= 1) body(0);
    else {
        std::vector<std::thread> th;
        for (int w = 1; w < nw; ++w) th.emplace_back(body, w);
        body(0);
        for (auto& t : th) t.join();
    }
    const int rc = status.load();
    if (rc) return set_error(rc, msg);
    return J2KCUDA_OK;
}

}  // namespace

extern "C" {

j2kcuda_sequence* j2kcuda_sequence_create(const j2kcuda_params* p, int device, int nworkers) { try {
    if (nworkers < 1) nworkers = 4;
    if (nworkers > 16) nworkers = 16;
    j2kcuda_sequence* s = new j2kcuda_sequence();
    s->nworkers = nworkers;
    s->device = device;
    if (p) {
        s->have_params = true;
        s->params = *p;
        if (device >= 0) s->params.device = device;
        else s->device = p->device;
    }
    s->devs.assign((size_t)nworkers, s->device);
    return s;
} catch (...) { set_error(J2KCUDA_ERR_NOMEM, "j2kcuda_sequence_create: out of memory or internal error"); return nullptr; } }

// Frames of a sequence over SEVERAL GPUs of the box from one process (what a C++ host such as the plug-in can call;
// bench.py uses one process per GPU instead).  Workers are spread round-robin over the devices and pull frames from one
// shared counter, so a faster or less loaded GPU simply takes more frames; nothing is exchanged between devices.
j2kcuda_sequence* j2kcuda_sequence_create_multi(const j2kcuda_params* p, const int* devices, int ndevices, int workers_per_device) { try {
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have == 0) { cudaGetLastError(); set_error(J2KCUDA_ERR_NO_DEVICE, "no CUDA device: libj2kcuda has no CPU path"); return nullptr; }
    if (!devices || ndevices < 1 || ndevices > 64) { set_error(J2KCUDA_ERR_PARAM, "bad device list"); return nullptr; }
    for (int i = 0; i < ndevices; ++i)
        if (devices[i] < 0 || devices[i] >= have) { set_error(J2KCUDA_ERR_PARAM, "device ordinal out of range"); return nullptr; }
    if (workers_per_device < 1) workers_per_device = 4;
    if (workers_per_device > 16) workers_per_device = 16;
    j2kcuda_sequence* s = new j2kcuda_sequence();
    s->nworkers = ndevices * workers_per_device;
    s->device = devices[0];
    if (p) { s->have_params = true; s->params = *p; }
    for (int w = 0; w < s->nworkers; ++w) s->devs.push_back(devices[w % ndevices]);
    return s;
} catch (...) { set_error(J2KCUDA_ERR_NOMEM, "j2kcuda_sequence_create_multi: out of memory or internal error"); return nullptr; } }

void j2kcuda_sequence_destroy(j2kcuda_sequence* s) {
    if (!s) return;
    for (auto* e : s->enc) j2kcuda_encoder_destroy(e);
    for (auto* d : s->dec) j2kcuda_decoder_destroy(d);
    delete s;
}

int j2kcuda_sequence_encode(j2kcuda_sequence* s, const j2kcuda_plane* const* frames, int nframes, int nplanes,
                            unsigned flags, uint8_t** outs, size_t* out_lens) { return j2k::guarded([&]() -> int {
    if (!s || !frames || !outs || !out_lens) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    if (!s->have_params) return set_error(J2KCUDA_ERR_PARAM, "sequence was created without encoder parameters");
    if (nframes <= 0) return J2KCUDA_OK;
    const int nw = s->nworkers < nframes ? s->nworkers : nframes;
    while ((int)s->enc.size() < nw) {            // contexts are created once and reused by later calls
        j2kcuda_params pw = s->params;
        pw.device = s->devs[s->enc.size()];
        j2kcuda_encoder* e = j2kcuda_encoder_create(&pw);
        if (!e) return J2KCUDA_ERR_PARAM;        // message already set
        encoder_set_throughput_mode(e, s->nworkers > 2);
        s->enc.push_back(e);
    }
    for (int i = 0; i < nframes; ++i) { outs[i] = nullptr; out_lens[i] = 0; }
    return run_workers(nw, nframes, [&](int w, int i) -> int {
        j2kcuda_encoder* e = s->enc[w];
        int rc;
        if (flags & J2KCUDA_SEQ_DEVICE_PIXELS)
            rc = j2kcuda_encoder_set_device_input(e, frames[i][0].buf, frames[i][0].rowbytes, frames[i][0].sample_type);
        else
            rc = j2kcuda_encoder_upload(e, frames[i], nplanes);
        if (!rc) rc = j2kcuda_encoder_run(e);
        if (!rc) rc = j2kcuda_encoder_finish(e, &outs[i], &out_lens[i]);
        if (!rc) {
            float ms[8]; uint64_t cnt[6];
            j2kcuda_encoder_timings(e, ms, 8); j2kcuda_encoder_counters(e, cnt, 6);
            std::lock_guard<std::mutex> g(s->mu);
            for (int k = 0; k < 8; ++k) s->enc_ms[k] += ms[k];
            for (int k = 0; k < 6; ++k) s->enc_cnt[k] += cnt[k];
            ++s->enc_frames;
        }
        return rc;
    });
}); }

int j2kcuda_sequence_decode(j2kcuda_sequence* s, const uint8_t* const* streams, const size_t* lens, int nframes,
                            unsigned reduce, const j2kcuda_plane* const* frames, int nplanes, unsigned flags) { return j2k::guarded([&]() -> int {
    if (!s || !streams || !lens || !frames) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    if (nframes <= 0) return J2KCUDA_OK;
    const int nw = s->nworkers < nframes ? s->nworkers : nframes;
    while ((int)s->dec.size() < nw) {
        j2kcuda_decoder* d = j2kcuda_decoder_create(s->devs[s->dec.size()]);
        if (!d) return J2KCUDA_ERR_NO_DEVICE;
        s->dec.push_back(d);
    }
    return run_workers(nw, nframes, [&](int w, int i) -> int {
        j2kcuda_decoder* d = s->dec[w];
        int rc = j2kcuda_decoder_prepare(d, streams[i], lens[i], reduce);
        if (!rc) rc = j2kcuda_decoder_run(d);
        if (!rc) {
            if (flags & J2KCUDA_SEQ_DEVICE_PIXELS) rc = j2kcuda_decoder_copy_to_device(d, frames[i][0].buf, frames[i][0].rowbytes);
            else rc = j2kcuda_decoder_download(d, frames[i], nplanes);
        }
        if (!rc) {
            float ms[8] = {0}; uint64_t cnt[6];
            j2kcuda_decoder_timings(d, ms, 7); j2kcuda_decoder_counters(d, cnt, 6);
            std::lock_guard<std::mutex> g(s->mu);
            for (int k = 0; k < 8; ++k) s->dec_ms[k] += ms[k];
            for (int k = 0; k < 6; ++k) s->dec_cnt[k] += cnt[k];
            ++s->dec_frames;
        }
        return rc;
    });
}); }

int j2kcuda_sequence_stats(j2kcuda_sequence* s, double* enc_ms, uint64_t* enc_cnt, double* dec_ms, uint64_t* dec_cnt,
                           uint64_t* frames, int reset) { return j2k::guarded([&]() -> int {
    if (!s) return set_error(J2KCUDA_ERR_PARAM, "null sequence");
    std::lock_guard<std::mutex> g(s->mu);
    for (int k = 0; k < 8; ++k) { if (enc_ms) enc_ms[k] = s->enc_ms[k]; if (dec_ms) dec_ms[k] = s->dec_ms[k]; }
    for (int k = 0; k < 6; ++k) { if (enc_cnt) enc_cnt[k] = s->enc_cnt[k]; if (dec_cnt) dec_cnt[k] = s->dec_cnt[k]; }
    if (frames) { frames[0] = s->enc_frames; frames[1] = s->dec_frames; }
    if (reset) {
        for (int k = 0; k < 8; ++k) s->enc_ms[k] = s->dec_ms[k] = 0;
        for (int k = 0; k < 6; ++k) s->enc_cnt[k] = s->dec_cnt[k] = 0;
        s->enc_frames = s->dec_frames = 0;
    }
    return J2KCUDA_OK;
}); }

int j2kcuda_encode_batch(const j2kcuda_params* p, const j2kcuda_plane* const* frames, int nframes, int nplanes,
                         uint8_t** outs, size_t* out_lens) { return j2k::guarded([&]() -> int {
    if (!p || !frames || !outs || !out_lens) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    j2kcuda_sequence* s = j2kcuda_sequence_create(p, p->device, 0);
    int rc = j2kcuda_sequence_encode(s, frames, nframes, nplanes, 0, outs, out_lens);
    j2kcuda_sequence_destroy(s);
    return rc;
}); }

int j2kcuda_decode_batch(const uint8_t* const* streams, const size_t* lens, int nframes, unsigned reduce,
                         const j2kcuda_plane* const* frames, int nplanes) { return j2k::guarded([&]() -> int {
    if (!streams || !lens || !frames) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    j2kcuda_sequence* s = j2kcuda_sequence_create(nullptr, -1, 0);
    int rc = j2kcuda_sequence_decode(s, streams, lens, nframes, reduce, frames, nplanes, 0);
    j2kcuda_sequence_destroy(s);
    return rc;
}); }

}  // extern "C"
