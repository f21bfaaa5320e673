// Entry points of include/j2kcuda.h whose device path is not built yet: they fail loudly (never a
// CPU fallback).  Each is replaced by its implementation as the decode path lands.
#include "../../include/j2kcuda.h"
#include "common.h"
using namespace j2k;
static int pending(const char* what) { return set_error(J2KCUDA_ERR_UNSUPPORTED, std::string(what) + ": not implemented in this build"); }
extern "C" {
int j2kcuda_encode_batch(const j2kcuda_params* p, const j2kcuda_plane* const* frames, int nframes, int nplanes, uint8_t** outs, size_t* out_lens) {
    if (!p || !frames || !outs || !out_lens) return set_error(J2KCUDA_ERR_PARAM, "null argument");
    j2kcuda_encoder* e = j2kcuda_encoder_create(p);
    if (!e) return J2KCUDA_ERR_PARAM;
    int rc = 0;
    for (int i = 0; i < nframes && !rc; ++i) {
        rc = j2kcuda_encoder_upload(e, frames[i], nplanes);
        if (!rc) rc = j2kcuda_encoder_run(e);
        if (!rc) rc = j2kcuda_encoder_finish(e, &outs[i], &out_lens[i]);
    }
    j2kcuda_encoder_destroy(e);
    return rc;
}
int j2kcuda_decode_batch(const uint8_t* const*, const size_t*, int, unsigned, const j2kcuda_plane* const*, int) { return pending("j2kcuda_decode_batch"); }
int j2kcuda_probe(const uint8_t*, size_t, j2kcuda_info*) { return pending("j2kcuda_probe"); }
int j2kcuda_decode(const uint8_t*, size_t, unsigned, const j2kcuda_plane*, int, j2kcuda_abort_fn, void*) { return pending("j2kcuda_decode"); }
j2kcuda_decoder* j2kcuda_decoder_create(int) { pending("j2kcuda_decoder_create"); return nullptr; }
void j2kcuda_decoder_destroy(j2kcuda_decoder*) {}
int j2kcuda_decoder_prepare(j2kcuda_decoder*, const uint8_t*, size_t, unsigned) { return pending("j2kcuda_decoder_prepare"); }
int j2kcuda_decoder_run(j2kcuda_decoder*) { return pending("j2kcuda_decoder_run"); }
int j2kcuda_decoder_download(j2kcuda_decoder*, const j2kcuda_plane*, int) { return pending("j2kcuda_decoder_download"); }
int j2kcuda_decoder_get_device_output(j2kcuda_decoder*, void**, int64_t*, int*) { return pending("j2kcuda_decoder_get_device_output"); }
int j2kcuda_decoder_timings(j2kcuda_decoder*, float*, int) { return 0; }
int j2kcuda_decoder_counters(j2kcuda_decoder*, uint64_t*, int) { return 0; }
int j2kcuda_stage_t1_decode(const uint8_t*, const int32_t*, int, int32_t*) { return pending("j2kcuda_stage_t1_decode"); }
int j2kcuda_t2_parse(const uint8_t*, size_t, unsigned, int32_t**, int*, uint8_t**, size_t*) { return pending("j2kcuda_t2_parse"); }
}
