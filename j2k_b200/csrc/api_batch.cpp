// Sequence encode on this process's device (frames of an image sequence, include/j2kcuda.h).
#include "../../include/j2kcuda.h"
#include "common.h"
using namespace j2k;
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
}
