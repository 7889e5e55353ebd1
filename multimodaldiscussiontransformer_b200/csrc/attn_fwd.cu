// placeholder until the tcgen05 kernel lands (replaced in the next commit)
#include "common.cuh"
using namespace mdt;
extern "C" int mdt_biased_attn_fwd(const void *, const void *, const void *, const mdt_attn_dims *, const void *, int,
                                   const uint8_t *, int, const float *, const uint8_t *, void *, float *, mdt_stream_t) {
    return fail(MDT_ENOSUP, "mdt_biased_attn_fwd: not built yet");
}
extern "C" int mdt_biased_attn_bwd(const void *, const void *, const void *, const void *, const void *, const float *,
                                   const mdt_attn_dims *, const void *, int, const uint8_t *, int, const float *,
                                   const uint8_t *, void *, void *, void *, const mdt_attn_dims *, float *, float *, void *,
                                   float *, mdt_stream_t) {
    return fail(MDT_ENOSUP, "mdt_biased_attn_bwd: not built yet");
}
