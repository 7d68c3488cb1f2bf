// placeholder until the verifier kernels land (next commit)
#include "ctx.h"
using namespace g2v;
extern "C" {
int gtsfm_ransac_samples(uint64_t, uint64_t, int, int, int32_t*) { return GTSFM_E_UNSUPPORTED; }
int gtsfm_verify_batch(gtsfm_ctx* c, const int32_t*, int, const uint32_t*, const int64_t*, const gtsfm_verify_params*, double*, double*,
                       int32_t*, int32_t*, uint8_t*, double*) { return fail(c, GTSFM_E_UNSUPPORTED, "verifier not built"); }
int gtsfm_verify_pair(gtsfm_ctx* c, const double*, int, const double*, int, const uint32_t*, int, const double*, const double*,
                      const gtsfm_verify_params*, double*, double*, int32_t*, int32_t*, uint8_t*, double*) { return fail(c, GTSFM_E_UNSUPPORTED, "verifier not built"); }
}
