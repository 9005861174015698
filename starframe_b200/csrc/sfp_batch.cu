// placeholder — replaced below
#include "sfp.h"
extern "C" {
int32_t sfp_batch_create(const SfpParams*, uint32_t, uint32_t, SfpBatch** out) { if (out) *out = nullptr; return SFP_E_INVALID; }
int32_t sfp_batch_destroy(SfpBatch*) { return SFP_OK; }
int32_t sfp_batch_upload(SfpBatch*, const SfpBodyDesc*) { return SFP_E_INVALID; }
int32_t sfp_batch_step(SfpBatch*, float, uint32_t) { return SFP_E_INVALID; }
int32_t sfp_batch_sync(SfpBatch*) { return SFP_E_INVALID; }
int32_t sfp_batch_download(SfpBatch*, SfpBodyState*) { return SFP_E_INVALID; }
int32_t sfp_batch_stats(SfpBatch*, SfpStats*) { return SFP_E_INVALID; }
const char* sfp_batch_last_error(SfpBatch*) { return "batch not built yet"; }
int32_t sfp_batch_last_step_ms(SfpBatch*, float*) { return SFP_E_INVALID; }
}
