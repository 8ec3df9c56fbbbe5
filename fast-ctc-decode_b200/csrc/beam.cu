// TEMPORARY: beam-search entry points land in the next commit.
#include "../../include/fctc.h"
extern "C" {
int64_t fctc_beam_scratch_bytes(const fctc_handle*, int64_t, int64_t, int32_t, int32_t, int32_t) { return 0; }
int fctc_beam_search_batch_device(fctc_handle*, const float*, int64_t, int64_t, int32_t, int32_t, float, int32_t, uint8_t*,
                                  int32_t*, int64_t, int64_t*, int32_t*, int32_t*, unsigned long long*, void*, int64_t,
                                  int32_t, void*) { return FCTC_E_UNSUPPORTED; }
int fctc_beam_search_batch(fctc_handle*, const float*, int64_t, int64_t, int32_t, int32_t, float, int32_t, uint8_t*,
                           int32_t*, int64_t, int64_t*, int32_t*, int32_t*, int64_t*) { return FCTC_E_UNSUPPORTED; }
int fctc_beam_search(fctc_handle*, const float*, int64_t, int32_t, int64_t, int64_t, int32_t, float, int32_t, uint8_t*,
                     int64_t*, int64_t, int64_t*) { return FCTC_E_UNSUPPORTED; }
}
