// mode-B raster placeholder: replaced by the real kernel in the next commit.
#include "tdod_internal.h"
extern "C" int tdod_score_segmentation(const uint8_t*, int, int, const float*, int64_t, int, float*, void*) {
    return tdod_fail(TDOD_ERANGE, "tdod_score_segmentation: not built yet");
}
