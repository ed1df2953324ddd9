// K-B placeholder: replaced by the real kernel in the next commit (entry points exist so the library links).
#include "tdod_internal.h"
extern "C" int64_t tdod_mask_rect_workspace_bytes(int64_t n_masks, int H, int W) { return 16; }
extern "C" int tdod_mask_rect(const uint8_t*, int64_t, int, int, float*, uint8_t*, void*, int64_t, void*) {
    return tdod_fail(TDOD_ERANGE, "tdod_mask_rect: not built yet");
}
