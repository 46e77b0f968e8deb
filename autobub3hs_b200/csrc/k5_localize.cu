// K5-K7 -- L3Localizer blob extraction (placeholder until the kernels land in this file).
#include "abub_internal.cuh"

int abub_loc_init(abub_ctx*) { return ABUB_OK; }
void abub_loc_free(abub_ctx*) {}

extern "C" int abub_localize(abub_ctx* ctx, int, int, int, int, int, int, abub_blob*, int, int*) {
    if (!ctx) return ABUB_E_INVALID;
    ctx->err = "abub_localize: not built yet";
    return ABUB_E_INVALID;
}
extern "C" int abub_localize_binary(abub_ctx* ctx, int, uint8_t*) {
    if (!ctx) return ABUB_E_INVALID;
    ctx->err = "abub_localize_binary: not built yet";
    return ABUB_E_INVALID;
}
