// deepmind coarse/fine variant (models/deepmind_version.py:75-165) — placeholder until the kernel lands.
#include "internal.h"

struct wrnn_dm_handle { std::string err; };
static thread_local std::string g_dm_err;

extern "C" {
int wrnn_dm_create(int32_t, int32_t, int, wrnn_dm_handle **out) {
    if (out) *out = nullptr;
    g_dm_err = "deepmind variant not built yet";
    return WRNN_ERR_STATE;
}
void wrnn_dm_destroy(wrnn_dm_handle *h) { delete h; }
const char *wrnn_dm_last_error(const wrnn_dm_handle *h) { return h ? h->err.c_str() : g_dm_err.c_str(); }
int wrnn_dm_load_weights(wrnn_dm_handle *, const wrnn_tensor *, int32_t) { return WRNN_ERR_STATE; }
int wrnn_dm_generate(wrnn_dm_handle *, int64_t, int32_t, const float *, const float *, uint64_t, int64_t *, int64_t *,
                     int64_t *, void *) { return WRNN_ERR_STATE; }
int wrnn_dm_last_timing(wrnn_dm_handle *, float *, int64_t *) { return WRNN_ERR_STATE; }
}
