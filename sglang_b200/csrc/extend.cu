// placeholder until the tcgen05 kernel lands (replaced in the next commit)
#include "common.cuh"
extern "C" int64_t b200attn_extend_workspace_bytes(int32_t, int64_t, int32_t) { return 0; }
extern "C" int b200attn_extend(const void*, int64_t, const void*, const void*, int64_t, int64_t, void*,
                               int64_t, float*, const int32_t*, const int32_t*, const int32_t*, int32_t,
                               int64_t, int32_t, int32_t, int32_t, int32_t, int32_t, int, int, float,
                               void*, int64_t, void*) {
  B2_UNSUPPORTED("extend: kernel not built yet");
}
