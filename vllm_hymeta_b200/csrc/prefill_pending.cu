// Placeholder entry points, replaced as the tcgen05 prefill kernels land.
#include "common.cuh"
extern "C" int hymeta_gla_prefill_varlen(const void*, const void*, const void*, const void*, void*, void*,
    const void*, const int32_t*, const int32_t*, const int32_t*, int, int, int, int, int, int,
    int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t,
    int64_t, int64_t, int, float, int, int, void*) { return HYMETA_ERR_UNSUPPORTED; }
