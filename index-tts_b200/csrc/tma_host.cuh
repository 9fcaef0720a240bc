// Host-side cuTensorMapEncodeTiled access without linking libcuda: the entry point is fetched through the runtime.
#pragma once
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>

namespace bvg {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess) return nullptr;
        return reinterpret_cast<EncodeTiledFn>(f);
    }();
    return fn;
}

// Programmatic dependent launch for the persistent kernels of the forward plan (BVG_PDL=0 switches it off for A/B runs).
static inline bool pdl_enabled() {
    static const bool v = []() {
        const char* e = getenv("BVG_PDL");
        return !(e != nullptr && e[0] == '0');
    }();
    return v;
}

static inline int device_sm_count() {
    static thread_local int cached_dev = -1, cached = 0;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev != cached_dev) {
        cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
        cached_dev = dev;
    }
    return cached;
}

}  // namespace bvg
