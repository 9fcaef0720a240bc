// libbvg.so: diagnostics, thread-local error state, launch accounting.
#include "bvg_common.cuh"

namespace bvg {
static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
void count_launch(int n) { g_launches += n; }
}  // namespace bvg

extern "C" const char* bvg_last_error(void) { return bvg::g_err; }
extern "C" int bvg_version(void) { return BVG_VERSION; }
extern "C" int64_t bvg_launch_count(void) { return bvg::g_launches; }
extern "C" void bvg_reset_launch_count(void) { bvg::g_launches = 0; }

extern "C" int bvg_device_check(int device) {
    if (device < 0) BVG_CUDA(cudaGetDevice(&device));
    int major = 0, minor = 0;
    BVG_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    BVG_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
    if (major != 10)
        return bvg::set_error(BVG_ERR_DEVICE,
                              "device %d has compute capability %d.%d; libbvg.so contains sm_100a code only "
                              "(tcgen05/TMEM/TMA) and has no fallback path",
                              device, major, minor);
    return BVG_OK;
}
