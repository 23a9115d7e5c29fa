// library-level entry points: version, last error, SM count
#include "common.cuh"
#include <stdarg.h>

namespace xnn {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
// SM count of the CURRENT device (cached per device), minus the SMs the host asked to leave free for a
// concurrent collective (xnn_set_sm_reserve): every persistent kernel sizes its grid from this.
static int g_sm_count[64] = {0};
static int g_sm_reserve = 0;
int num_sms() {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev >= 0 && dev < 64 && g_sm_count[dev] > 0) n = g_sm_count[dev];
    else {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
        if (dev >= 0 && dev < 64) g_sm_count[dev] = n;
    }
    const int r = g_sm_reserve;
    return (r > 0 && r < n) ? n - r : n;
}
}  // namespace xnn

extern "C" int xnn_version(void) { return 100; }
extern "C" const char* xnn_last_error(void) { return xnn::g_err; }
extern "C" int xnn_set_sm_reserve(int n) {
    XNN_CHECK_ARG(n >= 0 && n < 64, "reserve must be in [0, 64)");
    xnn::g_sm_reserve = n;
    return XNN_OK;
}
extern "C" int xnn_kernel_size_limit(int path) { return path == 0 ? 48 : 32; }
extern "C" int xnn_sm_count(int* out) {
    XNN_CHECK_ARG(out != nullptr, "out is NULL");
    int dev = 0, n = 0;
    XNN_CUDA(cudaGetDevice(&dev));
    XNN_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    *out = n;
    return XNN_OK;
}
