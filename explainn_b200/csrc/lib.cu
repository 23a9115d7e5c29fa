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
int num_sms() {
    static thread_local int cached = 0;
    if (cached > 0) return cached;
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
    cached = n;
    return n;
}
}  // namespace xnn

extern "C" int xnn_version(void) { return 100; }
extern "C" const char* xnn_last_error(void) { return xnn::g_err; }
extern "C" int xnn_sm_count(int* out) {
    XNN_CHECK_ARG(out != nullptr, "out is NULL");
    int dev = 0, n = 0;
    XNN_CUDA(cudaGetDevice(&dev));
    XNN_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    *out = n;
    return XNN_OK;
}
