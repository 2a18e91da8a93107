// capi.cu -- error reporting and device probing for libgs2d.so.
#include <cstring>

#include "gs2d_host.h"

namespace gs2d {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

}  // namespace gs2d

extern "C" {

int gs2d_version(void) { return 100; }

const char* gs2d_last_error(void) { return gs2d::g_err; }

int gs2d_device_arch(int device) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, device) != cudaSuccess) {
        (void)cudaGetLastError();
        return -1;
    }
    return p.major * 10 + p.minor;
}

}  // extern "C"
