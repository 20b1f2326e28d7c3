// Error reporting, launch accounting and device probing shared by the C ABI entry points.
#include <stdarg.h>
#include <string.h>

#include <atomic>

#include "common.h"

namespace fwi {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void count_launch(int64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

}  // namespace fwi

extern "C" {

const char *fwi_last_error(void) { return fwi::g_err; }

int fwi_version(void) { return 4; }

int64_t fwi_kernel_launches(void) { return fwi::g_launches.load(std::memory_order_relaxed); }

int fwi_device_ok(int device) {
    cudaDeviceProp prop;
    FWI_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        fwi::set_error("device %d is sm_%d%d; this library is built for sm_100a only", device,
                       prop.major, prop.minor);
        return FWI_ERR_NODEVICE;
    }
    return FWI_OK;
}

}  // extern "C"
