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

int fwi_version(void) { return 7; }

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

// ---- measurement aid: the HBM traffic of one forward step without the stencil -----------------
// um = 2u - um + w*u, q = u over n4 float4 (w indexed modulo wn4): reads u, um, w; writes um, q.
// Used by scripts/dev_stream.py to find what the memory system sustains for this read/write mix.
namespace {
__global__ void __launch_bounds__(256) stream_step_kernel(const float4 *__restrict__ u, float4 *__restrict__ um,
                                                          const float4 *__restrict__ w, float4 *__restrict__ q,
                                                          long long n4, long long wn4) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 a = __ldg(u + i), b = um[i], c = __ldg(w + i % wn4);
        float4 r;
        r.x = fmaf(c.x, a.x, 2.0f * a.x - b.x); r.y = fmaf(c.y, a.y, 2.0f * a.y - b.y);
        r.z = fmaf(c.z, a.z, 2.0f * a.z - b.z); r.w = fmaf(c.w, a.w, 2.0f * a.w - b.w);
        um[i] = r;
        __stcs(q + i, a);
    }
}
}  // namespace

extern "C" int fwi_debug_stream_step(const float *u, float *um, const float *w, float *q, int64_t n, int64_t wn,
                                     int blocks, void *stream) {
    stream_step_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const float4 *>(u), reinterpret_cast<float4 *>(um), reinterpret_cast<const float4 *>(w),
        reinterpret_cast<float4 *>(q), n / 4, wn / 4);
    FWI_CHECK_CUDA(cudaGetLastError());
    return FWI_OK;
}
