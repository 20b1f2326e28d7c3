// Internal helpers shared by the translation units of libfwi_b200.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/fwi_b200.h"

namespace fwi {

void set_error(const char *fmt, ...);
void count_launch(int64_t n = 1);

#define FWI_CHECK_ARG(cond, ...)                                                        \
    do {                                                                                \
        if (!(cond)) {                                                                  \
            fwi::set_error(__VA_ARGS__);                                                \
            return FWI_ERR_ARG;                                                         \
        }                                                                               \
    } while (0)

#define FWI_CHECK_CUDA(expr)                                                            \
    do {                                                                                \
        cudaError_t e__ = (expr);                                                       \
        if (e__ != cudaSuccess) {                                                       \
            fwi::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__),     \
                           __FILE__, __LINE__);                                         \
            return FWI_ERR_CUDA;                                                        \
        }                                                                               \
    } while (0)

// Central-difference weights.  d2[h][k]: k = 0 is the centre weight of the 2nd derivative of
// accuracy 2h; d1[h][k-1] multiplies (f[+k] - f[-k]).
static const double kD2[5][5] = {
    {0, 0, 0, 0, 0},
    {-2.0, 1.0, 0, 0, 0},
    {-5.0 / 2.0, 4.0 / 3.0, -1.0 / 12.0, 0, 0},
    {-49.0 / 18.0, 3.0 / 2.0, -3.0 / 20.0, 1.0 / 90.0, 0},
    {-205.0 / 72.0, 8.0 / 5.0, -1.0 / 5.0, 8.0 / 315.0, -1.0 / 560.0}};
static const double kD1[5][4] = {
    {0, 0, 0, 0},
    {1.0 / 2.0, 0, 0, 0},
    {2.0 / 3.0, -1.0 / 12.0, 0, 0},
    {3.0 / 4.0, -3.0 / 20.0, 1.0 / 60.0, 0},
    {4.0 / 5.0, -1.0 / 5.0, 4.0 / 105.0, -1.0 / 280.0}};

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src, int src_bytes) {
    unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gmem_src),
                 "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

}  // namespace fwi
