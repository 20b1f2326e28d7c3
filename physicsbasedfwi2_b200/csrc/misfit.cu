// Trace misfit fused with its own backward: the step on the receiver side of the propagator in
// every physics-guided training iteration of the reference (models/networks.py:5467-5476 and the
// backward the autograd engine runs for it at :5491):
//
//     d     = pred - direct                                (direct-wave subtraction, :5467)
//     m     = max_t |d|            per trace (shot, receiver) (:5468)
//     n     = d / (m + 1e-10)                              (:5470)
//     loss  = mean |n - obs|   (L1Loss, :5476)   or   mean (n - obs)^2  (MSELoss)
//     dloss/dpred                                           (what lossinner.backward() feeds the adjoint)
//
// In the reference this is ~10 separate torch kernels forward and as many backward, each a full
// pass over rec[nt, S, R] (1.8 GB per GPU batch at BASELINE config 2).  Here: three passes over
// the traces (max; sums; gradient) and a final reduction, no atomics, fixed summation order, so
// the result is deterministic.  HBM-bound, no tensor cores.
//
// Layout: [nt][M] with M = S*R traces contiguous; a thread owns one trace inside one time chunk,
// so a warp reads 32 adjacent traces per time sample (coalesced).
#include <algorithm>
#include <mutex>

#include "common.h"

namespace {

struct MParams {
    int nt, M, C, chunk;        // time samples, traces, time chunks, samples per chunk
    int kind, normalize;        // 0 = L1, 1 = MSE
    float eps, inv_n;           // 1e-10, 1 / (nt * M)
    const float *scale;         // upstream gradient of the loss (device scalar) or NULL = 1
    const float *pred, *direct, *obs;
    float *pmax, *pacc, *pl;    // [C][M] partials
    int *parg;                  // [C][M] first index of the chunk maximum
    float *grad;                // [nt][M]
    float *loss;                // [1]
};

__device__ __forceinline__ float sgn(float x) { return (x > 0.f) - (x < 0.f); }

__global__ void __launch_bounds__(128) misfit_max(const MParams p) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x, c = blockIdx.y;
    if (m >= p.M) return;
    const int t0 = c * p.chunk, t1 = min(p.nt, t0 + p.chunk);
    float best = -1.f;
    int arg = t0;
    for (int t = t0; t < t1; ++t) {
        const size_t i = (size_t)t * p.M + m;
        float d = __ldg(p.pred + i);
        if (p.direct) d = d - __ldg(p.direct + i);
        const float a = fabsf(d);
        if (a > best) { best = a; arg = t; }
    }
    p.pmax[(size_t)c * p.M + m] = best;
    p.parg[(size_t)c * p.M + m] = arg;
}

// trace maximum and its first index from the chunk partials (every chunk's threads redo this small loop)
__device__ __forceinline__ void trace_max(const MParams &p, int m, float &mx, int &arg) {
    mx = -1.f; arg = 0;
    for (int c = 0; c < p.C; ++c) {
        const float v = p.pmax[(size_t)c * p.M + m];
        if (v > mx) { mx = v; arg = p.parg[(size_t)c * p.M + m]; }
    }
}

template <bool GRAD>
__global__ void __launch_bounds__(128) misfit_pass(const MParams p) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x, c = blockIdx.y;
    if (m >= p.M) return;
    float mx = 1.f, inv = 1.f, A = 0.f;
    int arg = -1;
    const float scale = (GRAD && p.scale) ? __ldg(p.scale) : 1.f;
    if (p.normalize) {
        trace_max(p, m, mx, arg);
        inv = 1.0f / (mx + p.eps);
        if (GRAD)
            for (int k = 0; k < p.C; ++k) A = A + p.pacc[(size_t)k * p.M + m];
    }
    const int t0 = c * p.chunk, t1 = min(p.nt, t0 + p.chunk);
    float l = 0.f, acc = 0.f;
    for (int t = t0; t < t1; ++t) {
        const size_t i = (size_t)t * p.M + m;
        float d = __ldg(p.pred + i);
        if (p.direct) d = d - __ldg(p.direct + i);
        const float n = d * inv;
        const float r = n - __ldg(p.obs + i);
        const float gn = p.kind == 0 ? sgn(r) * p.inv_n : (2.0f * r) * p.inv_n;
        if (!GRAD) {
            l = l + (p.kind == 0 ? fabsf(r) : r * r);
            acc = acc + gn * d;
        } else {
            float g = gn * inv;
            if (t == arg) g = g - sgn(d) * (A * inv) * inv;      // through m = |d[t*]|
            __stcs(p.grad + i, scale * g);
        }
    }
    if (!GRAD) {
        p.pl[(size_t)c * p.M + m] = l;
        p.pacc[(size_t)c * p.M + m] = acc;
    }
}

// one block: fixed-order sum of the [C][M] loss partials
__global__ void __launch_bounds__(1024) misfit_finish(const MParams p) {
    __shared__ float sh[1024];
    const size_t n = (size_t)p.C * p.M;
    float s = 0.f;
    for (size_t i = threadIdx.x; i < n; i += blockDim.x) s = s + p.pl[i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int w = 512; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) sh[threadIdx.x] = sh[threadIdx.x] + sh[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) p.loss[0] = sh[0] * p.inv_n;
}

int chunks_for(int nt, int M) {
    const int bm = (M + 127) / 128;
    int C = (4 * 148 + bm - 1) / bm;
    C = std::max(1, std::min(C, 64));
    return std::min(C, std::max(1, nt));
}

}  // namespace

extern "C" int64_t fwi_trace_misfit_workspace_bytes(const fwi_misfit_desc *d) {
    if (!d || d->nt < 1 || d->n_traces < 1) return -1;
    return 4ll * chunks_for(d->nt, d->n_traces) * d->n_traces * 4;
}

static int fill(MParams &p, const fwi_misfit_desc *d, const float *pred, const float *direct, const float *obs,
                void *workspace) {
    FWI_CHECK_ARG(d != nullptr, "descriptor is NULL");
    FWI_CHECK_ARG(d->nt >= 1 && d->n_traces >= 1, "nt and n_traces must be >= 1");
    FWI_CHECK_ARG(d->kind == 0 || d->kind == 1, "kind must be 0 (L1) or 1 (MSE)");
    FWI_CHECK_ARG(pred && obs && workspace, "pred, obs and workspace must not be NULL");
    p.nt = d->nt; p.M = d->n_traces;
    p.C = chunks_for(p.nt, p.M);
    p.chunk = (p.nt + p.C - 1) / p.C;
    p.kind = d->kind; p.normalize = d->normalize;
    p.eps = d->eps; p.inv_n = (float)(1.0 / ((double)p.nt * p.M)); p.scale = nullptr;
    p.pred = pred; p.direct = direct; p.obs = obs;
    const size_t cm = (size_t)p.C * p.M;
    p.pmax = static_cast<float *>(workspace);
    p.pacc = p.pmax + cm;
    p.pl = p.pacc + cm;
    p.parg = reinterpret_cast<int *>(p.pl + cm);
    p.grad = nullptr; p.loss = nullptr;
    return FWI_OK;
}

extern "C" int fwi_trace_misfit(const fwi_misfit_desc *d, const float *pred, const float *direct, const float *obs,
                                float *loss, void *workspace, void *stream) {
    MParams p;
    int rc = fill(p, d, pred, direct, obs, workspace);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(loss != nullptr, "loss must not be NULL");
    p.loss = loss;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const dim3 grid((p.M + 127) / 128, p.C), block(128);
    if (p.normalize) { misfit_max<<<grid, block, 0, st>>>(p); fwi::count_launch(); }
    misfit_pass<false><<<grid, block, 0, st>>>(p);
    misfit_finish<<<1, 1024, 0, st>>>(p);
    fwi::count_launch(2);
    FWI_CHECK_CUDA(cudaGetLastError());
    return FWI_OK;
}

extern "C" int fwi_trace_misfit_backward(const fwi_misfit_desc *d, const float *pred, const float *direct,
                                         const float *obs, const float *grad_loss, float *grad_pred, void *workspace,
                                         void *stream) {
    MParams p;
    int rc = fill(p, d, pred, direct, obs, workspace);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(grad_pred != nullptr, "grad_pred must not be NULL");
    p.grad = grad_pred; p.scale = grad_loss;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const dim3 grid((p.M + 127) / 128, p.C), block(128);
    misfit_pass<true><<<grid, block, 0, st>>>(p);
    fwi::count_launch();
    FWI_CHECK_CUDA(cudaGetLastError());
    return FWI_OK;
}

// ---- allocation-free variant of fwi_trace_misfit --------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot allocate device
// memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The [4][C][M] partials
// that fwi_trace_misfit keeps in a caller-provided workspace live in a file-scope device array of
// FWI_TRACE_MISFIT_STATIC_MAX_BYTES (4 MiB; every GPU that loads the library holds it until exit).  All calls share the
// array: it keeps the previous call's contents (every element the kernels read is written first by the same call, so
// nothing has to start at zero and nothing is cleared), and two calls that overlap on one GPU race.
namespace {

static __device__ __align__(256) unsigned char g_misfit_scratch[FWI_TRACE_MISFIT_STATIC_MAX_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mutex;
void *g_scratch_addr[kMaxDevices];      // device address of g_misfit_scratch on device i (0 = not looked up yet)

// In host code `g_misfit_scratch` is a host-side stand-in, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, once per device, with that device current.
int misfit_scratch(void **ptr) {
    int dev = -1;
    FWI_CHECK_CUDA(cudaGetDevice(&dev));
    FWI_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "device ordinal %d out of range (max %d)", dev, kMaxDevices);
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_addr[dev]) {
        void *p = nullptr;
        FWI_CHECK_CUDA(cudaGetSymbolAddress(&p, g_misfit_scratch));
        FWI_CHECK_ARG(p != nullptr, "cudaGetSymbolAddress returned NULL for the static misfit scratch");
        g_scratch_addr[dev] = p;
    }
    *ptr = g_scratch_addr[dev];
    return FWI_OK;
}

}  // namespace

extern "C" int fwi_trace_misfit_noalloc(const fwi_misfit_desc *d, const float *pred, const float *direct, const float *obs,
                                        float *loss, void *stream) {
    FWI_CHECK_ARG(d != nullptr, "descriptor is NULL");
    const int64_t need = fwi_trace_misfit_workspace_bytes(d);
    FWI_CHECK_ARG(need > 0, "nt and n_traces must be >= 1");
    FWI_CHECK_ARG(need <= (int64_t)FWI_TRACE_MISFIT_STATIC_MAX_BYTES,
                  "fwi_trace_misfit_noalloc: nt = %d, n_traces = %d need %lld bytes of scratch, the static array holds %lld "
                  "(use fwi_trace_misfit with a caller-provided workspace)",
                  d->nt, d->n_traces, (long long)need, (long long)FWI_TRACE_MISFIT_STATIC_MAX_BYTES);
    void *scratch = nullptr;
    int rc = misfit_scratch(&scratch);
    if (rc != FWI_OK) return rc;
    return fwi_trace_misfit(d, pred, direct, obs, loss, scratch, stream);
}
