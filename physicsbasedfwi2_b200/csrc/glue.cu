// The steps either side of the elastic gradient in the reference's `prop` methods, as kernels:
//   * the frequency stage of d.add_fwi_stage(fc_low=0.0, fc_high=10.0, ...) (models/networks.py:9147): a Butterworth
//     low-pass applied to every trace of the recorded / synthetic seismograms (zero-phase: forward then backward in time);
//   * the gradient post-processing of models/networks.py:9194-9240: depth weight z^1.5 on the density gradient, zeroed
//     water rows, per-parameter scaling max|model| / max|gradient|, the 0.5 factor and 1e-10 offset of rho - fused into
//     two passes over the three gradients instead of ~30 numpy / torch statements.
// (DENISE's file protocol - SU seismograms, binary models - is replaced by tensors; physicsbasedfwi2_b200/denise_glue.py.)
#include <math.h>

#include <algorithm>

#include "common.h"

namespace {

constexpr int kMaxSections = 8;

struct Biquads {
    int n;
    double b0[kMaxSections], b1[kMaxSections], b2[kMaxSections], a1[kMaxSections], a2[kMaxSections];
};

// Butterworth low-pass of even order as second-order sections (bilinear transform with pre-warping):
// analog section s^2 + 2 zeta_k s + 1, zeta_k = sin(pi (2k + 1) / (2 N)); K = tan(pi fc dt)
int design_lowpass(int order, double fc, double dt, Biquads *q) {
    FWI_CHECK_ARG(order >= 2 && order % 2 == 0 && order / 2 <= kMaxSections, "Butterworth order must be even, 2..%d (got %d)",
                  2 * kMaxSections, order);
    FWI_CHECK_ARG(dt > 0 && fc > 0 && fc < 0.5 / dt, "cut-off %g Hz must lie in (0, Nyquist = %g Hz)", fc, 0.5 / dt);
    const double K = tan(M_PI * fc * dt);
    q->n = order / 2;
    for (int k = 0; k < q->n; ++k) {
        const double zeta = sin(M_PI * (2.0 * k + 1.0) / (2.0 * order));
        const double a0 = 1.0 + 2.0 * zeta * K + K * K;
        q->b0[k] = K * K / a0; q->b1[k] = 2.0 * K * K / a0; q->b2[k] = K * K / a0;
        q->a1[k] = 2.0 * (K * K - 1.0) / a0;
        q->a2[k] = (1.0 - 2.0 * zeta * K + K * K) / a0;
    }
    return FWI_OK;
}

// one thread per trace; traces are the fast axis ([nt][n_traces]) so a warp reads 128 contiguous bytes per sample.
// State in double (the poles of a 10 Hz cut-off at 1 kHz sampling sit at 1 - 0.03: fp32 recursions lose 3 digits).
// dir = +1: t = 0 .. nt-1, dir = -1: t = nt-1 .. 0; zero initial state.
__global__ void __launch_bounds__(128) sos_pass_kernel(const Biquads q, const float *__restrict__ in, float *__restrict__ out,
                                                       const int nt, const int n_traces, const int dir) {
    const int tr = blockIdx.x * blockDim.x + threadIdx.x;
    if (tr >= n_traces) return;
    double z1[kMaxSections], z2[kMaxSections];
#pragma unroll
    for (int k = 0; k < kMaxSections; ++k) z1[k] = z2[k] = 0.0;
    for (int i = 0; i < nt; ++i) {
        const int t = dir > 0 ? i : nt - 1 - i;
        double x = (double)in[(size_t)t * n_traces + tr];
#pragma unroll
        for (int k = 0; k < kMaxSections; ++k) {
            if (k < q.n) {               // transposed direct form II
                const double y = q.b0[k] * x + z1[k];
                z1[k] = q.b1[k] * x - q.a1[k] * y + z2[k];
                z2[k] = q.b2[k] * x - q.a2[k] * y;
                x = y;
            }
        }
        out[(size_t)t * n_traces + tr] = (float)x;
    }
}

// ---- gradient post-processing -----------------------------------------------------------------------------------
struct PostParams {
    int nz, ny, water_rows;
    float depth_pow;       // rho gradient *= z^depth_pow (z = row index)
    float rho_factor, rho_offset;
};

__device__ __forceinline__ void atomic_max_nonneg(float *addr, float v) {
    atomicMax(reinterpret_cast<int *>(addr), __float_as_int(v));      // order of floats >= 0 is the order of their bits
}

// pass 1: weight / mask in place; maxima of |gradient| (after weight and mask) and of |model| -> stats[0..5]
__global__ void __launch_bounds__(256) grad_post_pass1(const PostParams p, float *__restrict__ gvp, float *__restrict__ gvs,
                                                       float *__restrict__ grho, const float *__restrict__ vp,
                                                       const float *__restrict__ vs, const float *__restrict__ rho,
                                                       float *__restrict__ stats) {
    const size_t n = (size_t)p.nz * p.ny;
    float m[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int z = (int)(i / p.ny);
        float a = gvp[i], b = gvs[i], c = grho[i] * powf((float)z, p.depth_pow);
        if (z < p.water_rows) a = b = c = 0.f;
        gvp[i] = a; gvs[i] = b; grho[i] = c;
        m[0] = fmaxf(m[0], fabsf(a)); m[1] = fmaxf(m[1], fabsf(b)); m[2] = fmaxf(m[2], fabsf(c));
        m[3] = fmaxf(m[3], fabsf(vp[i])); m[4] = fmaxf(m[4], fabsf(vs[i])); m[5] = fmaxf(m[5], fabsf(rho[i]));
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m[k] = fmaxf(m[k], __shfl_xor_sync(0xffffffffu, m[k], o));
        if ((threadIdx.x & 31) == 0) atomic_max_nonneg(stats + k, m[k]);
    }
}

// pass 2: gradient *= max|model| / max|gradient|; rho additionally * rho_factor + rho_offset
__global__ void __launch_bounds__(256) grad_post_pass2(const PostParams p, float *__restrict__ gvp, float *__restrict__ gvs,
                                                       float *__restrict__ grho, const float *__restrict__ stats) {
    const size_t n = (size_t)p.nz * p.ny;
    const float r1 = stats[3] / stats[0], r2 = stats[4] / stats[1], r3 = stats[5] / stats[2];
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        gvp[i] = 1.0f * gvp[i] * r1;
        gvs[i] = 1.0f * gvs[i] * r2;
        grho[i] = p.rho_factor * grho[i] * r3 + p.rho_offset;
    }
}

}  // namespace

extern "C" int fwi_butterworth_lowpass(const fwi_butter_desc *d, const float *in, float *out, void *stream) {
    FWI_CHECK_ARG(d && in && out, "NULL argument");
    FWI_CHECK_ARG(d->nt >= 1 && d->n_traces >= 1, "empty traces");
    FWI_CHECK_ARG(d->fc_low == 0.0, "only low-pass stages are supported (fc_low must be 0, as at models/networks.py:9147)");
    Biquads q;
    int rc = design_lowpass(d->order, d->fc_high, d->dt, &q);
    if (rc != FWI_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int blocks = (d->n_traces + 127) / 128;
    sos_pass_kernel<<<blocks, 128, 0, st>>>(q, in, out, d->nt, d->n_traces, +1);
    if (d->zero_phase) sos_pass_kernel<<<blocks, 128, 0, st>>>(q, out, out, d->nt, d->n_traces, -1);
    FWI_CHECK_CUDA(cudaGetLastError());
    fwi::count_launch(d->zero_phase ? 2 : 1);
    return FWI_OK;
}

extern "C" int fwi_elastic_gradient_post(const fwi_gradpost_desc *d, float *grad_vp, float *grad_vs, float *grad_rho,
                                         const float *vp, const float *vs, const float *rho, float *stats, void *stream) {
    FWI_CHECK_ARG(d && grad_vp && grad_vs && grad_rho && vp && vs && rho && stats, "NULL argument");
    FWI_CHECK_ARG(d->nz >= 1 && d->ny >= 1 && d->water_rows >= 0, "bad extents");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    PostParams p;
    p.nz = d->nz; p.ny = d->ny; p.water_rows = d->water_rows;
    p.depth_pow = (float)d->depth_pow; p.rho_factor = (float)d->rho_factor; p.rho_offset = (float)d->rho_offset;
    FWI_CHECK_CUDA(cudaMemsetAsync(stats, 0, 6 * sizeof(float), st));
    const int blocks = (int)std::min<size_t>(((size_t)d->nz * d->ny + 255) / 256, 1184);
    grad_post_pass1<<<blocks, 256, 0, st>>>(p, grad_vp, grad_vs, grad_rho, vp, vs, rho, stats);
    grad_post_pass2<<<blocks, 256, 0, st>>>(p, grad_vp, grad_vs, grad_rho, stats);
    FWI_CHECK_CUDA(cudaGetLastError());
    fwi::count_launch(2);
    return FWI_OK;
}
