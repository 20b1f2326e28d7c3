// Device-side preparation and post-processing of one propagator call: what deepwave's Python layer does around its
// native time loop for  Propagator({'vp': v}, dx)(src, x_s, x_r, dt)  (reference call sites models/networks.py:5449,
// 5464, backward :5491) - pad the model, scale it to w = v^2 dt^2, build the PML profiles, turn metre coordinates into
// cells and sorted injection lists, and fold the image of the adjoint pass back to a model gradient.  Round 1 did this
// with ~125 eager torch micro-kernels per gradient (1.3 ms of a 9.9 ms C1 gradient); here it is three launches.
// All three are deterministic (fixed summation orders, no atomics on data).
#include <math.h>

#include <algorithm>

#include "common.h"

namespace {

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

struct ModelGeom {
    int nz, ny, h, pw, P, nzp, nyp_global, c0, c1, nyp, pitch, col0;
};

int make_model_geom(const fwi_model_desc *m, ModelGeom *g) {
    FWI_CHECK_ARG(m != nullptr, "model descriptor is NULL");
    FWI_CHECK_ARG(m->order == 2 || m->order == 4 || m->order == 6 || m->order == 8, "order must be 2, 4, 6 or 8 (got %d)", m->order);
    FWI_CHECK_ARG(m->nz >= 1 && m->ny >= 1 && m->pml_width >= 0, "bad model extents / pml_width");
    FWI_CHECK_ARG(m->dz > 0 && m->dy > 0, "dz, dy must be > 0");
    g->nz = m->nz; g->ny = m->ny; g->h = m->order / 2; g->pw = m->pml_width; g->P = g->pw + g->h;
    g->nzp = m->nz + 2 * g->P; g->nyp_global = m->ny + 2 * g->P;
    g->c0 = m->c0; g->c1 = m->c1 > 0 ? m->c1 : g->nyp_global;
    FWI_CHECK_ARG(0 <= g->c0 && g->c0 < g->c1 && g->c1 <= g->nyp_global, "bad column window [%d, %d)", g->c0, g->c1);
    g->nyp = g->c1 - g->c0;
    fwi_acoustic_desc d = {};
    d.nzp = g->nzp; d.nyp = g->nyp; d.order = m->order; d.pml_width = m->pml_width; d.n_shots = 1;
    d.dz = m->dz; d.dy = m->dy; d.dt = 1.0;
    fwi_acoustic_layout lay;
    int rc = fwi_acoustic_get_layout(&d, &lay);
    if (rc != FWI_OK) return rc;
    g->pitch = lay.pitch; g->col0 = lay.col0;
    return FWI_OK;
}

// quadratic PML ramp of oracle/torch_ref.py pml_profile / _host.sigma_profile, evaluated in double like they do
__device__ __forceinline__ float sigma_at(int i, int n, int h, int pw, double top) {
    if (pw <= 0) return 0.f;
    const double below = (double)(h + pw) - (double)i;
    const double above = (double)i - (double)(n - h - pw - 1);
    double depth = fmax(below, above);
    depth = fmin(fmax(depth, 0.0), (double)pw);
    const double r = depth / (double)pw;
    return (float)(top * (r * r));
}

// w[z][col0 + y] = (v[clamp(z - P)][clamp(c0 + y - P)])^2 * dt2 on the window; everything else of the pitched row 0;
// block (0, 0) also writes the two profiles
__global__ void __launch_bounds__(256) prepare_model_kernel(const ModelGeom g, const float *__restrict__ v, float *__restrict__ w,
                                                            float *__restrict__ sigz, float *__restrict__ sigy, const float dt2,
                                                            const double top_z, const double top_y) {
    const int z = blockIdx.y;
    const int zc = min(max(z - g.P, 0), g.nz - 1);
    for (int col = blockIdx.x * blockDim.x + threadIdx.x; col < g.pitch; col += gridDim.x * blockDim.x) {
        const int y = col - g.col0;
        float out = 0.f;
        if (y >= 0 && y < g.nyp) {
            const int yc = min(max(g.c0 + y - g.P, 0), g.ny - 1);
            const float val = __ldg(v + (size_t)zc * g.ny + yc);
            out = __fmul_rn(__fmul_rn(val, val), dt2);
        }
        w[(size_t)z * g.pitch + col] = out;
        if (z == 0) {
            float s = 0.f;
            if (y >= 0 && y < g.nyp) s = sigma_at(g.c0 + y, g.nyp_global, g.h, g.pw, top_y);
            sigy[col] = s;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) sigz[z] = sigma_at(z, g.nzp, g.h, g.pw, top_z);
}

// One CTA per shot.  Phase 1: metres -> clamped device cell + out-of-model flag.  Phase 2: stable rank of every point by
// cell (count of smaller keys + equal keys that come earlier) -> sorted lists.  Phase 3: row offsets by binary search.
__global__ void __launch_bounds__(256) plan_points_kernel(const ModelGeom g, const float *__restrict__ x, const int n,
                                                          const int loc_round, const double inv_dz, const double dz,
                                                          const double dy, int *__restrict__ cell, int *__restrict__ row_start,
                                                          int *__restrict__ sorted_cell, int *__restrict__ sorted_col,
                                                          int *__restrict__ bad) {
    (void)inv_dz;
    const int s = blockIdx.x;
    int *key = cell + (size_t)s * n;
    int any_bad = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const double mz = (double)x[((size_t)s * n + i) * 2 + 0] / dz;
        const double my = (double)x[((size_t)s * n + i) * 2 + 1] / dy;
        const long long cz = (long long)(loc_round ? rint(mz) : floor(mz + 1e-6));
        const long long cy = (long long)(loc_round ? rint(my) : floor(my + 1e-6));
        if (cz < 0 || cz >= g.nz || cy < 0 || cy >= g.ny) any_bad = 1;
        const int zc = (int)min(max(cz, 0LL), (long long)g.nz - 1) + g.P;
        int yc = (int)min(max(cy, 0LL), (long long)g.ny - 1) + g.P - g.c0;
        yc = min(max(yc, g.h), g.nyp - g.h - 1) + g.col0;
        key[i] = zc * g.pitch + yc;
    }
    if (any_bad) atomicOr(bad, 1);
    __syncthreads();
    int *sc = sorted_cell + (size_t)s * n;
    int *so = sorted_col + (size_t)s * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int k = key[i];
        int rank = 0;
        for (int j = 0; j < n; ++j) {
            const int kj = key[j];
            rank += (kj < k) || (kj == k && j < i);
        }
        sc[rank] = k;
        so[rank] = i;
    }
    __syncthreads();
    int *rs = row_start + (size_t)s * (g.nzp + 1);
    for (int z = threadIdx.x; z <= g.nzp; z += blockDim.x) {
        const int bound = z * g.pitch;
        int lo = 0, hi = n;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (sc[mid] < bound) lo = mid + 1; else hi = mid;
        }
        rs[z] = s * n + lo;
    }
}

// gsum[cell] = sum over shots (shot order) of image[s][cell]
__global__ void __launch_bounds__(256) image_sum_kernel(const float *__restrict__ image, const int n_shots, const size_t stride,
                                                        float *__restrict__ gsum) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < stride; i += (size_t)gridDim.x * blockDim.x) {
        float acc = 0.f;
        for (int s = 0; s < n_shots; ++s) acc += image[(size_t)s * stride + i];
        gsum[i] = acc;
    }
}

// term[p] = sum_n gs[n][p] * amp[n][p], one warp per point p of the flat [S * n_src] list (lane-strided, fixed tree)
__global__ void __launch_bounds__(256) source_term_kernel(const float *__restrict__ gs, const float *__restrict__ amp,
                                                          const int nsteps, const int n_pts, float *__restrict__ term) {
    const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= n_pts) return;
    float acc = 0.f;
    for (int n = lane; n < nsteps; n += 32) acc = fmaf(gs[(size_t)n * n_pts + p], amp[(size_t)n * n_pts + p], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) term[p] = acc;
}

// gsum[cell(p)] += term[p]: points sharing a cell (also across shots) are added by the first of them, in list order
__global__ void __launch_bounds__(256) source_scatter_kernel(const int *__restrict__ src_cell, const float *__restrict__ term,
                                                             const int n_pts, float *__restrict__ gsum) {
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n_pts; p += gridDim.x * blockDim.x) {
        const int c = src_cell[p];
        bool head = true;
        for (int k = 0; k < p; ++k)
            if (src_cell[k] == c) { head = false; break; }
        if (!head) continue;
        float acc = gsum[c];
        for (int k = p; k < n_pts; ++k)
            if (src_cell[k] == c) acc += term[k];
        gsum[c] = acc;
    }
}

// grad_v[z][y] = sum over the padded cells that replicate (z, y) of (2 / v[z][y]) * gsum[cell]: rows of the pad first,
// then columns (the order of _host.replicate_pad_adjoint)
__global__ void __launch_bounds__(256) fold_kernel(const ModelGeom g, const float *__restrict__ gsum, const float *__restrict__ v,
                                                   float *__restrict__ grad_v) {
    const int y = blockIdx.x * blockDim.x + threadIdx.x, z = blockIdx.y;
    if (y >= g.ny) return;
    const float f = 2.0f / v[(size_t)z * g.ny + y];
    const int z0 = z == 0 ? 0 : z + g.P, z1 = z == g.nz - 1 ? g.nzp : z + g.P + 1;          // padded rows mapped to z
    const int y0 = y == 0 ? 0 : y + g.P, y1 = y == g.ny - 1 ? g.nyp : y + g.P + 1;          // padded columns mapped to y
    const int zc = z + g.P, yc = y + g.P;
    auto colsum = [&](const int c) {
        const float *col = gsum + g.col0 + c;
        float pad = 0.f;
        bool any = false;
        for (int r = z0; r < z1; ++r)
            if (r != zc) { pad += f * col[(size_t)r * g.pitch]; any = true; }
        const float centre = f * col[(size_t)zc * g.pitch];
        return any ? centre + pad : centre;
    };
    float pad = 0.f;
    bool any = false;
    for (int c = y0; c < y1; ++c)
        if (c != yc) { pad += colsum(c); any = true; }
    const float centre = colsum(yc);
    grad_v[(size_t)z * g.ny + y] = any ? centre + pad : centre;
}

}  // namespace

extern "C" int fwi_acoustic_prepare_model(const fwi_model_desc *m, const float *v, float *w, float *sigz, float *sigy,
                                          void *stream) {
    ModelGeom g;
    int rc = make_model_geom(m, &g);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(v && w && sigz && sigy, "NULL array");
    FWI_CHECK_ARG(m->vmax > 0 && isfinite(m->vmax), "vmax must be positive and finite");
    const float dt32 = (float)m->dt;
    const float dt2 = dt32 * dt32;
    const double top_z = g.pw > 0 ? 3.0 * m->vmax * log(1000.0) / (2.0 * (g.pw * m->dz)) : 0.0;
    const double top_y = g.pw > 0 ? 3.0 * m->vmax * log(1000.0) / (2.0 * (g.pw * m->dy)) : 0.0;
    dim3 grid(std::min(ceil_div(g.pitch, 256), 64), g.nzp);
    prepare_model_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(g, v, w, sigz, sigy, dt2, top_z, top_y);
    FWI_CHECK_CUDA(cudaGetLastError());
    fwi::count_launch();
    return FWI_OK;
}

extern "C" int fwi_acoustic_plan_points(const fwi_model_desc *m, const float *x, int32_t n_shots, int32_t n, int32_t loc_round,
                                        int32_t *cell, int32_t *row_start, int32_t *sorted_cell, int32_t *sorted_col,
                                        int32_t *bad, void *stream) {
    ModelGeom g;
    int rc = make_model_geom(m, &g);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(n_shots >= 1 && n >= 0, "bad point counts");
    FWI_CHECK_ARG(row_start && bad, "NULL array");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n == 0) {
        FWI_CHECK_CUDA(cudaMemsetAsync(row_start, 0, sizeof(int32_t) * (size_t)n_shots * (g.nzp + 1), st));
        return FWI_OK;
    }
    FWI_CHECK_ARG(x && cell && sorted_cell && sorted_col, "NULL array");
    FWI_CHECK_ARG((int64_t)g.nzp * g.pitch < ((int64_t)1 << 31), "grid too large for 32-bit cells");
    plan_points_kernel<<<n_shots, 256, 0, st>>>(g, x, n, loc_round, 1.0 / m->dz, m->dz, m->dy, cell, row_start, sorted_cell,
                                                sorted_col, bad);
    FWI_CHECK_CUDA(cudaGetLastError());
    fwi::count_launch();
    return FWI_OK;
}

extern "C" int64_t fwi_acoustic_fold_workspace_floats(const fwi_model_desc *m, int32_t n_shots, int32_t n_src) {
    ModelGeom g;
    if (make_model_geom(m, &g) != FWI_OK) return -1;
    return (int64_t)g.nzp * g.pitch + (int64_t)std::max(n_shots, 1) * std::max(n_src, 1);
}

extern "C" int fwi_acoustic_fold_gradient(const fwi_model_desc *m, int32_t n_shots, const float *image, const float *gs,
                                          const float *amp, const int32_t *src_cell, int32_t n_src, int32_t nsteps,
                                          const float *v, float *workspace, float *grad_v, void *stream) {
    ModelGeom g;
    int rc = make_model_geom(m, &g);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(g.c0 == 0 && g.c1 == g.nyp_global, "the gradient fold works on the whole grid, not on a column window");
    FWI_CHECK_ARG(n_shots >= 1 && image && v && workspace && grad_v, "NULL array / bad shot count");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t stride = (size_t)g.nzp * g.pitch;
    float *gsum = workspace, *term = workspace + stride;
    image_sum_kernel<<<std::min<int>(ceil_div((int)std::min<size_t>(stride, 1u << 30), 256), 1184), 256, 0, st>>>(image, n_shots, stride, gsum);
    int launches = 1;
    if (gs && amp && src_cell && n_src > 0 && nsteps > 0) {
        const int n_pts = n_shots * n_src;
        source_term_kernel<<<ceil_div(n_pts * 32, 256), 256, 0, st>>>(gs, amp, nsteps, n_pts, term);
        source_scatter_kernel<<<std::min(ceil_div(n_pts, 256), 64), 256, 0, st>>>(src_cell, term, n_pts, gsum);
        launches += 2;
    }
    dim3 grid(ceil_div(g.ny, 256), g.nz);
    fold_kernel<<<grid, 256, 0, st>>>(g, gsum, v, grad_v);
    FWI_CHECK_CUDA(cudaGetLastError());
    fwi::count_launch(launches + 1);
    return FWI_OK;
}
