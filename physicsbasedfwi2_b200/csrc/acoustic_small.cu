// Cluster-resident acoustic time loop for small grids (the reference's own cases: 100x300 with 10
// shots, models/networks.py:2711-2745; Auto22's 151x200 with 18 shots, :5339-5358).
//
// On those grids one launch of the general step kernel (acoustic.cu) is latency-bound: the work of a
// step is a few microseconds, but every step pays a kernel hand-over and a TMA round trip.  Here the
// WHOLE time loop of a call is one kernel.  A shot is owned by one thread-block cluster; CTA r of the
// cluster keeps a band of rows of u, u-, pz, py (plus order/2 halo rows) in its shared memory for the
// entire loop, the CTAs exchange their edge rows through distributed shared memory after every step,
// and two cluster barriers per step replace the kernel boundary.  HBM sees only what the algorithm
// needs: the snapshot q (written once per step forward, read once per step in the adjoint), the
// L2-resident w = v^2 dt^2 and image partials, and the receiver traces.
//
// Same arithmetic, per cell and in the same order, as acoustic.cu and oracle/cpu_ref_impl.h (interior
// formula inside, full PML formula in the frame), so results agree with the general kernel to the bit
// on the forward pass.  Selected by the host (acoustic.cu: run_impl) when a shot fits; FWI_B200_SMALL=0
// turns it off.
#include <cooperative_groups.h>
#include <stdlib.h>

#include <algorithm>

#include "common.h"
#include "acoustic_small.h"

namespace cg = cooperative_groups;

namespace fwi {

namespace {

constexpr int kNT = 512;

template <int H>
__global__ void __launch_bounds__(kNT, 1) acoustic_small_kernel(const SmallParams p) {
    extern __shared__ __align__(16) float sm[];
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int s = blockIdx.x / p.C;
    const int tid = threadIdx.x;
    const int W = p.nyp;                                   // shared-memory row width
    const int r0 = H + rank * p.rows_per;                  // first / past-last computational row of this CTA
    const int r1 = min(p.nzp - H, r0 + p.rows_per);
    const int rows = r1 - r0;                              // may be <= 0 for a trailing CTA of a short grid
    const int LR = p.rows_per + 2 * H;                     // local rows (with halos), same for every CTA
    const int fld = LR * W;
    float *U = sm, *UM = sm + fld, *PZ = sm + 2 * fld, *PY = sm + 3 * fld;
    const int ncols = p.nyp - 2 * H;
    const int ncell = max(rows, 0) * ncols;
    const float inv_ncols = 1.0f / (float)ncols;
    // PML frame (widened by H: cells that can see a memory variable), as in cpu_ref_impl.h
    const int zlo = H + p.pw + H, zhi = p.nzp - H - p.pw - H;
    int ylo = H + p.pw + H, yhi = p.nyp - H - p.pw - H;
    const bool all_pml = p.pw > 0 && yhi <= ylo;
    const size_t gshot = (size_t)s * p.stride;
    const size_t gfld = (size_t)p.S * p.stride;

    // ---- load the state: rows [r0 - H, r0 - H + LR) of u, u-, pz, py (zero outside the grid) ----
    for (int t = tid; t < fld; t += kNT) {
        const int lz = t / W, x = t - lz * W;
        const int z = r0 - H + lz;
        const bool in = z >= 0 && z < p.nzp;
        const size_t g = gshot + (size_t)(in ? z : 0) * p.pitch + p.col0 + x;
        U[t] = in ? p.wave[g] : 0.f;
        UM[t] = in ? p.wave[gfld + g] : 0.f;
        if (p.wave_pml) {
            PZ[t] = in ? p.wave[2 * gfld + g] : 0.f;
            PY[t] = in ? p.wave[4 * gfld + g] : 0.f;
        }
    }
    cluster.sync();

    for (int it = 0; it < p.nsteps; ++it) {
        const int n = p.adjoint ? p.nsteps - 1 - it : it;
        // ---- sampling (before the step): the points of this shot that lie in my rows ----
        for (int r = tid; r < p.n_smp; r += kNT) {
            const int cell = __ldg(p.smp_cell + s * p.n_smp + r);
            const int z = cell / p.pitch, y = cell - z * p.pitch;
            if (z >= r0 && z < r1) {
                float val = U[(z - r0 + H) * W + (y - p.col0)];
                if (p.smp_div_a) val = val * (1.0f + p.hdt * (__ldg(p.sigz + z) + __ldg(p.sigy + y)));
                p.smp_out[((size_t)n * p.S + s) * p.n_smp + r] = val;
            }
        }
        // ---- phase A: q and the new field (into UM, cell by cell in place) ----
        const float *qr = p.mode == 2 ? p.q_read + (size_t)n * gfld + gshot : nullptr;
        float *qw = p.mode == 1 ? p.q_write + (size_t)n * gfld + gshot : nullptr;
        float *G = p.mode == 2 ? p.image + gshot : nullptr;
        for (int c = tid; c < ncell; c += kNT) {
            const int lr = (int)(((float)c + 0.5f) * inv_ncols);
            const int x = c - lr * ncols + H;
            const int z = r0 + lr;
            const int l = (lr + H) * W + x;
            const size_t g = (size_t)z * p.pitch + p.col0 + x;
            const float u = U[l];
            float lap = p.c0 * u;
#pragma unroll
            for (int k = 1; k <= H; ++k) lap = fmaf(p.cy2[k], U[l - k] + U[l + k], lap);
#pragma unroll
            for (int k = 1; k <= H; ++k) lap = fmaf(p.cz2[k], U[l - k * W] + U[l + k * W], lap);
            const bool pml = p.pw > 0 && (all_pml || z < zlo || z >= zhi || x < ylo || x >= yhi);
            const float wv = __ldg(p.w + g);
            if (!pml) {
                if (qr) G[g] = fmaf(u, __ldcs(qr + g), G[g]);
                if (qw) __stcs(qw + g, lap);
                UM[l] = fmaf(wv, lap, fmaf(2.0f, u, -UM[l]));
            } else {
                const float sz = __ldg(p.sigz + z), sy = __ldg(p.sigy + p.col0 + x);
                float dpz = 0.f, dpy = 0.f;
#pragma unroll
                for (int k = 1; k <= H; ++k) dpz = fmaf(p.cz1[k], PZ[l + k * W] - PZ[l - k * W], dpz);
#pragma unroll
                for (int k = 1; k <= H; ++k) dpy = fmaf(p.cy1[k], PY[l + k] - PY[l - k], dpy);
                const float q = (lap + dpz) + dpy;
                const float hs = p.hdt * (sz + sy);
                const float a = 1.0f / (1.0f + hs);
                const float cu = a * fmaf(-p.dt2, sz * sy, 2.0f);
                const float cum = a * (hs - 1.0f);
                const float aw = a * wv;
                if (qr) G[g] = fmaf(u, __ldcs(qr + g), G[g]);
                if (qw) __stcs(qw + g, q);
                UM[l] = fmaf(aw, q, fmaf(cu, u, cum * UM[l]));
            }
        }
        cluster.sync();      // everyone is done reading u / pz halos of this step
        // ---- phase B: memory variables of the frame cells, in place (u is still the old field) ----
        if (p.wave_pml) {
            for (int c = tid; c < ncell; c += kNT) {
                const int lr = (int)(((float)c + 0.5f) * inv_ncols);
                const int x = c - lr * ncols + H;
                const int z = r0 + lr;
                if (!(all_pml || z < zlo || z >= zhi || x < ylo || x >= yhi)) continue;
                const int l = (lr + H) * W + x;
                const float sz = __ldg(p.sigz + z), sy = __ldg(p.sigy + p.col0 + x);
                float duz = 0.f, duy = 0.f;
#pragma unroll
                for (int k = 1; k <= H; ++k) duz = fmaf(p.cz1[k], U[l + k * W] - U[l - k * W], duz);
#pragma unroll
                for (int k = 1; k <= H; ++k) duy = fmaf(p.cy1[k], U[l + k] - U[l - k], duy);
                PZ[l] = fmaf(p.dt * (sy - sz), duz, fmaf(-p.dt, sz, 1.0f) * PZ[l]);
                PY[l] = fmaf(p.dt * (sz - sy), duy, fmaf(-p.dt, sy, 1.0f) * PY[l]);
            }
        }
        // ---- injection into the new field: list entries of my rows, runs of equal cells in list order ----
        if (p.n_inj > 0 && rows > 0) {
            const int *rs = p.inj_row_start + (size_t)s * (p.nzp + 1);
            const int e0 = __ldg(rs + r0), e1 = __ldg(rs + r1);
            const float *amp = p.inj_amp + ((size_t)n * p.S + s) * p.n_inj;
            for (int e = e0 + tid; e < e1; e += kNT) {
                const int cell = __ldg(p.inj_cell + e);
                if (e > e0 && __ldg(p.inj_cell + e - 1) == cell) continue;      // not the first of its run
                const int z = cell / p.pitch, y = cell - z * p.pitch;
                float coef = __ldg(p.w + cell);
                if (p.inj_scale_a) coef = (1.0f / (1.0f + p.hdt * (__ldg(p.sigz + z) + __ldg(p.sigy + y)))) * coef;
                const int l = (z - r0 + H) * W + (y - p.col0);
                float v = UM[l];
                for (int k = e; k < e1 && __ldg(p.inj_cell + k) == cell; ++k)
                    v = fmaf(coef, __ldg(amp + __ldg(p.inj_col + k)), v);
                UM[l] = v;
            }
        }
        __syncthreads();
        // ---- push my edge rows of the new field and of pz into the neighbours' halo rows (DSMEM) ----
        if (rows > 0) {
            const int nedge = H * W;
            if (rank > 0) {                      // my first H rows -> the lower halo of the CTA above
                float *nUM = cluster.map_shared_rank(UM, rank - 1), *nPZ = cluster.map_shared_rank(PZ, rank - 1);
                for (int t = tid; t < nedge; t += kNT) {
                    const int src = H * W + t, dst = (H + p.rows_per) * W + t;
                    nUM[dst] = UM[src];
                    if (p.wave_pml) nPZ[dst] = PZ[src];
                }
            }
            if (rank < p.C - 1 && r1 < p.nzp - H) {     // my last H rows -> the upper halo of the CTA below
                float *nUM = cluster.map_shared_rank(UM, rank + 1), *nPZ = cluster.map_shared_rank(PZ, rank + 1);
                for (int t = tid; t < nedge; t += kNT) {
                    const int src = rows * W + t, dst = t;
                    nUM[dst] = UM[src];
                    if (p.wave_pml) nPZ[dst] = PZ[src];
                }
            }
        }
        cluster.sync();      // halos of the new field are in place
        float *tmp = U; U = UM; UM = tmp;
    }

    // ---- store the state back: own rows; fields in the slots the general kernel would leave them in ----
    const int odd = p.nsteps & 1;
    float *gu = p.wave + (odd ? gfld : 0) + gshot, *gum = p.wave + (odd ? 0 : gfld) + gshot;
    float *gpz = p.wave_pml ? p.wave + (odd ? 3 : 2) * gfld + gshot : nullptr;
    float *gpy = p.wave_pml ? p.wave + (odd ? 5 : 4) * gfld + gshot : nullptr;
    for (int c = tid; c < ncell; c += kNT) {
        const int lr = (int)(((float)c + 0.5f) * inv_ncols);
        const int x = c - lr * ncols + H;
        const int l = (lr + H) * W + x;
        const size_t g = (size_t)(r0 + lr) * p.pitch + p.col0 + x;
        gu[g] = U[l];
        gum[g] = UM[l];
        if (p.wave_pml) { gpz[g] = PZ[l]; gpy[g] = PY[l]; }
    }
}

template <int H>
int try_launch(const SmallParams &p0, cudaStream_t st, bool *launched) {
    *launched = false;
    auto kern = acoustic_small_kernel<H>;
    static int max_smem = -1;
    if (max_smem < 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 0);
    }
    const int nrows = p0.nzp - 2 * H, nf = p0.wave_pml ? 4 : 2;
    static const int cmax = getenv("FWI_B200_SMALL_CLUSTER") ? atoi(getenv("FWI_B200_SMALL_CLUSTER")) : 8;
    for (int C = std::min(cmax, 8); C >= 1; C >>= 1) {
        SmallParams p = p0;
        p.C = C;
        p.rows_per = (nrows + C - 1) / C;
        if (p.rows_per < H) continue;
        if ((C - 1) * p.rows_per >= nrows && C > 1) continue;          // a CTA would own no row
        const size_t smem = (size_t)nf * (p.rows_per + 2 * H) * p.nyp * sizeof(float);
        if (smem > (size_t)max_smem) continue;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(p.S * C);
        cfg.blockDim = dim3(kNT);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int nclusters = 0;
        if (cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg) != cudaSuccess) { cudaGetLastError(); continue; }
        if (nclusters < p.S) continue;                 // every shot's cluster must be resident at once
        cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
        if (e != cudaSuccess) {
            set_error("acoustic_small launch failed: %s", cudaGetErrorString(e));
            return FWI_ERR_CUDA;
        }
        count_launch();
        *launched = true;
        return FWI_OK;
    }
    return FWI_OK;       // does not fit: the caller uses the general kernel
}

}  // namespace

// Runs the whole time loop in one cluster-resident kernel when the problem fits; *launched says so.
int small_run(const SmallParams &p, void *stream, bool *launched) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (p.h) {
    case 1: return try_launch<1>(p, st, launched);
    case 2: return try_launch<2>(p, st, launched);
    case 3: return try_launch<3>(p, st, launched);
    default: return try_launch<4>(p, st, launched);
    }
}

}  // namespace fwi
