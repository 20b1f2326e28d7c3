// Cluster-resident acoustic time loop for small grids (the reference's own cases: 100x300 with 10
// shots, models/networks.py:2711-2745; Auto22's 151x200 with 18 shots, :5339-5358).
//
// On those grids one launch of the general step kernel (acoustic.cu) is latency-bound: the work of a
// step is a few microseconds, but every step pays a kernel hand-over and a TMA round trip.  Here the
// WHOLE time loop of a call is one kernel.  A shot is owned by one thread-block cluster; CTA r of the
// cluster keeps a band of rows of u, u-, pz, py (plus order/2 halo rows) in its shared memory for the
// entire loop, the CTAs exchange their edge rows through distributed shared memory after every step
// (asynchronous remote stores that signal an mbarrier of the receiver: a CTA waits for its two
// neighbours, not for the cluster).  HBM sees only what the algorithm
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

// Cell c of a CTA (row-major over its rows x computational columns) -> local row / padded column.
struct CellMap {
    int ncols, H, W, r0, pitch, col0;
    float inv;
    __device__ __forceinline__ void at(int c, int &lr, int &x) const {
        lr = (int)(((float)c + 0.5f) * inv);
        x = c - lr * ncols + H;
    }
    __device__ __forceinline__ int loc(int lr, int x) const { return (lr + H) * W + x; }
    __device__ __forceinline__ size_t glob(int lr, int x) const { return (size_t)(r0 + lr) * pitch + col0 + x; }
};

// ---- halo rows through asynchronous remote stores (p.async_halo) ---------------------------------------------
// The edge rows a CTA pushes into its neighbours' halo rows are written with st.async, which signals the bytes it has
// delivered on an mbarrier in the RECEIVER's shared memory; the receiver waits on that barrier for the bytes of both
// neighbours.  A step then synchronises a CTA with its two neighbours only, and nothing waits for a release fence (the
// fence of cluster.sync also waited for the step's snapshot stores to reach L2: 10 % of the step, as much again for the
// barrier itself - profiles/r02e_small_full.txt).  Two barriers alternate by step parity: a neighbour can be at most
// one step ahead (its step n + 1 needs my rows of step n), so the bytes of step n + 2 cannot arrive before the wait of
// step n has been passed.
__device__ __forceinline__ unsigned smem_addr(const void *p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ unsigned map_to_rank(unsigned addr, unsigned rank) {
    unsigned r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_async_f32(unsigned remote_addr, float v, unsigned remote_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(remote_addr),
                 "r"(__float_as_uint(v)), "r"(remote_bar)
                 : "memory");
}
__device__ __forceinline__ void halo_bar_expect(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void halo_bar_wait(unsigned bar, unsigned parity) {
    unsigned done = 0;
    for (unsigned spin = 0; !done; ++spin) {
        asm volatile(
            "{\n"
            ".reg .pred P1;\n"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%1], %2;\n"
            "selp.u32 %0, 1, 0, P1;\n"
            "}\n"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (spin > (1u << 24)) __trap();       // (seconds: a lost signal must not hang the device)
    }
}

// CPT = cells per thread (compile-time bound: per-cell constants and accumulators live in registers
// for the whole time loop); ADJ = adjoint pass with imaging (reads the snapshot, accumulates the image).
template <int H, int CPT, bool ADJ>
__global__ void __launch_bounds__(kNT, 1) acoustic_small_kernel(const SmallParams p) {
    extern __shared__ __align__(16) float sm[];
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int s = blockIdx.x / p.C;
    const int tid = threadIdx.x;
    const int W = p.nyp;                                   // shared-memory row width
    const int r0 = p.row_start[rank], r1 = p.row_start[rank + 1];   // my computational rows
    const int rows = r1 - r0;
    const int LR = p.rows_per + 2 * H;                     // local rows (with halos), same for every CTA
    const int fld = LR * W;
    // pz is double-buffered (PZ read, PZN written, swapped every step): a neighbour may then store the halo rows
    // of the new pz while this CTA still reads the old ones, and the barrier between the two phases of a step
    // can be a CTA barrier instead of a cluster barrier
    // (py likewise, so that the memory variables are advanced in the same sweep as the field: no second phase)
    float *U = sm, *UM = sm + fld, *PZ = sm + 2 * fld, *PZN = sm + 3 * fld, *PY = sm + 4 * fld, *PYN = sm + 5 * fld;
    // PML profiles of my rows / of all columns: cluster.sync invalidates L1, so per-step __ldg of them would
    // pay an L2 round trip on the critical path of every step
    float *SZ = sm + (p.wave_pml ? 6 : 2) * fld, *SY = SZ + LR;
    CellMap cm;
    cm.ncols = p.nyp - 2 * H; cm.H = H; cm.W = W; cm.r0 = r0; cm.pitch = p.pitch; cm.col0 = p.col0;
    cm.inv = 1.0f / (float)cm.ncols;
    const int ncell = max(rows, 0) * cm.ncols;
    // PML frame (widened by H: cells that can see a memory variable), as in cpu_ref_impl.h
    const int zlo = H + p.pw + H, zhi = p.nzp - H - p.pw - H;
    const int ylo = H + p.pw + H, yhi = p.nyp - H - p.pw - H;
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
            PZN[t] = 0.f;
            PY[t] = in ? p.wave[4 * gfld + g] : 0.f;
            PYN[t] = 0.f;
        }
    }
    for (int t = tid; t < LR + W; t += kNT) {
        if (t < LR) {
            const int z = r0 - H + t;
            SZ[t] = (p.wave_pml && z >= 0 && z < p.nzp) ? __ldg(p.sigz + z) : 0.f;
        } else {
            SY[t - LR] = p.wave_pml ? __ldg(p.sigy + p.col0 + (t - LR)) : 0.f;
        }
    }
    // ---- per-cell constants and accumulators in registers ----
    // cell descriptor: bits 0..11 padded column x, bits 12..23 local row lr, bit 30 = cell of the PML frame,
    // -1 = no cell (the index arithmetic and the frame test are done once, not every step)
    float wreg[CPT], Greg[ADJ ? CPT : 1], qreg[ADJ ? CPT : 1];
    int cell[CPT];
    float *G = ADJ ? p.image + gshot : nullptr;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const int c = tid + k * kNT;
        wreg[k] = 0.f;
        cell[k] = -1;
        if (ADJ) { Greg[k] = 0.f; qreg[k] = 0.f; }
        if (c < ncell) {
            int lr, x;
            cm.at(c, lr, x);
            const int z = r0 + lr;
            const bool pml = p.pw > 0 && (all_pml || z < zlo || z >= zhi || x < ylo || x >= yhi);
            cell[k] = x | (lr << 12) | (pml ? (1 << 30) : 0);
            wreg[k] = __ldg(p.w + cm.glob(lr, x));
            if (ADJ) Greg[k] = G[cm.glob(lr, x)];
        }
    }
    // ---- this thread's sampling point and injection entry (constant over the time loop) ----
    int smp_l = -1;                 // local offset of sampling point r = tid if it lies in my rows
    float smp_scale = 1.0f;
    if (tid < p.n_smp) {
        const int cell = __ldg(p.smp_cell + s * p.n_smp + tid);
        const int z = cell / p.pitch, y = cell - z * p.pitch;
        if (z >= r0 && z < r1) {
            smp_l = (z - r0 + H) * W + (y - p.col0);
            if (p.smp_div_a) smp_scale = 1.0f + p.hdt * (__ldg(p.sigz + z) + __ldg(p.sigy + y));
        }
    }
    int e0 = 0, e1 = 0;             // list entries of my rows
    if (p.n_inj > 0 && rows > 0) {
        const int *rs = p.inj_row_start + (size_t)s * (p.nzp + 1);
        e0 = __ldg(rs + r0); e1 = __ldg(rs + r1);
    }
    int inj_l = -1, inj_col = 0, inj_run = 0;      // entry e0 + tid: first of a run of equal cells?
    float inj_coef = 0.f;
    if (e0 + tid < e1) {
        const int e = e0 + tid;
        const int cell = __ldg(p.inj_cell + e);
        if (!(e > e0 && __ldg(p.inj_cell + e - 1) == cell)) {
            const int z = cell / p.pitch, y = cell - z * p.pitch;
            inj_coef = __ldg(p.w + cell);
            if (p.inj_scale_a) inj_coef = (1.0f / (1.0f + p.hdt * (__ldg(p.sigz + z) + __ldg(p.sigy + y)))) * inj_coef;
            inj_l = (z - r0 + H) * W + (y - p.col0);
            inj_col = __ldg(p.inj_col + e);
            inj_run = 1;
            while (e + inj_run < e1 && __ldg(p.inj_cell + e + inj_run) == cell) ++inj_run;
        }
    }
    const int n_first = p.adjoint ? p.nsteps - 1 : 0;
    if (ADJ) {
        const float *qr = p.q_read + (size_t)n_first * gfld + gshot;
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const int c = tid + k * kNT;
            if (cell[k] >= 0) qreg[k] = __ldcs(qr + cm.glob((cell[k] >> 12) & 0xfff, cell[k] & 0xfff));
        }
    }
    // (two mbarriers behind the profiles, in the dynamic allocation: static shared memory would count against the opt-in limit)
    unsigned long long *halo_bar = reinterpret_cast<unsigned long long *>(sm + (((p.wave_pml ? 6 : 2) * fld + LR + W + 1) & ~1));
    if (p.async_halo && tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&halo_bar[0])) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&halo_bar[1])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // bytes my neighbours deliver per step: H rows of the field (and of pz) from each of them
    const unsigned halo_bytes = (unsigned)((rank > 0 ? 1 : 0) + (rank < p.C - 1 ? 1 : 0)) * (unsigned)(H * W) * 4u * (p.wave_pml ? 2u : 1u);
    const unsigned sm_up = rank > 0 ? map_to_rank(smem_addr(sm), rank - 1) : 0u;            // the neighbours' dynamic shared memory
    const unsigned sm_dn = rank < p.C - 1 ? map_to_rank(smem_addr(sm), rank + 1) : 0u;
    cluster.sync();

    for (int it = 0; it < p.nsteps; ++it) {
        const int n = p.adjoint ? p.nsteps - 1 - it : it;
        const unsigned hb = smem_addr(&halo_bar[it & 1]);
        if (p.async_halo && tid == 0) halo_bar_expect(hb, halo_bytes);
        // the amplitude this thread injects at the end of the step: in flight during the stencil work
        float amp0 = 0.f;
        const float *amp = p.n_inj > 0 ? p.inj_amp + ((size_t)n * p.S + s) * p.n_inj : nullptr;
        if (inj_run > 0) amp0 = __ldg(amp + inj_col);
        // ---- sampling (before the step) ----
        if (smp_l >= 0) p.smp_out[((size_t)n * p.S + s) * p.n_smp + tid] = U[smp_l] * smp_scale;
        for (int r = tid + kNT; r < p.n_smp; r += kNT) {           // more points than threads: the slow way
            const int cell = __ldg(p.smp_cell + s * p.n_smp + r);
            const int z = cell / p.pitch, y = cell - z * p.pitch;
            if (z >= r0 && z < r1) {
                float val = U[(z - r0 + H) * W + (y - p.col0)];
                if (p.smp_div_a) val = val * (1.0f + p.hdt * (__ldg(p.sigz + z) + __ldg(p.sigy + y)));
                p.smp_out[((size_t)n * p.S + s) * p.n_smp + r] = val;
            }
        }
        // ---- phase A: q and the new field (into UM, cell by cell in place) ----
        float *qw = p.mode == 1 ? p.q_write + (size_t)n * gfld + gshot : nullptr;
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            if (cell[k] < 0) break;
            const int x = cell[k] & 0xfff, lr = (cell[k] >> 12) & 0xfff;
            const int l = cm.loc(lr, x);
            const float u = U[l];
            float lap = p.c0 * u;
#pragma unroll
            for (int j = 1; j <= H; ++j) lap = fmaf(p.cy2[j], U[l - j] + U[l + j], lap);
#pragma unroll
            for (int j = 1; j <= H; ++j) lap = fmaf(p.cz2[j], U[l - j * W] + U[l + j * W], lap);
            const bool pml = (cell[k] >> 30) & 1;
            if (ADJ) Greg[k] = fmaf(u, qreg[k], Greg[k]);
            if (!pml) {
                if (qw) __stcs(qw + cm.glob(lr, x), lap);
                UM[l] = fmaf(wreg[k], lap, fmaf(2.0f, u, -UM[l]));
            } else {
                const float sz = SZ[lr + H], sy = SY[x];
                float dpz = 0.f, dpy = 0.f;
#pragma unroll
                for (int j = 1; j <= H; ++j) dpz = fmaf(p.cz1[j], PZ[l + j * W] - PZ[l - j * W], dpz);
#pragma unroll
                for (int j = 1; j <= H; ++j) dpy = fmaf(p.cy1[j], PY[l + j] - PY[l - j], dpy);
                const float q = (lap + dpz) + dpy;
                const float hs = p.hdt * (sz + sy);
                const float a = __frcp_rn(1.0f + hs);          // correctly rounded: the same value as 1.0f / (1.0f + hs)
                const float cu = a * fmaf(-p.dt2, sz * sy, 2.0f);
                const float cum = a * (hs - 1.0f);
                const float aw = a * wreg[k];
                float duz = 0.f, duy = 0.f;
#pragma unroll
                for (int j = 1; j <= H; ++j) duz = fmaf(p.cz1[j], U[l + j * W] - U[l - j * W], duz);
#pragma unroll
                for (int j = 1; j <= H; ++j) duy = fmaf(p.cy1[j], U[l + j] - U[l - j], duy);
                UM[l] = fmaf(aw, q, fmaf(cu, u, cum * UM[l]));
                PZN[l] = fmaf(p.dt * (sy - sz), duz, fmaf(-p.dt, sz, 1.0f) * PZ[l]);
                PYN[l] = fmaf(p.dt * (sz - sy), duy, fmaf(-p.dt, sy, 1.0f) * PY[l]);
                if (qw) __stcs(qw + cm.glob(lr, x), q);
            }
        }
        if (ADJ && it + 1 < p.nsteps) {      // next step's snapshot: in flight across the barriers
            const float *qr = p.q_read + (size_t)(n - 1) * gfld + gshot;
#pragma unroll
            for (int k = 0; k < CPT; ++k) {
                const int c = tid + k * kNT;
                if (cell[k] >= 0) qreg[k] = __ldcs(qr + cm.glob((cell[k] >> 12) & 0xfff, cell[k] & 0xfff));
            }
        }
        __syncthreads();     // the new field is complete: the injection below adds to cells other threads wrote
        // ---- injection into the new field: runs of equal cells are applied in list order ----
        if (inj_run > 0) {
            float v = fmaf(inj_coef, amp0, UM[inj_l]);
            for (int k = 1; k < inj_run; ++k) v = fmaf(inj_coef, __ldg(amp + __ldg(p.inj_col + e0 + tid + k)), v);
            UM[inj_l] = v;
        }
        for (int e = e0 + tid + kNT; e < e1; e += kNT) {           // more entries than threads: the slow way
            const int cell = __ldg(p.inj_cell + e);
            if (__ldg(p.inj_cell + e - 1) == cell) continue;
            const int z = cell / p.pitch, y = cell - z * p.pitch;
            float coef = __ldg(p.w + cell);
            if (p.inj_scale_a) coef = (1.0f / (1.0f + p.hdt * (__ldg(p.sigz + z) + __ldg(p.sigy + y)))) * coef;
            const int l = (z - r0 + H) * W + (y - p.col0);
            float v = UM[l];
            for (int k = e; k < e1 && __ldg(p.inj_cell + k) == cell; ++k) v = fmaf(coef, __ldg(amp + __ldg(p.inj_col + k)), v);
            UM[l] = v;
        }
        __syncthreads();
        // ---- push my edge rows of the new field and of pz into the neighbours' halo rows (DSMEM) ----
        if (p.async_halo) {
            const int nedge = H * W;
            const unsigned o_um = (unsigned)(UM - sm) * 4u, o_pz = (unsigned)(PZN - sm) * 4u;
            if (rank > 0) {                      // my first H rows -> the lower halo of the CTA above
                const unsigned rbar = map_to_rank(hb, rank - 1);
                const int dst0 = (H + (r0 - p.row_start[rank - 1])) * W;
                for (int t = tid; t < nedge; t += kNT) {
                    st_async_f32(sm_up + o_um + (unsigned)(dst0 + t) * 4u, UM[H * W + t], rbar);
                    if (p.wave_pml) st_async_f32(sm_up + o_pz + (unsigned)(dst0 + t) * 4u, PZN[H * W + t], rbar);
                }
            }
            if (rank < p.C - 1) {                // my last H rows -> the upper halo of the CTA below
                const unsigned rbar = map_to_rank(hb, rank + 1);
                for (int t = tid; t < nedge; t += kNT) {
                    st_async_f32(sm_dn + o_um + (unsigned)t * 4u, UM[rows * W + t], rbar);
                    if (p.wave_pml) st_async_f32(sm_dn + o_pz + (unsigned)t * 4u, PZN[rows * W + t], rbar);
                }
            }
            halo_bar_wait(hb, (unsigned)(it >> 1) & 1u);     // my halo rows of the new field / pz have arrived
        } else {
            if (rows > 0) {
                const int nedge = H * W;
                if (rank > 0) {                      // my first H rows -> the lower halo of the CTA above
                    float *nUM = cluster.map_shared_rank(UM, rank - 1), *nPZ = cluster.map_shared_rank(PZN, rank - 1);
                    for (int t = tid; t < nedge; t += kNT) {
                        const int src = H * W + t, dst = (H + (r0 - p.row_start[rank - 1])) * W + t;
                        nUM[dst] = UM[src];
                        if (p.wave_pml) nPZ[dst] = PZN[src];
                    }
                }
                if (rank < p.C - 1) {                // my last H rows -> the upper halo of the CTA below
                    float *nUM = cluster.map_shared_rank(UM, rank + 1), *nPZ = cluster.map_shared_rank(PZN, rank + 1);
                    for (int t = tid; t < nedge; t += kNT) {
                        const int src = rows * W + t, dst = t;
                        nUM[dst] = UM[src];
                        if (p.wave_pml) nPZ[dst] = PZN[src];
                    }
                }
            }
            cluster.sync();      // halos of the new field and of the new pz are in place
        }
        float *tmp = U; U = UM; UM = tmp;
        tmp = PZ; PZ = PZN; PZN = tmp;
        tmp = PY; PY = PYN; PYN = tmp;
    }

    if (p.async_halo) cluster.sync();      // no CTA exits while remote stores into it may be in flight
    // ---- store the state back: own rows; fields in the slots the general kernel would leave them in ----
    const int odd = p.nsteps & 1;
    float *gu = p.wave + (odd ? gfld : 0) + gshot, *gum = p.wave + (odd ? 0 : gfld) + gshot;
    float *gpz = p.wave_pml ? p.wave + (odd ? 3 : 2) * gfld + gshot : nullptr;
    float *gpy = p.wave_pml ? p.wave + (odd ? 5 : 4) * gfld + gshot : nullptr;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        if (cell[k] < 0) break;
        const int x = cell[k] & 0xfff, lr = (cell[k] >> 12) & 0xfff;
        const int l = cm.loc(lr, x);
        const size_t g = cm.glob(lr, x);
        gu[g] = U[l];
        gum[g] = UM[l];
        if (p.wave_pml) { gpz[g] = PZ[l]; gpy[g] = PY[l]; }
        if (ADJ) G[g] = Greg[k];
    }
}

template <int H, int CPT, bool ADJ>
int try_launch(const SmallParams &p0, cudaStream_t st, bool *launched) {
    auto kern = acoustic_small_kernel<H, CPT, ADJ>;
    static int smem_of[64];               // per device: function attributes are per device (0 = not set yet)
    int dev = 0;
    cudaGetDevice(&dev);
    int max_smem = (dev >= 0 && dev < 64) ? smem_of[dev] : 0;
    if (max_smem == 0) {
        cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (dev >= 0 && dev < 64) smem_of[dev] = max_smem;
    }
    const int nf = p0.wave_pml ? 6 : 2;
    if (p0.nyp >= 4096 || p0.nzp >= 4096) return FWI_OK;      // cell descriptors pack row and column in 12 bits each
    const int env_async = getenv("FWI_B200_SMALL_ASYNC") ? atoi(getenv("FWI_B200_SMALL_ASYNC")) : 1;
    const int cmax = getenv("FWI_B200_SMALL_CLUSTER") ? atoi(getenv("FWI_B200_SMALL_CLUSTER")) : 8;
    for (int C = std::min(cmax, 8); C >= 1; --C) {        // the largest cluster whose shots are all co-resident
        SmallParams p = p0;
        p.C = C;
        p.async_halo = env_async;
        // cost-balanced split of the computational rows: a row of the PML frame costs about 2.6 interior rows,
        // an interior row 1 + 1.6 x (share of its columns in the frame)
        const int zlo = 2 * H + p.pw, zhi = p.nzp - 2 * H - p.pw, ncols = p.nyp - 2 * H;
        const double mid = 1.0 + (p.pw > 0 ? 1.6 * std::min(1.0, 2.0 * (p.pw + 2 * H) / (double)ncols) : 0.0);
        auto cost = [&](int z) { return (p.pw > 0 && (z < zlo || z >= zhi)) ? 2.6 : mid; };
        double total = 0;
        for (int z = H; z < p.nzp - H; ++z) total += cost(z);
        p.row_start[0] = H;
        {
            double acc = 0;
            int r = 1;
            for (int z = H; z < p.nzp - H && r < C; ++z) {
                acc += cost(z);
                if (acc >= total * r / C) p.row_start[r++] = z + 1;
            }
            for (; r <= C; ++r) p.row_start[r] = p.nzp - H;
        }
        p.row_start[C] = p.nzp - H;
        bool ok = true;
        p.rows_per = 0;
        for (int r = 0; r < C; ++r) {
            const int n = p.row_start[r + 1] - p.row_start[r];
            if (n < H) ok = false;                                       // pushes come from own rows only
            p.rows_per = std::max(p.rows_per, n);
        }
        if (!ok) continue;
        if ((long long)p.rows_per * ncols > (long long)CPT * kNT) continue;   // too many cells per thread
        // fields, the two profiles, and two 8-byte mbarriers on an 8-byte boundary
        const size_t smem = ((((size_t)nf * (p.rows_per + 2 * H) * p.nyp + (p.rows_per + 2 * H) + p.nyp + 1) & ~(size_t)1) + 4) * sizeof(float);
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

// smallest register footprint first: 12, 16, then 24 cells per thread
template <int H, bool ADJ>
int try_cpt(const SmallParams &p, cudaStream_t st, bool *launched) {
    int rc = try_launch<H, 12, ADJ>(p, st, launched);
    if (rc != FWI_OK || *launched) return rc;
    rc = try_launch<H, 16, ADJ>(p, st, launched);
    if (rc != FWI_OK || *launched) return rc;
    return try_launch<H, 24, ADJ>(p, st, launched);
}

template <int H>
int try_h(const SmallParams &p, cudaStream_t st, bool *launched) {
    *launched = false;
    return p.mode == 2 ? try_cpt<H, true>(p, st, launched) : try_cpt<H, false>(p, st, launched);
}

}  // namespace

// Runs the whole time loop in one cluster-resident kernel when the problem fits; *launched says so.
int small_run(const SmallParams &p, void *stream, bool *launched) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (p.h) {
    case 1: return try_h<1>(p, st, launched);
    case 2: return try_h<2>(p, st, launched);
    case 3: return try_h<3>(p, st, launched);
    default: return try_h<4>(p, st, launched);
    }
}

}  // namespace fwi
