// 2-D P-SV elastic velocity-stress propagator (staggered grid, C-PML) for sm_100a: forward and
// exact discrete adjoint.
//
// Replaces what the reference reaches through pyapi_denise.Denise.grad(model, src, rec)
// (reference models/networks.py:9173, gradients read back at :9185-9188): the elastic forward
// modelling, the adjoint modelling and the correlation into gradients.  DENISE does this on CPUs
// with MPI and files; here one call advances all shots of a batch on the GPU, tensors in / tensors
// out.  Scheme and notation: oracle/elastic_ref.py (the CPU restatement the tests check against).
//
// Fields per shot on the padded grid [nzp][nyp]: vy, vz, sxx, szz, syz and eight C-PML memory
// variables (one per spatial derivative).  A time step is two kernels, F1 (velocities from stress
// derivatives) and F2 (stresses from velocity derivatives): each reads only fields the other one
// writes, so both update in place.  With `snap` the forward pass stores the five fields the
// imaging condition needs (fy, fz: the force terms; e1, e2, e3: the strain rates).
// The adjoint of a step is four kernels: B2a / B1a are pointwise (adjoint of the parameter
// multiplications and of the memory-variable recursions, imaging sums, and the fields Y that the
// transposed difference operators act on), B2b / B1b gather (D+)^T = -D-, (D-)^T = -D+ of Y.
// One thread per cell; these kernels are HBM/L2 bound and use no tensor cores.
#include <stdlib.h>

#include "common.h"

namespace {

struct EParams {
    int nzp, nyp, S, h, pml;
    long long cells;            // nzp * nyp
    float dt;
    float cz[4], cy[4];         // staggered coefficients c_k / dz, c_k / dy
    const float *l2m, *lam, *mu, *by, *bz;
    const float *pz, *py;       // profiles [4][nzp] / [4][nyp]: a, b at integer nodes, a, b at half nodes
    float *f[13];               // vy vz sxx szz syz psi0..psi7 (adjoint pass: their adjoints)
    float *snap;                // this step's [5][S][cells]: fy fz e1 e2 e3 (NULL: not stored / read)
    float *scratch;             // [4][S][cells]
    float *grad;                // [5][S][cells]: l2m lam mu by bz images
    float *psi_in[8], *psi_out[8];  // fused kernels: memory variables read / written this step (ping-pong)
    float *fo[5];                   // fused forward: vy vz sxx szz syz written this step (p.f[0..4] are read)
    int fs_row;                     // free surface: first model row (stress imaging there, nothing above it is updated); 0 = none
};

// Every kernel of the time loop is launched with programmatic stream serialization: its blocks
// may start (index arithmetic, parameter loads) while the previous kernel drains, and wait here
// before touching any field the previous kernel writes.
__device__ __forceinline__ void pdl_wait() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // the next kernel may become resident early
    asm volatile("griddepcontrol.wait;" ::: "memory");                // ... but nobody reads before the previous grid is done
}

template <typename... KArgs, typename... Args>
inline void launch_pdl_smem(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

constexpr int VY = 0, VZ = 1, SXX = 2, SZZ = 3, SYZ = 4, PSI = 5;

// Wavefields are read with plain (coherent) loads everywhere in this file, never with __ldg.  ld.global.nc declares the
// data read-only for the lifetime of the grid, and both nvcc and ptxas move such loads above griddepcontrol.wait, asm
// "memory" clobber or not (an address made to depend on an opaque zero produced after the wait does not help: ptxas
// folds the zero) - but under programmatic dependent launch the previous kernel of the time loop may still be writing
// the field then.  Seen in the SASS of one build of the free-surface F2 (six LDG.E.CONSTANT of vy / vz ahead of ACQBULK)
// and as traces that depended on timing.  __ldg stays on what no kernel of the time loop writes: material grids,
// profiles, index lists, amplitudes.
template <int H>
__device__ __forceinline__ float dplus(const float *f, long long i, long long st, const float *c) {
    float s = c[0] * (f[i + st] - f[i]);
#pragma unroll
    for (int k = 2; k <= H; ++k) s = s + c[k - 1] * (f[i + k * st] - f[i - (k - 1) * st]);
    return s;
}
template <int H>
__device__ __forceinline__ float dminus(const float *f, long long i, long long st, const float *c) {
    float s = c[0] * (f[i] - f[i - st]);
#pragma unroll
    for (int k = 2; k <= H; ++k) s = s + c[k - 1] * (f[i + (k - 1) * st] - f[i - k * st]);
    return s;
}

// z-derivatives of szz / syz next to a free surface at row fs (stress imaging: the values above the surface are minus
// their mirror images; szz mirrors about the row itself, syz - half a cell lower - about the surface between its rows)
template <int H>
__device__ __forceinline__ float dplus_z_szz_fs(const float *f, int i, int j, int nyp, int fs, const float *c) {
    auto at = [&](int r) { return r >= fs ? f[(long long)r * nyp + j] : -f[(long long)(2 * fs - r) * nyp + j]; };
    float s = c[0] * (at(i + 1) - at(i));
#pragma unroll
    for (int k = 2; k <= H; ++k) s = s + c[k - 1] * (at(i + k) - at(i - (k - 1)));
    return s;
}
template <int H>
__device__ __forceinline__ float dminus_z_syz_fs(const float *f, int i, int j, int nyp, int fs, const float *c) {
    auto at = [&](int r) { return r >= fs ? f[(long long)r * nyp + j] : -f[(long long)(2 * fs - r - 1) * nyp + j]; };
    float s = c[0] * (at(i) - at(i - 1));
#pragma unroll
    for (int k = 2; k <= H; ++k) s = s + c[k - 1] * (at(i + (k - 1)) - at(i - k));
    return s;
}

#define ELASTIC_CELL()                                                              \
    const int j = blockIdx.x * blockDim.x + threadIdx.x + p.h;                      \
    const int i = blockIdx.y * blockDim.y + threadIdx.y + p.h;                      \
    const int s = blockIdx.z;                                                       \
    if (i >= p.nzp - p.h || j >= p.nyp - p.h || i < p.fs_row) return;               \
    const long long c0 = (long long)i * p.nyp + j;                                  \
    const long long o = (long long)s * p.cells + c0;                                \
    const long long fs = (long long)p.S * p.cells;                                  \
    const bool ypml = p.pml > 0 && (j < p.h + p.pml || j >= p.nyp - p.h - p.pml - 1); \
    const bool zpml = p.pml > 0 && (i < p.h + p.pml || i >= p.nzp - p.h - p.pml - 1); \
    (void)fs; (void)ypml; (void)zpml;                                               \
    pdl_wait();

// profile accessors: row 0 a, 1 b (integer nodes), 2 a, 3 b (half nodes)
#define AZ(i) __ldg(p.pz + (i))
#define BZ(i) __ldg(p.pz + p.nzp + (i))
#define AZH(i) __ldg(p.pz + 2 * p.nzp + (i))
#define BZH(i) __ldg(p.pz + 3 * p.nzp + (i))
#define AY(j) __ldg(p.py + (j))
#define BY(j) __ldg(p.py + p.nyp + (j))
#define AYH(j) __ldg(p.py + 2 * p.nyp + (j))
#define BYH(j) __ldg(p.py + 3 * p.nyp + (j))

// ---- forward ------------------------------------------------------------------------------
#ifdef FWI_EL_MINBLK         // minimum blocks per SM of F1 / F2 (a register budget): a tuning knob of scripts/dev_variants.sh
#define FWI_EL_FB 256, FWI_EL_MINBLK
#else
#define FWI_EL_FB 256
#endif
// Every operand of a cell is requested before the first one is used: the memory-variable stores of the PML branches
// (which may alias anything, as far as the compiler knows) would otherwise keep the loads of the updated fields and of the
// material grids behind them, and a thread would pay two or three memory round trips in a row instead of one
// (profiles/r02e_el_fwd_full.txt: 60 % of the stall samples sat on those late loads).
template <int H, bool SAVE, bool FS>
__global__ void __launch_bounds__(FWI_EL_FB) elastic_f1(const EParams p) {
    ELASTIC_CELL();
    const float *sxx = p.f[SXX] + (long long)s * p.cells, *szz = p.f[SZZ] + (long long)s * p.cells,
                *syz = p.f[SYZ] + (long long)s * p.cells;
    const float vy0 = p.f[VY][o], vz0 = p.f[VZ][o];
    const float by = __ldg(p.by + c0), bz = __ldg(p.bz + c0);
    float ps0 = 0.f, ps1 = 0.f, ps2 = 0.f, ps3 = 0.f;
    if (ypml) { ps0 = p.f[PSI + 0][o]; ps2 = p.f[PSI + 2][o]; }
    if (zpml) { ps1 = p.f[PSI + 1][o]; ps3 = p.f[PSI + 3][o]; }
    const bool near_fs = FS && i < p.fs_row + H;      // (FS is a template flag: the grids without a free surface keep their code)
    float d1 = dplus<H>(sxx, c0, 1, p.cy);
    float d2 = near_fs ? dminus_z_syz_fs<H>(syz, i, j, p.nyp, p.fs_row, p.cz) : dminus<H>(syz, c0, p.nyp, p.cz);
    float d3 = dminus<H>(syz, c0, 1, p.cy);
    float d4 = near_fs ? dplus_z_szz_fs<H>(szz, i, j, p.nyp, p.fs_row, p.cz) : dplus<H>(szz, c0, p.nyp, p.cz);
    if (ypml) {
        float q0 = BYH(j) * ps0 + AYH(j) * d1; p.f[PSI + 0][o] = q0; d1 = d1 + q0;
        float q2 = BY(j) * ps2 + AY(j) * d3; p.f[PSI + 2][o] = q2; d3 = d3 + q2;
    }
    if (zpml) {
        float q1 = BZ(i) * ps1 + AZ(i) * d2; p.f[PSI + 1][o] = q1; d2 = d2 + q1;
        float q3 = BZH(i) * ps3 + AZH(i) * d4; p.f[PSI + 3][o] = q3; d4 = d4 + q3;
    }
    const float fy = d1 + d2, fz = d3 + d4;
    p.f[VY][o] = vy0 + (p.dt * by) * fy;
    p.f[VZ][o] = vz0 + (p.dt * bz) * fz;
    if (SAVE) {
        __stcs(p.snap + o, fy);
        __stcs(p.snap + fs + o, fz);
    }
}

template <int H, bool SAVE, bool FS>
__global__ void __launch_bounds__(FWI_EL_FB) elastic_f2(const EParams p) {
    ELASTIC_CELL();
    const float *vy = p.f[VY] + (long long)s * p.cells, *vz = p.f[VZ] + (long long)s * p.cells;
    const float sxx0 = p.f[SXX][o], szz0 = p.f[SZZ][o], syz0 = p.f[SYZ][o];       // (all loads first: see elastic_f1)
    const float l2m = __ldg(p.l2m + c0), lam = __ldg(p.lam + c0), mu = __ldg(p.mu + c0);
    float ps4 = 0.f, ps5 = 0.f, ps6 = 0.f, ps7 = 0.f;
    if (ypml) { ps4 = p.f[PSI + 4][o]; ps7 = p.f[PSI + 7][o]; }
    if (zpml) { ps5 = p.f[PSI + 5][o]; ps6 = p.f[PSI + 6][o]; }
    float d5 = dminus<H>(vy, c0, 1, p.cy), d6 = dminus<H>(vz, c0, p.nyp, p.cz);
    float d7 = dplus<H>(vy, c0, p.nyp, p.cz), d8 = dplus<H>(vz, c0, 1, p.cy);
    if (ypml) {
        float q4 = BY(j) * ps4 + AY(j) * d5; p.f[PSI + 4][o] = q4; d5 = d5 + q4;
        float q7 = BYH(j) * ps7 + AYH(j) * d8; p.f[PSI + 7][o] = q7; d8 = d8 + q7;
    }
    if (zpml) {
        float q5 = BZ(i) * ps5 + AZ(i) * d6; p.f[PSI + 5][o] = q5; d6 = d6 + q5;
        float q6 = BZH(i) * ps6 + AZH(i) * d7; p.f[PSI + 6][o] = q6; d7 = d7 + q6;
    }
    const float e1 = d5, e2 = d6, e3 = d7 + d8;
    if (FS && i == p.fs_row) {
        // on the free surface szz stays zero, which removes e2 from the sxx update: (l2m - lam^2 / l2m) e1
        p.f[SXX][o] = sxx0 + p.dt * ((l2m - lam * lam / l2m) * e1);
    } else {
        p.f[SXX][o] = sxx0 + p.dt * (l2m * e1 + lam * e2);
        p.f[SZZ][o] = szz0 + p.dt * (lam * e1 + l2m * e2);
    }
    p.f[SYZ][o] = syz0 + (p.dt * mu) * e3;
    if (SAVE) {
        __stcs(p.snap + 2 * fs + o, e1);
        __stcs(p.snap + 3 * fs + o, e2);
        __stcs(p.snap + 4 * fs + o, e3);
    }
}

// ---- adjoint ------------------------------------------------------------------------------
// adjoint of F2, pointwise part (stress nodes)
__global__ void __launch_bounds__(256) elastic_b2a(const EParams p) {
    ELASTIC_CELL();
    const float asxx = p.f[SXX][o], aszz = p.f[SZZ][o], asyz = p.f[SYZ][o];
    const float l2m = __ldg(p.l2m + c0), lam = __ldg(p.lam + c0), mu = __ldg(p.mu + c0);
    const float ae1 = p.dt * (l2m * asxx + lam * aszz), ae2 = p.dt * (lam * asxx + l2m * aszz), ae3 = (p.dt * mu) * asyz;
    const float e1 = __ldcs(p.snap + 2 * fs + o), e2 = __ldcs(p.snap + 3 * fs + o), e3 = __ldcs(p.snap + 4 * fs + o);
    p.grad[o] = p.grad[o] + p.dt * (asxx * e1 + aszz * e2);
    p.grad[fs + o] = p.grad[fs + o] + p.dt * (asxx * e2 + aszz * e1);
    p.grad[2 * fs + o] = p.grad[2 * fs + o] + p.dt * (asyz * e3);
    float y1 = ae1, y2 = ae2, y3 = ae3, y4 = ae3;
    if (ypml) {
        const float x4 = ae1 + p.f[PSI + 4][o]; p.f[PSI + 4][o] = BY(j) * x4; y1 = ae1 + AY(j) * x4;
        const float x7 = ae3 + p.f[PSI + 7][o]; p.f[PSI + 7][o] = BYH(j) * x7; y4 = ae3 + AYH(j) * x7;
    }
    if (zpml) {
        const float x5 = ae2 + p.f[PSI + 5][o]; p.f[PSI + 5][o] = BZ(i) * x5; y2 = ae2 + AZ(i) * x5;
        const float x6 = ae3 + p.f[PSI + 6][o]; p.f[PSI + 6][o] = BZH(i) * x6; y3 = ae3 + AZH(i) * x6;
    }
    p.scratch[o] = y1;
    p.scratch[fs + o] = y2;
    p.scratch[2 * fs + o] = y3;
    p.scratch[3 * fs + o] = y4;
}

// adjoint of F2, transposed differences (velocity nodes):
//   avy += (D-y)^T Y1 + (D+z)^T Y3 = -D+y Y1 - D-z Y3;  avz += -D+z Y2 - D-y Y4
template <int H>
__global__ void __launch_bounds__(256) elastic_b2b(const EParams p) {
    ELASTIC_CELL();
    const float *y1 = p.scratch + (long long)s * p.cells, *y2 = y1 + fs, *y3 = y2 + fs, *y4 = y3 + fs;
    p.f[VY][o] = p.f[VY][o] - (dplus<H>(y1, c0, 1, p.cy) + dminus<H>(y3, c0, p.nyp, p.cz));
    p.f[VZ][o] = p.f[VZ][o] - (dplus<H>(y2, c0, p.nyp, p.cz) + dminus<H>(y4, c0, 1, p.cy));
}

// adjoint of F1, pointwise part (velocity nodes)
__global__ void __launch_bounds__(256) elastic_b1a(const EParams p) {
    ELASTIC_CELL();
    const float avy = p.f[VY][o], avz = p.f[VZ][o];
    const float agy = (p.dt * __ldg(p.by + c0)) * avy, agz = (p.dt * __ldg(p.bz + c0)) * avz;
    const float fy = __ldcs(p.snap + o), fz = __ldcs(p.snap + fs + o);
    p.grad[3 * fs + o] = p.grad[3 * fs + o] + p.dt * (avy * fy);
    p.grad[4 * fs + o] = p.grad[4 * fs + o] + p.dt * (avz * fz);
    float y1 = agy, y2 = agy, y3 = agz, y4 = agz;
    if (ypml) {
        const float x0 = agy + p.f[PSI + 0][o]; p.f[PSI + 0][o] = BYH(j) * x0; y1 = agy + AYH(j) * x0;
        const float x2 = agz + p.f[PSI + 2][o]; p.f[PSI + 2][o] = BY(j) * x2; y3 = agz + AY(j) * x2;
    }
    if (zpml) {
        const float x1 = agy + p.f[PSI + 1][o]; p.f[PSI + 1][o] = BZ(i) * x1; y2 = agy + AZ(i) * x1;
        const float x3 = agz + p.f[PSI + 3][o]; p.f[PSI + 3][o] = BZH(i) * x3; y4 = agz + AZH(i) * x3;
    }
    p.scratch[o] = y1;
    p.scratch[fs + o] = y2;
    p.scratch[2 * fs + o] = y3;
    p.scratch[3 * fs + o] = y4;
}

// adjoint of F1, transposed differences (stress nodes):
//   asxx += (D+y)^T Y1 = -D-y Y1;  asyz += (D-z)^T Y2 + (D-y)^T Y3 = -D+z Y2 - D+y Y3;  aszz += -D-z Y4
template <int H>
__global__ void __launch_bounds__(256) elastic_b1b(const EParams p) {
    ELASTIC_CELL();
    const float *y1 = p.scratch + (long long)s * p.cells, *y2 = y1 + fs, *y3 = y2 + fs, *y4 = y3 + fs;
    p.f[SXX][o] = p.f[SXX][o] - dminus<H>(y1, c0, 1, p.cy);
    p.f[SYZ][o] = p.f[SYZ][o] - (dplus<H>(y2, c0, p.nyp, p.cz) + dplus<H>(y3, c0, 1, p.cy));
    p.f[SZZ][o] = p.f[SZZ][o] - dminus<H>(y4, c0, p.nyp, p.cz);
}

// ---- fused adjoint kernels ------------------------------------------------------------------
// B2 = B2a + B2b and B1 = B1a + B1b in one kernel each: a CTA evaluates the pointwise stage on its
// 64 x 32 tile plus a halo of H cells, keeps the four fields Y the transposed differences act on in
// shared memory, and gathers them for its own cells.  The `scratch` round trip of the two-kernel
// form (4 fields written and read back per half step) disappears.  Halo cells need the memory
// variables of the neighbouring tile as they were BEFORE this step, so those are double-buffered
// (psi_in / psi_out, swapped every step; only PML cells ever touch them).
constexpr int kTX = 64, kTZ = 32;

template <int H>
__device__ __forceinline__ float sm_dplus(const float *y, int st, const float *c) {
    float s = c[0] * (y[st] - y[0]);
#pragma unroll
    for (int k = 2; k <= H; ++k) s = s + c[k - 1] * (y[k * st] - y[-(k - 1) * st]);
    return s;
}
template <int H>
__device__ __forceinline__ float sm_dminus(const float *y, int st, const float *c) {
    float s = c[0] * (y[0] - y[-st]);
#pragma unroll
    for (int k = 2; k <= H; ++k) s = s + c[k - 1] * (y[(k - 1) * st] - y[-k * st]);
    return s;
}

template <int H, int NT, bool FS>
__global__ void __launch_bounds__(NT, NT <= 512 ? 1024 / NT : 1) elastic_b2_fused(const EParams p) {
    constexpr int SX = kTX + 2 * H, SZ = kTZ + 2 * H;
    constexpr int N1 = (SX * SZ + NT - 1) / NT, N2 = (kTX * kTZ) / NT;
    static_assert((kTX * kTZ) % NT == 0, "tile must be a multiple of the CTA size");
    __shared__ float Y[4][SZ * SX];
    const int s = blockIdx.z;
    const int zs = FS ? p.fs_row : p.h;                                     // first computational row
    const int j0 = blockIdx.x * kTX + p.h, i0 = blockIdx.y * kTZ + zs;      // first own cell
    const long long fs = (long long)p.S * p.cells, so = (long long)s * p.cells;
    pdl_wait();
    // stage 1: the fields Y the transposed differences act on, for the tile and its halo (loads only,
    // apart from the memory variables of own PML cells)
#pragma unroll
    for (int q = 0; q < N1; ++q) {
        const int t = threadIdx.x + q * NT;
        if (t >= SX * SZ) break;
        const int lz = t / SX, lx = t - lz * SX;
        const int i = i0 - H + lz, j = j0 - H + lx;
        float y1 = 0.f, y2 = 0.f, y3 = 0.f, y4 = 0.f;
        if (i >= zs && i < p.nzp - p.h && j >= p.h && j < p.nyp - p.h) {
            const long long c0 = (long long)i * p.nyp + j, o = so + c0;
            const bool ypml = p.pml > 0 && (j < p.h + p.pml || j >= p.nyp - p.h - p.pml - 1);
            const bool zpml = p.pml > 0 && (i < p.h + p.pml || i >= p.nzp - p.h - p.pml - 1);
            const float asxx = p.f[SXX][o], aszz = p.f[SZZ][o], asyz = p.f[SYZ][o];
            const float l2m = __ldg(p.l2m + c0), lam = __ldg(p.lam + c0), mu = __ldg(p.mu + c0);
            float ae1 = p.dt * (l2m * asxx + lam * aszz), ae2 = p.dt * (lam * asxx + l2m * aszz);
            const float ae3 = (p.dt * mu) * asyz;
            if (FS && i == p.fs_row) {                // free-surface row: sxx += dt (l2m - lam^2 / l2m) e1, szz is not updated
                ae1 = p.dt * ((l2m - lam * lam / l2m) * asxx);
                ae2 = 0.f;
            }
            y1 = ae1; y2 = ae2; y3 = ae3; y4 = ae3;
            if (ypml || zpml) {
                const bool own = lz >= H && lz < H + kTZ && lx >= H && lx < H + kTX;
                if (ypml) {
                    const float x4 = ae1 + p.psi_in[4][o]; y1 = ae1 + AY(j) * x4;
                    const float x7 = ae3 + p.psi_in[7][o]; y4 = ae3 + AYH(j) * x7;
                    if (own) { p.psi_out[4][o] = BY(j) * x4; p.psi_out[7][o] = BYH(j) * x7; }
                }
                if (zpml) {
                    const float x5 = ae2 + p.psi_in[5][o]; y2 = ae2 + AZ(i) * x5;
                    const float x6 = ae3 + p.psi_in[6][o]; y3 = ae3 + AZH(i) * x6;
                    if (own) { p.psi_out[5][o] = BZ(i) * x5; p.psi_out[6][o] = BZH(i) * x6; }
                }
            }
        }
        Y[0][t] = y1; Y[1][t] = y2; Y[2][t] = y3; Y[3][t] = y4;
    }
    // stage 2 operands that do not depend on stage 1: fetched before the barrier so they are in flight
    float sx[N2], sz[N2], sy[N2], e1[N2], e2[N2], e3[N2], g0[N2], g1[N2], g2[N2], vy[N2], vz[N2];
    long long oo[N2];
#pragma unroll
    for (int q = 0; q < N2; ++q) {
        const int t = threadIdx.x + q * NT;
        const int tz = t / kTX, tx = t - tz * kTX;
        const int i = i0 + tz, j = j0 + tx;
        const bool in = i < p.nzp - p.h && j < p.nyp - p.h;
        const long long o = so + (long long)(in ? i : p.h) * p.nyp + (in ? j : p.h);
        oo[q] = in ? o : -1;
        sx[q] = p.f[SXX][o]; sz[q] = p.f[SZZ][o]; sy[q] = p.f[SYZ][o];
        e1[q] = __ldcs(p.snap + 2 * fs + o); e2[q] = __ldcs(p.snap + 3 * fs + o); e3[q] = __ldcs(p.snap + 4 * fs + o);
        g0[q] = p.grad[o]; g1[q] = p.grad[fs + o]; g2[q] = p.grad[2 * fs + o];
        vy[q] = p.f[VY][o]; vz[q] = p.f[VZ][o];
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < N2; ++q) {
        if (oo[q] < 0) continue;
        const long long o = oo[q];
        const int t = threadIdx.x + q * NT;
        const int tz = t / kTX, tx = t - tz * kTX;
        const int l = (tz + H) * SX + tx + H;
        if (FS && i0 + tz == p.fs_row) {
            // images of the free-surface row: d/d l2m and d/d lam of (l2m - lam^2 / l2m) e1
            const long long c0 = o - so;
            const float l2m = __ldg(p.l2m + c0), lam = __ldg(p.lam + c0);
            const float r = lam / l2m;
            p.grad[o] = g0[q] + p.dt * (sx[q] * e1[q]) * (1.0f + r * r);
            p.grad[fs + o] = g1[q] - p.dt * (sx[q] * e1[q]) * (2.0f * r);
        } else {
            p.grad[o] = g0[q] + p.dt * (sx[q] * e1[q] + sz[q] * e2[q]);
            p.grad[fs + o] = g1[q] + p.dt * (sx[q] * e2[q] + sz[q] * e1[q]);
        }
        p.grad[2 * fs + o] = g2[q] + p.dt * (sy[q] * e3[q]);
        p.f[VY][o] = vy[q] - (sm_dplus<H>(&Y[0][l], 1, p.cy) + sm_dminus<H>(&Y[2][l], SX, p.cz));
        p.f[VZ][o] = vz[q] - (sm_dplus<H>(&Y[1][l], SX, p.cz) + sm_dminus<H>(&Y[3][l], 1, p.cy));
    }
}

template <int H, int NT, bool FS>
__global__ void __launch_bounds__(NT, NT <= 512 ? 1024 / NT : 1) elastic_b1_fused(const EParams p) {
    constexpr int SX = kTX + 2 * H, SZ = kTZ + 2 * H;
    constexpr int N1 = (SX * SZ + NT - 1) / NT, N2 = (kTX * kTZ) / NT;
    __shared__ float Y[4][SZ * SX];
    const int s = blockIdx.z;
    const int zs = FS ? p.fs_row : p.h;
    const int j0 = blockIdx.x * kTX + p.h, i0 = blockIdx.y * kTZ + zs;
    const long long fs = (long long)p.S * p.cells, so = (long long)s * p.cells;
    pdl_wait();
#pragma unroll
    for (int q = 0; q < N1; ++q) {
        const int t = threadIdx.x + q * NT;
        if (t >= SX * SZ) break;
        const int lz = t / SX, lx = t - lz * SX;
        const int i = i0 - H + lz, j = j0 - H + lx;
        float y1 = 0.f, y2 = 0.f, y3 = 0.f, y4 = 0.f;
        if (i >= zs && i < p.nzp - p.h && j >= p.h && j < p.nyp - p.h) {
            const long long c0 = (long long)i * p.nyp + j, o = so + c0;
            const bool ypml = p.pml > 0 && (j < p.h + p.pml || j >= p.nyp - p.h - p.pml - 1);
            const bool zpml = p.pml > 0 && (i < p.h + p.pml || i >= p.nzp - p.h - p.pml - 1);
            const float avy = p.f[VY][o], avz = p.f[VZ][o];
            const float agy = (p.dt * __ldg(p.by + c0)) * avy, agz = (p.dt * __ldg(p.bz + c0)) * avz;
            y1 = agy; y2 = agy; y3 = agz; y4 = agz;
            if (ypml || zpml) {
                const bool own = lz >= H && lz < H + kTZ && lx >= H && lx < H + kTX;
                if (ypml) {
                    const float x0 = agy + p.psi_in[0][o]; y1 = agy + AYH(j) * x0;
                    const float x2 = agz + p.psi_in[2][o]; y3 = agz + AY(j) * x2;
                    if (own) { p.psi_out[0][o] = BYH(j) * x0; p.psi_out[2][o] = BY(j) * x2; }
                }
                if (zpml) {
                    const float x1 = agy + p.psi_in[1][o]; y2 = agy + AZ(i) * x1;
                    const float x3 = agz + p.psi_in[3][o]; y4 = agz + AZH(i) * x3;
                    if (own) { p.psi_out[1][o] = BZ(i) * x1; p.psi_out[3][o] = BZH(i) * x3; }
                }
            }
        }
        Y[0][t] = y1; Y[1][t] = y2; Y[2][t] = y3; Y[3][t] = y4;
    }
    float vy[N2], vz[N2], fy[N2], fz[N2], g3[N2], g4[N2], sx[N2], sy[N2], sz[N2];
    long long oo[N2];
#pragma unroll
    for (int q = 0; q < N2; ++q) {
        const int t = threadIdx.x + q * NT;
        const int tz = t / kTX, tx = t - tz * kTX;
        const int i = i0 + tz, j = j0 + tx;
        const bool in = i < p.nzp - p.h && j < p.nyp - p.h;
        const long long o = so + (long long)(in ? i : p.h) * p.nyp + (in ? j : p.h);
        oo[q] = in ? o : -1;
        vy[q] = p.f[VY][o]; vz[q] = p.f[VZ][o];
        fy[q] = __ldcs(p.snap + o); fz[q] = __ldcs(p.snap + fs + o);
        g3[q] = p.grad[3 * fs + o]; g4[q] = p.grad[4 * fs + o];
        sx[q] = p.f[SXX][o]; sy[q] = p.f[SYZ][o]; sz[q] = p.f[SZZ][o];
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < N2; ++q) {
        if (oo[q] < 0) continue;
        const long long o = oo[q];
        const int t = threadIdx.x + q * NT;
        const int tz = t / kTX, tx = t - tz * kTX;
        const int l = (tz + H) * SX + tx + H;
        p.grad[3 * fs + o] = g3[q] + p.dt * (vy[q] * fy[q]);
        p.grad[4 * fs + o] = g4[q] + p.dt * (vz[q] * fz[q]);
        float nsy = sy[q] - (sm_dplus<H>(&Y[1][l], SX, p.cz) + sm_dplus<H>(&Y[2][l], 1, p.cy));
        float nsz = sz[q] - sm_dminus<H>(&Y[3][l], SX, p.cz);
        if (FS && blockIdx.y == 0 && tz <= H) {
            // free surface: F1 read the stresses above the surface as minus their mirror images, so a row next to the
            // surface also receives, with the opposite sign, what the transposed differences deliver to its mirror row
            // (the Y of rows above the surface are zero: those rows are not part of the system)
            const int col = tx + H;
            auto y_at = [&](const int f, const int lz) { return lz >= H ? Y[f][lz * SX + col] : 0.f; };   // smem row lz <-> grid row i0 - H + lz
            const int gz = H - tz - 1;                      // smem row of the ghost row 2 fs - r - 1 of syz (r = fs + tz)
            if (gz >= 0) {
                float t = p.cz[0] * (y_at(1, gz + 1) - y_at(1, gz));
#pragma unroll
                for (int k = 2; k <= H; ++k) t = t + p.cz[k - 1] * (y_at(1, gz + k) - y_at(1, gz - (k - 1)));
                nsy = nsy + t;                              // + (D+z Y2)[ghost]
            }
            const int hz = H - tz;                          // smem row of the ghost row 2 fs - r of szz
            if (tz >= 1 && hz >= 0) {
                float t = p.cz[0] * (y_at(3, hz) - y_at(3, hz - 1));
#pragma unroll
                for (int k = 2; k <= H; ++k) t = t + p.cz[k - 1] * (y_at(3, hz + (k - 1)) - y_at(3, hz - k));
                nsz = nsz + t;                              // + (D-z Y4)[ghost]
            }
        }
        p.f[SXX][o] = sx[q] - sm_dminus<H>(&Y[0][l], 1, p.cy);
        p.f[SYZ][o] = nsy;
        p.f[SZZ][o] = nsz;
    }
}

// ---- fused forward kernel -----------------------------------------------------------------------
// F1 + F2 in one kernel: a CTA loads the three stresses of its 64 x 32 tile plus 2H halo cells into
// shared memory, computes the new velocities on the tile plus H halo cells into shared memory (own
// cells also go to HBM, with the force terms of the snapshot), then the new stresses of its own cells
// from the shared velocities.  Each field crosses HBM once per step (80 B per cell-update instead
// of ~100 B).  Halo cells recompute what the neighbouring CTA also computes, from the values BEFORE the
// step, so everything the halo reads is double-buffered: p.f[0..4] / psi_in are read, p.fo / psi_out
// written, and the host swaps them every step.  psi4..7 belong to own cells only: in place.
template <int H, int NT, bool SAVE>
__global__ void __launch_bounds__(NT, NT <= 512 ? 1024 / NT : 1) elastic_f_fused(const EParams p) {
    constexpr int SX0 = kTX + 4 * H, SZ0 = kTZ + 4 * H;      // stress tile
    constexpr int SX1 = kTX + 2 * H, SZ1 = kTZ + 2 * H;      // velocity tile
    constexpr int N0 = (SX0 * SZ0 + NT - 1) / NT, N1 = (SX1 * SZ1 + NT - 1) / NT, N2 = (kTX * kTZ) / NT;
    extern __shared__ __align__(16) float fsm[];
    float *Sxx = fsm, *Szz = fsm + SX0 * SZ0, *Syz = fsm + 2 * SX0 * SZ0;
    float *Vy = fsm + 3 * SX0 * SZ0, *Vz = Vy + SX1 * SZ1;
    const int s = blockIdx.z;
    const int j0 = blockIdx.x * kTX + p.h, i0 = blockIdx.y * kTZ + p.h;     // first own cell
    const long long fs = (long long)p.S * p.cells, so = (long long)s * p.cells;
    pdl_wait();
    // stage 0: stresses of the tile and 2H halo cells (zero outside the grid)
#pragma unroll
    for (int q = 0; q < N0; ++q) {
        const int t = threadIdx.x + q * NT;
        if (t >= SX0 * SZ0) break;
        const int lz = t / SX0, lx = t - lz * SX0;
        const int i = i0 - 2 * H + lz, j = j0 - 2 * H + lx;
        float a = 0.f, b = 0.f, c = 0.f;
        if (i >= 0 && i < p.nzp && j >= 0 && j < p.nyp) {
            const long long o = so + (long long)i * p.nyp + j;
            a = p.f[SXX][o]; b = p.f[SZZ][o]; c = p.f[SYZ][o];
        }
        Sxx[t] = a; Szz[t] = b; Syz[t] = c;
    }
    // operands of stage 1 that do not depend on stage 0: in flight across the barrier
    float pvy[N1], pvz[N1], pby[N1], pbz[N1];
#pragma unroll
    for (int q = 0; q < N1; ++q) {
        const int t = threadIdx.x + q * NT;
        const int lz = t / SX1, lx = t - lz * SX1;
        const int i = i0 - H + lz, j = j0 - H + lx;
        const bool in = t < SX1 * SZ1 && i >= p.h && i < p.nzp - p.h && j >= p.h && j < p.nyp - p.h;
        const long long c0 = (long long)(in ? i : p.h) * p.nyp + (in ? j : p.h);
        pvy[q] = p.f[VY][so + c0]; pvz[q] = p.f[VZ][so + c0];
        pby[q] = __ldg(p.by + c0); pbz[q] = __ldg(p.bz + c0);
    }
    __syncthreads();
    // stage 1: new velocities on the tile and H halo cells
#pragma unroll
    for (int q = 0; q < N1; ++q) {
        const int t = threadIdx.x + q * NT;
        if (t >= SX1 * SZ1) break;
        const int lz = t / SX1, lx = t - lz * SX1;
        const int i = i0 - H + lz, j = j0 - H + lx;
        float vy = 0.f, vz = 0.f;
        if (i >= p.h && i < p.nzp - p.h && j >= p.h && j < p.nyp - p.h) {
            const bool own = lz >= H && lz < H + kTZ && lx >= H && lx < H + kTX;
            const long long c0 = (long long)i * p.nyp + j, o = so + c0;
            const int l = (lz + H) * SX0 + lx + H;
            float d1 = sm_dplus<H>(Sxx + l, 1, p.cy), d2 = sm_dminus<H>(Syz + l, SX0, p.cz);
            float d3 = sm_dminus<H>(Syz + l, 1, p.cy), d4 = sm_dplus<H>(Szz + l, SX0, p.cz);
            const bool ypml = p.pml > 0 && (j < p.h + p.pml || j >= p.nyp - p.h - p.pml - 1);
            const bool zpml = p.pml > 0 && (i < p.h + p.pml || i >= p.nzp - p.h - p.pml - 1);
            if (ypml) {
                const float q0 = BYH(j) * p.psi_in[0][o] + AYH(j) * d1; d1 = d1 + q0;
                const float q2 = BY(j) * p.psi_in[2][o] + AY(j) * d3; d3 = d3 + q2;
                if (own) { p.psi_out[0][o] = q0; p.psi_out[2][o] = q2; }
            }
            if (zpml) {
                const float q1 = BZ(i) * p.psi_in[1][o] + AZ(i) * d2; d2 = d2 + q1;
                const float q3 = BZH(i) * p.psi_in[3][o] + AZH(i) * d4; d4 = d4 + q3;
                if (own) { p.psi_out[1][o] = q1; p.psi_out[3][o] = q3; }
            }
            const float fy = d1 + d2, fz = d3 + d4;
            vy = pvy[q] + (p.dt * pby[q]) * fy;
            vz = pvz[q] + (p.dt * pbz[q]) * fz;
            if (own) {
                p.fo[VY][o] = vy; p.fo[VZ][o] = vz;
                if (SAVE) { __stcs(p.snap + o, fy); __stcs(p.snap + fs + o, fz); }
            }
        }
        Vy[t] = vy; Vz[t] = vz;
    }
    float pl2m[N2], plam[N2], pmu[N2];
#pragma unroll
    for (int q = 0; q < N2; ++q) {
        const int t = threadIdx.x + q * NT;
        const int tz = t / kTX, tx = t - tz * kTX;
        const int i = min(i0 + tz, p.nzp - p.h - 1), j = min(j0 + tx, p.nyp - p.h - 1);
        const long long c0 = (long long)i * p.nyp + j;
        pl2m[q] = __ldg(p.l2m + c0); plam[q] = __ldg(p.lam + c0); pmu[q] = __ldg(p.mu + c0);
    }
    __syncthreads();
    // stage 2: new stresses of the own cells
#pragma unroll
    for (int q = 0; q < N2; ++q) {
        const int t = threadIdx.x + q * NT;
        const int tz = t / kTX, tx = t - tz * kTX;
        const int i = i0 + tz, j = j0 + tx;
        if (i >= p.nzp - p.h || j >= p.nyp - p.h) continue;
        const long long c0 = (long long)i * p.nyp + j, o = so + c0;
        const int l = (tz + H) * SX1 + tx + H;
        float d5 = sm_dminus<H>(Vy + l, 1, p.cy), d6 = sm_dminus<H>(Vz + l, SX1, p.cz);
        float d7 = sm_dplus<H>(Vy + l, SX1, p.cz), d8 = sm_dplus<H>(Vz + l, 1, p.cy);
        const bool ypml = p.pml > 0 && (j < p.h + p.pml || j >= p.nyp - p.h - p.pml - 1);
        const bool zpml = p.pml > 0 && (i < p.h + p.pml || i >= p.nzp - p.h - p.pml - 1);
        if (ypml) {
            const float q4 = BY(j) * p.f[PSI + 4][o] + AY(j) * d5; p.f[PSI + 4][o] = q4; d5 = d5 + q4;
            const float q7 = BYH(j) * p.f[PSI + 7][o] + AYH(j) * d8; p.f[PSI + 7][o] = q7; d8 = d8 + q7;
        }
        if (zpml) {
            const float q5 = BZ(i) * p.f[PSI + 5][o] + AZ(i) * d6; p.f[PSI + 5][o] = q5; d6 = d6 + q5;
            const float q6 = BZH(i) * p.f[PSI + 6][o] + AZH(i) * d7; p.f[PSI + 6][o] = q6; d7 = d7 + q6;
        }
        const float e1 = d5, e2 = d6, e3 = d7 + d8;
        const float l2m = pl2m[q], lam = plam[q], mu = pmu[q];
        const int ls = (tz + 2 * H) * SX0 + tx + 2 * H;
        p.fo[SXX][o] = Sxx[ls] + p.dt * (l2m * e1 + lam * e2);
        p.fo[SZZ][o] = Szz[ls] + p.dt * (lam * e1 + l2m * e2);
        p.fo[SYZ][o] = Syz[ls] + (p.dt * mu) * e3;
        if (SAVE) {
            __stcs(p.snap + 2 * fs + o, e1);
            __stcs(p.snap + 3 * fs + o, e2);
            __stcs(p.snap + 4 * fs + o, e3);
        }
    }
}

// ---- point kernels (one launch per time step: the injection that ends a step and the sampling that begins the next
// one touch different fields, so they share a launch; deterministic: points that share a cell are chained by `link`
// (bits 0..29: index of the next point in the same cell + 1, 0 = none; bit 30 set: not the first of its chain) and
// added by the first one in list order) ----
struct PointArgs {
    long long cells;
    int S;
    // sampling: out_a[t] = scale * (a[c] (+ b[c] if sum)) * (coef ? coef[cell] : 1); out_b[t] = scale * b[c] (two-component)
    int n_smp;
    const int *smp_cell;
    const float *sa, *sb, *s_coef;
    float *out_a, *out_b;
    int smp_sum;
    float s_scale;
    // image term of force sources (adjoint): g_img[c] += s_scale * img_amp[t] * sa[c] (chains of smp_link: one owner per cell)
    float *g_img;
    const float *img_amp;
    const int *smp_link;
    // injection: a[c] += scale * coef * amp_a[t]; b[c] += scale * amp_b[t] (b may be NULL)
    int n_inj;
    const int *inj_cell, *inj_link;
    float *ia, *ib;
    const float *amp_a, *amp_b, *i_coef;
    float i_scale;
};

__global__ void __launch_bounds__(128) elastic_points(const PointArgs q) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int n_i = q.S * q.n_inj, n_s = q.S * q.n_smp;
    pdl_wait();
    if (t < n_i) {
        int l = __ldg(q.inj_link + t);
        if (!(l & 0x40000000)) {
            const int s = t / q.n_inj;
            const int cell = __ldg(q.inj_cell + t);
            const long long c = (long long)s * q.cells + cell;
            const float k = q.i_scale * (q.i_coef ? __ldg(q.i_coef + cell) : 1.0f);
            float va = q.ia[c] + k * __ldg(q.amp_a + t), vb = q.ib ? q.ib[c] + k * __ldg(q.amp_b + t) : 0.f;
            while ((l & 0x3fffffff) != 0) {
                const int m = s * q.n_inj + (l & 0x3fffffff) - 1;
                va = va + k * __ldg(q.amp_a + m);
                if (q.ib) vb = vb + k * __ldg(q.amp_b + m);
                l = __ldg(q.inj_link + m);
            }
            q.ia[c] = va;
            if (q.ib) q.ib[c] = vb;
        }
    } else if (t < n_i + n_s) {
        const int u = t - n_i;
        const int s = u / q.n_smp;
        const int cell = __ldg(q.smp_cell + u);
        const long long c = (long long)s * q.cells + cell;
        const float k = q.s_scale * (q.s_coef ? __ldg(q.s_coef + cell) : 1.0f);
        const float a = q.sa[c];
        if (q.smp_sum) q.out_a[u] = k * (a + q.sb[c]);
        else {
            q.out_a[u] = k * a;
            if (q.out_b) q.out_b[u] = k * q.sb[c];
        }
        // force sources, adjoint: d(dt b s)/db at the source cell; the first point of a chain adds for all of them
        if (q.g_img) {
            int l = __ldg(q.smp_link + u);
            if (!(l & 0x40000000)) {
                float g = q.g_img[c] + (q.s_scale * __ldg(q.img_amp + u)) * a;
                while ((l & 0x3fffffff) != 0) {
                    const int m = s * q.n_smp + (l & 0x3fffffff) - 1;
                    g = g + (q.s_scale * __ldg(q.img_amp + m)) * a;
                    l = __ldg(q.smp_link + m);
                }
                q.g_img[c] = g;
            }
        }
    }
}

inline void launch_points(const PointArgs &q, cudaStream_t st) {
    const long long n = (long long)q.S * (q.n_inj + q.n_smp);
    if (n == 0) return;
    launch_pdl(elastic_points, dim3((unsigned)((n + 127) / 128)), dim3(128), st, q);
    fwi::count_launch();
}

int check(const fwi_elastic_desc *d) {
    FWI_CHECK_ARG(d != nullptr, "descriptor is NULL");
    FWI_CHECK_ARG(d->order == 2 || d->order == 4 || d->order == 6 || d->order == 8,
                  "order must be 2, 4, 6 or 8 (got %d)", d->order);
    const int h = d->order / 2;
    FWI_CHECK_ARG(d->pml_width >= 0, "pml_width must be >= 0");
    FWI_CHECK_ARG(d->nzp > 2 * (h + d->pml_width) && d->nyp > 2 * (h + d->pml_width), "padded grid %dx%d too small",
                  d->nzp, d->nyp);
    FWI_CHECK_ARG(d->n_shots >= 1 && d->nsteps >= 0 && d->n_src >= 0 && d->n_rec >= 0, "bad counts");
    FWI_CHECK_ARG(d->dz > 0 && d->dy > 0 && d->dt != 0, "dz, dy must be > 0 and dt != 0");
    return FWI_OK;
}

const double kStag[5][4] = {{0, 0, 0, 0},
                            {1.0, 0, 0, 0},
                            {9.0 / 8.0, -1.0 / 24.0, 0, 0},
                            {75.0 / 64.0, -25.0 / 384.0, 3.0 / 640.0, 0},
                            {1225.0 / 1024.0, -245.0 / 3072.0, 49.0 / 5120.0, -5.0 / 7168.0}};

void fill(EParams &p, const fwi_elastic_desc *d, const fwi_elastic_buffers *b) {
    p.nzp = d->nzp; p.nyp = d->nyp; p.S = d->n_shots; p.h = d->order / 2; p.pml = d->pml_width;
    p.cells = (long long)d->nzp * d->nyp;
    p.dt = (float)d->dt;
    for (int k = 0; k < 4; ++k) {
        p.cz[k] = (float)(kStag[p.h][k] * (1.0 / d->dz));
        p.cy[k] = (float)(kStag[p.h][k] * (1.0 / d->dy));
    }
    p.l2m = b->l2m; p.lam = b->lam; p.mu = b->mu; p.by = b->by; p.bz = b->bz;
    p.pz = b->prof_z; p.py = b->prof_y;
    const long long fld = (long long)d->n_shots * p.cells;
    for (int k = 0; k < 13; ++k) p.f[k] = b->wave + k * fld;
    p.scratch = b->scratch; p.grad = b->grad; p.snap = nullptr;
    p.fs_row = d->free_surface ? d->pml_width + d->order / 2 : 0;
}

// thread-block shape of the stencil kernels (x along y, the fast axis); tuning knobs for scripts/
dim3 block_shape() {
    static const int bx = getenv("FWI_B200_EL_BX") ? atoi(getenv("FWI_B200_EL_BX")) : 32;
    static const int by = getenv("FWI_B200_EL_BY") ? atoi(getenv("FWI_B200_EL_BY")) : 8;
    return dim3(bx, by);
}

#define ELASTIC_LAUNCH_H(kern, ...)                                      \
    switch (p.h) {                                                       \
    case 1: launch_pdl(kern<1, ##__VA_ARGS__>, grid, block, st, p); break;    \
    case 2: launch_pdl(kern<2, ##__VA_ARGS__>, grid, block, st, p); break;    \
    case 3: launch_pdl(kern<3, ##__VA_ARGS__>, grid, block, st, p); break;    \
    default: launch_pdl(kern<4, ##__VA_ARGS__>, grid, block, st, p); break;   \
    }

}  // namespace

extern "C" int64_t fwi_elastic_wave_floats(const fwi_elastic_desc *d) {
    if (check(d) != FWI_OK) return -1;
    return 22ll * d->n_shots * d->nzp * d->nyp;      // 13 fields + the second copy of vy vz sxx szz syz psi0..3 (fused forward)
}

namespace {
template <int H, bool SAVE>
void launch_f_fused(const EParams &p, dim3 grid, cudaStream_t st) {
    constexpr size_t smem = (3 * (kTX + 4 * H) * (kTZ + 4 * H) + 2 * (kTX + 2 * H) * (kTZ + 2 * H)) * sizeof(float);
    auto kern = elastic_f_fused<H, 512, SAVE>;
    static bool done[64] = {};           // function attributes are per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !done[dev]) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (dev >= 0 && dev < 64) done[dev] = true;
    }
    launch_pdl_smem(kern, grid, dim3(512), smem, st, p);
}
template <bool SAVE>
void launch_f_fused_h(const EParams &p, dim3 grid, cudaStream_t st) {
    switch (p.h) {
    case 1: launch_f_fused<1, SAVE>(p, grid, st); break;
    case 2: launch_f_fused<2, SAVE>(p, grid, st); break;
    case 3: launch_f_fused<3, SAVE>(p, grid, st); break;
    default: launch_f_fused<4, SAVE>(p, grid, st); break;
    }
}
}  // namespace

extern "C" int fwi_elastic_forward(const fwi_elastic_desc *d, const fwi_elastic_buffers *b, void *stream) {
    int rc = check(d);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(b && b->l2m && b->lam && b->mu && b->by && b->bz && b->prof_z && b->prof_y && b->wave,
                  "parameter grids, profiles and wave workspace must not be NULL");
    FWI_CHECK_ARG(d->n_src == 0 || (b->src_cell && b->src_link && b->src_amp), "src_cell/src_link/src_amp required when n_src > 0");
    FWI_CHECK_ARG(d->n_rec == 0 || (b->rec_cell && b->rec_out), "rec_cell/rec_out required when n_rec > 0");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    EParams p;
    fill(p, d, b);
    const dim3 block = block_shape();
    const dim3 grid((d->nyp - 2 * p.h + block.x - 1) / block.x, (d->nzp - 2 * p.h + block.y - 1) / block.y, d->n_shots);
    const long long fld = (long long)d->n_shots * p.cells;
    const long long nrec = (long long)d->n_shots * d->n_rec, nsrc = (long long)d->n_shots * d->n_src;
    // F1 + F2 as one kernel with double-buffered fields (elastic_f_fused): opt-in with FWI_B200_EL_FUSED_FWD=1.
    // It moves 80 B per cell-update instead of ~100 B, but its three barrier-separated stages measured
    // 57.6 us per step on C3 against 42.4 us for the two kernels, so the two-kernel form stays the default.
    const bool fused = getenv("FWI_B200_EL_FUSED_FWD") && atoi(getenv("FWI_B200_EL_FUSED_FWD")) == 1;
    FWI_CHECK_ARG(!(fused && d->free_surface), "the opt-in fused forward step (FWI_B200_EL_FUSED_FWD=1) has no free surface");
    const dim3 fgrid((d->nyp - 2 * p.h + kTX - 1) / kTX, (d->nzp - 2 * p.h + kTZ - 1) / kTZ, d->n_shots);
    float *setA[9], *setB[9];
    for (int k = 0; k < 5; ++k) { setA[k] = b->wave + k * fld; setB[k] = b->wave + (13 + k) * fld; }
    for (int k = 0; k < 4; ++k) { setA[5 + k] = b->wave + (PSI + k) * fld; setB[5 + k] = b->wave + (18 + k) * fld; }
    // point work of a step: sample the receivers before it, inject the sources after it.  The injection of step n and
    // the sampling of step n + 1 touch different fields and share one launch.
    FWI_CHECK_ARG(d->source_kind >= 0 && d->source_kind <= 2, "source_kind must be 0 (explosive), 1 (vertical force) or 2 (horizontal force)");
    auto fill_sample = [&](PointArgs &q, int n) {
        q.n_smp = d->n_rec; q.smp_cell = b->rec_cell; q.sa = p.f[VY]; q.sb = p.f[VZ]; q.s_coef = nullptr;
        q.out_a = b->rec_out + n * nrec; q.out_b = b->rec_out + ((long long)d->nsteps + n) * nrec;
        q.smp_sum = 0; q.s_scale = 1.0f; q.g_img = nullptr;
    };
    for (int n = 0; n < d->nsteps; ++n) {
        if (fused) {
            float **in = (n & 1) ? setB : setA, **out = (n & 1) ? setA : setB;
            for (int k = 0; k < 5; ++k) { p.f[k] = in[k]; p.fo[k] = out[k]; }
            for (int k = 0; k < 4; ++k) { p.psi_in[k] = in[5 + k]; p.psi_out[k] = out[5 + k]; }
        }
        if (n == 0 && d->n_rec) {
            PointArgs q = {};
            q.cells = p.cells; q.S = d->n_shots;
            fill_sample(q, 0);
            launch_points(q, st);
        }
        p.snap = b->snap ? b->snap + (long long)n * 5 * fld : nullptr;
        if (fused) {
            if (p.snap) launch_f_fused_h<true>(p, fgrid, st);
            else launch_f_fused_h<false>(p, fgrid, st);
            fwi::count_launch();
        } else {
            if (p.fs_row) {
                if (p.snap) { ELASTIC_LAUNCH_H(elastic_f1, true, true); ELASTIC_LAUNCH_H(elastic_f2, true, true); }
                else { ELASTIC_LAUNCH_H(elastic_f1, false, true); ELASTIC_LAUNCH_H(elastic_f2, false, true); }
            } else {
                if (p.snap) { ELASTIC_LAUNCH_H(elastic_f1, true, false); ELASTIC_LAUNCH_H(elastic_f2, true, false); }
                else { ELASTIC_LAUNCH_H(elastic_f1, false, false); ELASTIC_LAUNCH_H(elastic_f2, false, false); }
            }
            fwi::count_launch(2);
        }
        {
            float **cur = fused ? p.fo : p.f;       // the fields this step has just written
            PointArgs q = {};
            q.cells = p.cells; q.S = d->n_shots;
            if (d->n_src) {
                q.n_inj = d->n_src; q.inj_cell = b->src_cell; q.inj_link = b->src_link;
                q.amp_a = q.amp_b = b->src_amp + n * nsrc; q.i_scale = p.dt; q.i_coef = nullptr;
                if (d->source_kind == 0) { q.ia = cur[SXX]; q.ib = cur[SZZ]; }
                else if (d->source_kind == 1) { q.ia = cur[VZ]; q.ib = nullptr; q.i_coef = b->bz; }
                else { q.ia = cur[VY]; q.ib = nullptr; q.i_coef = b->by; }
            }
            if (d->n_rec && n + 1 < d->nsteps) {
                fill_sample(q, n + 1);
                q.sa = cur[VY]; q.sb = cur[VZ];
                // a force source writes the velocity the sampling of the next step reads: keep the two apart then
                if (d->n_src && d->source_kind != 0) {
                    PointArgs qi = q;
                    qi.n_smp = 0;
                    launch_points(qi, st);
                    q.n_inj = 0;
                }
            }
            launch_points(q, st);
        }
        FWI_CHECK_CUDA(cudaGetLastError());
    }
    return FWI_OK;
}

extern "C" int fwi_elastic_backward(const fwi_elastic_desc *d, const fwi_elastic_buffers *b, void *stream) {
    int rc = check(d);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(b && b->l2m && b->lam && b->mu && b->by && b->bz && b->prof_z && b->prof_y && b->wave,
                  "parameter grids, profiles and wave workspace must not be NULL");
    FWI_CHECK_ARG(b->snap && b->scratch && b->grad, "snap, scratch and grad are required by the adjoint pass");
    FWI_CHECK_ARG(d->n_rec == 0 || (b->rec_cell && b->rec_link && b->grad_rec), "rec_cell/rec_link/grad_rec required when n_rec > 0");
    FWI_CHECK_ARG(d->n_src == 0 || (b->src_cell && b->grad_src), "src_cell/grad_src required when n_src > 0");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    EParams p;
    fill(p, d, b);
    const dim3 block = block_shape();
    const dim3 grid((d->nyp - 2 * p.h + block.x - 1) / block.x, (d->nzp - 2 * p.h + block.y - 1) / block.y, d->n_shots);
    const long long fld = (long long)d->n_shots * p.cells;
    const long long nrec = (long long)d->n_shots * d->n_rec, nsrc = (long long)d->n_shots * d->n_src;
    // fused adjoint (default): B2a+B2b and B1a+B1b as one kernel each; FWI_B200_EL_FUSED=0 keeps the four-kernel form
    static const bool fused = !(getenv("FWI_B200_EL_FUSED") && atoi(getenv("FWI_B200_EL_FUSED")) == 0);
    static const int fnt = getenv("FWI_B200_EL_NT") ? atoi(getenv("FWI_B200_EL_NT")) : 512;
    const dim3 fblock(fnt == 256 || fnt == 1024 ? fnt : 512);
    FWI_CHECK_ARG(fused || !d->free_surface, "the four-kernel adjoint (FWI_B200_EL_FUSED=0) has no free surface");
    const int zs = p.fs_row ? p.fs_row : p.h;
    const dim3 fgrid((d->nyp - 2 * p.h + kTX - 1) / kTX, (d->nzp - p.h - zs + kTZ - 1) / kTZ, d->n_shots);
    FWI_CHECK_ARG(d->source_kind >= 0 && d->source_kind <= 2, "source_kind must be 0, 1 or 2");
    FWI_CHECK_ARG(d->source_kind == 0 || d->n_src == 0 || (b->src_amp && b->src_link),
                  "force sources: the adjoint pass needs src_amp and src_link (image term of the buoyancy)");
    // adjoint of a step, in reverse: sample the adjoint field at the sources (-> grad_src[n]), B2, B1, inject grad_rec[n].
    // The injection of step n and the sampling of step n - 1 share one launch (different fields, except for force
    // sources, which are sampled from the velocities the injection writes: two launches then).
    auto fill_sample = [&](PointArgs &q, int n) {
        q.n_smp = d->n_src; q.smp_cell = b->src_cell; q.out_a = b->grad_src + n * nsrc; q.out_b = nullptr;
        q.s_scale = p.dt; q.s_coef = nullptr; q.g_img = nullptr;
        if (d->source_kind == 0) { q.sa = p.f[SXX]; q.sb = p.f[SZZ]; q.smp_sum = 1; }
        else {
            const int comp = d->source_kind == 1 ? VZ : VY;
            q.sa = p.f[comp]; q.sb = nullptr; q.smp_sum = 0;
            q.s_coef = d->source_kind == 1 ? b->bz : b->by;
            q.g_img = b->grad + (d->source_kind == 1 ? 4 : 3) * fld;
            q.img_amp = b->src_amp + n * nsrc; q.smp_link = b->src_link;
        }
    };
    for (int n = d->nsteps - 1; n >= 0; --n) {
        if (n == d->nsteps - 1 && d->n_src) {
            PointArgs q = {};
            q.cells = p.cells; q.S = d->n_shots;
            fill_sample(q, n);
            launch_points(q, st);
        }
        p.snap = const_cast<float *>(b->snap) + (long long)n * 5 * fld;
        if (fused) {
            // memory variables ping-pong between the wave workspace and `scratch` (see elastic_b2_fused)
            const int par = ((d->nsteps - 1 - n) + d->adj_parity) & 1;
            for (int k = 0; k < 8; ++k) {
                float *a = p.f[PSI + k], *c = b->scratch + k * fld;
                p.psi_in[k] = par ? c : a;
                p.psi_out[k] = par ? a : c;
            }
            const dim3 block = fblock, grid = fgrid;
            if (p.fs_row) { ELASTIC_LAUNCH_H(elastic_b2_fused, 512, true); ELASTIC_LAUNCH_H(elastic_b1_fused, 512, true); }
            else if (fnt == 256) { ELASTIC_LAUNCH_H(elastic_b2_fused, 256, false); ELASTIC_LAUNCH_H(elastic_b1_fused, 256, false); }
            else if (fnt == 1024) { ELASTIC_LAUNCH_H(elastic_b2_fused, 1024, false); ELASTIC_LAUNCH_H(elastic_b1_fused, 1024, false); }
            else { ELASTIC_LAUNCH_H(elastic_b2_fused, 512, false); ELASTIC_LAUNCH_H(elastic_b1_fused, 512, false); }
            fwi::count_launch(2);
        } else {
            launch_pdl(elastic_b2a, grid, block, st, p);
            ELASTIC_LAUNCH_H(elastic_b2b);
            launch_pdl(elastic_b1a, grid, block, st, p);
            ELASTIC_LAUNCH_H(elastic_b1b);
            fwi::count_launch(4);
        }
        {
            PointArgs q = {};
            q.cells = p.cells; q.S = d->n_shots;
            if (d->n_rec) {
                q.n_inj = d->n_rec; q.inj_cell = b->rec_cell; q.inj_link = b->rec_link; q.ia = p.f[VY]; q.ib = p.f[VZ];
                q.amp_a = b->grad_rec + n * nrec; q.amp_b = b->grad_rec + ((long long)d->nsteps + n) * nrec;
                q.i_scale = 1.0f; q.i_coef = nullptr;
            }
            if (d->n_src && n > 0) {
                fill_sample(q, n - 1);
                if (d->n_rec && d->source_kind != 0) {
                    PointArgs qi = q;
                    qi.n_smp = 0; qi.g_img = nullptr;
                    launch_points(qi, st);
                    q.n_inj = 0;
                }
            }
            launch_points(q, st);
        }
        FWI_CHECK_CUDA(cudaGetLastError());
    }
    return FWI_OK;
}
