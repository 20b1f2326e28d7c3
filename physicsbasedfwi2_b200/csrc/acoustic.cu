// 2-D constant-density acoustic propagator for sm_100a: one fused kernel per time step.
//
// Replaces the time loop of deepwave.scalar.Propagator.forward / backward as called at
// reference models/networks.py:5464 (forward) and :5491 (adjoint).  Not a port: deepwave's
// kernels are one-thread-per-cell global-memory stencils; here a CTA owns a 128-column tile,
// stages u with its halo in shared memory (cp.async, double buffered over the shots of a
// group), each thread owns a float4 of columns x RPT rows, and the step fuses
//   receiver sampling  +  stencil/PML update  +  source injection  +  snapshot write
//   (forward)  or  snapshot read + imaging accumulation (adjoint)
// so a launch moves each field through HBM once.  The image and v^2dt^2 stay in registers
// across the shots a CTA walks, so their traffic is amortised over the group.
//
// Arithmetic order per cell follows oracle/cpu_ref_impl.h (explicit fma, no contraction).
#include "common.h"

namespace {

constexpr int kBY = 128;      // tile width in cells: 32 lanes x float4
constexpr int kLM = 32;       // device column of the first computational column (128 B aligned)
constexpr int kSP = kBY + 8;  // shared-memory row pitch in floats: 4-float halo on both sides

struct StepParams {
    int nzp, pitch;
    int zend, yend;          // end (exclusive) of the computational region, device coordinates
    int zlo, zhi, ylo, yhi;  // region with sigma = 0 and no PML memory variable within reach
    long long stride;        // floats per shot per field
    int S, spg;
    int tiles_y, n_tiles;
    int pml;
    const float *u;
    float *um;
    const float *pz, *py;
    float *pzn, *pyn;
    const float *w, *sigz, *sigy;
    float *qw;
    const float *qr;
    float *G;
    const int *inj_start, *inj_cell, *inj_col;
    const float *inj_amp;
    int n_inj, inj_scale_a;
    const int *smp_cell;
    float *smp_out;
    int n_smp, smp_div_a;
    float c0, cz2[5], cy2[5], cz1[5], cy1[5], dt, dt2, hdt;
};

__device__ __forceinline__ void f4_to_arr(const float4 v, float *a) {
    a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = v.w;
}
__device__ __forceinline__ float4 arr_to_f4(const float *a) { return make_float4(a[0], a[1], a[2], a[3]); }

// MODE 0: advance only.  MODE 1: advance and write q (forward with gradient).
// MODE 2: read q, accumulate the image, advance (adjoint with model gradient).
template <int H, int MODE, int NW, int RPT>
__global__ void __launch_bounds__(NW * 32, (NW <= 8 ? 2 : 1))
acoustic_step_kernel(const __grid_constant__ StepParams p) {
    constexpr int BZ = NW * RPT;
    constexpr int SROWS = BZ + 2 * H;
    constexpr int NT = NW * 32;
    constexpr int SP4 = kSP / 4;
    extern __shared__ float4 smem4[];
    float *smem = reinterpret_cast<float *>(smem4);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, g = blockIdx.y;
    const int tz = tile / p.tiles_y, ty = tile - tz * p.tiles_y;
    const int z0 = H + tz * BZ, x0 = kLM + ty * kBY;

    // ---- sampling duty: record the current field at the sampling cells (before the step) ----
    {
        const int total = p.S * p.n_smp;
        const int nblk = gridDim.x * gridDim.y;
        for (int item = (blockIdx.y * gridDim.x + blockIdx.x) * NT + tid; item < total;
             item += nblk * NT) {
            const int s = item / p.n_smp;
            const int cell = p.smp_cell[item];
            float val = p.u[(size_t)s * p.stride + cell];
            if (p.smp_div_a) {
                const int z = cell / p.pitch, y = cell - z * p.pitch;
                const float hs = p.hdt * (p.sigz[z] + p.sigy[y]);
                val = val * (1.0f + hs);
            }
            p.smp_out[item] = val;
        }
    }

    const int s0 = g * p.spg;
    const int s1 = min(p.S, s0 + p.spg);
    if (s0 >= s1) return;

    auto load_tile = [&](int s, int buf) {
        const float *src = p.u + (size_t)s * p.stride;
        float *dst = smem + buf * (SROWS * kSP);
        for (int idx = tid; idx < SROWS * SP4; idx += NT) {
            const int i = idx / SP4, j = idx - i * SP4;
            const int z = z0 - H + i, c = x0 - 4 + 4 * j;
            const bool ok = (z < p.nzp) && (c + 3 < p.pitch);
            const float *gp = ok ? src + (size_t)z * p.pitch + c : src;
            fwi::cp_async16(dst + i * kSP + 4 * j, gp, ok ? 16 : 0);
        }
        fwi::cp_async_commit();
    };

    const int col = x0 + 4 * lane;
    const int zr = z0 + warp * RPT;
    const bool any_col = col < p.yend;
    bool cm[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) cm[j] = (col + j) < p.yend;
    bool rowok[RPT];
#pragma unroll
    for (int r = 0; r < RPT; ++r) rowok[r] = any_col && (zr + r) < p.zend;

    const bool pml_tile =
        p.pml && !(z0 >= p.zlo && z0 + BZ <= p.zhi && x0 >= p.ylo && x0 + kBY <= p.yhi);

    // per-thread constants across the shots of the group
    float wv[RPT][4];
    float Gacc[RPT][4];
#pragma unroll
    for (int r = 0; r < RPT; ++r) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rowok[r]) t = __ldg(reinterpret_cast<const float4 *>(p.w + (size_t)(zr + r) * p.pitch + col));
        f4_to_arr(t, wv[r]);
#pragma unroll
        for (int j = 0; j < 4; ++j) Gacc[r][j] = 0.f;
    }
    float sy[4] = {0.f, 0.f, 0.f, 0.f};
    if (pml_tile && any_col) f4_to_arr(__ldg(reinterpret_cast<const float4 *>(p.sigy + col)), sy);

    load_tile(s0, 0);
    for (int s = s0; s < s1; ++s) {
        const int buf = (s - s0) & 1;
        if (s + 1 < s1) load_tile(s + 1, buf ^ 1);
        const size_t sb = (size_t)s * p.stride;

        // global operands of this shot that need no staging: issue before waiting on the tile
        float umv[RPT][4], qrv[RPT][4];
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
            if (rowok[r]) {
                const size_t off = sb + (size_t)(zr + r) * p.pitch + col;
                a = *reinterpret_cast<const float4 *>(p.um + off);
                if (MODE == 2) b = __ldg(reinterpret_cast<const float4 *>(p.qr + off));
            }
            f4_to_arr(a, umv[r]);
            f4_to_arr(b, qrv[r]);
        }

        if (s + 1 < s1) fwi::cp_async_wait<1>(); else fwi::cp_async_wait<0>();
        __syncthreads();
        const float *t = smem + buf * (SROWS * kSP);

        // z window of this thread's float4 column: smem rows warp*RPT .. warp*RPT + RPT + 2H - 1
        float W[RPT + 2 * H][4];
#pragma unroll
        for (int i = 0; i < RPT + 2 * H; ++i)
            f4_to_arr(*reinterpret_cast<const float4 *>(t + (warp * RPT + i) * kSP + 4 + 4 * lane), W[i]);

        float un[RPT][4], qv[RPT][4], cw[RPT][4];
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            const float *rowp = t + (warp * RPT + r + H) * kSP + 4 * lane;
            float v[12];
            f4_to_arr(*reinterpret_cast<const float4 *>(rowp), v);
            f4_to_arr(*reinterpret_cast<const float4 *>(rowp + 8), v + 8);
#pragma unroll
            for (int j = 0; j < 4; ++j) v[4 + j] = W[r + H][j];

            float lap[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float l = p.c0 * v[4 + j];
#pragma unroll
                for (int k = 1; k <= H; ++k) l = fmaf(p.cy2[k], v[4 + j - k] + v[4 + j + k], l);
#pragma unroll
                for (int k = 1; k <= H; ++k) l = fmaf(p.cz2[k], W[r + H - k][j] + W[r + H + k][j], l);
                lap[j] = l;
            }

            if (!pml_tile) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float uc = v[4 + j];
                    if (MODE == 2) Gacc[r][j] = fmaf(uc, qrv[r][j], Gacc[r][j]);
                    qv[r][j] = lap[j];
                    un[r][j] = fmaf(wv[r][j], lap[j], fmaf(2.0f, uc, -umv[r][j]));
                    cw[r][j] = wv[r][j];
                }
            } else {
                const int z = zr + r;
                const float sz = (z < p.nzp) ? __ldg(p.sigz + z) : 0.f;
                const size_t off = sb + (size_t)z * p.pitch + col;
                float dpz[4] = {0.f, 0.f, 0.f, 0.f}, duz[4] = {0.f, 0.f, 0.f, 0.f};
                float dpy[4] = {0.f, 0.f, 0.f, 0.f}, duy[4] = {0.f, 0.f, 0.f, 0.f};
                float pzc[4] = {0.f, 0.f, 0.f, 0.f}, pyv[12];
#pragma unroll
                for (int j = 0; j < 12; ++j) pyv[j] = 0.f;
                if (rowok[r]) {
                    f4_to_arr(__ldg(reinterpret_cast<const float4 *>(p.pz + off)), pzc);
#pragma unroll
                    for (int k = 1; k <= H; ++k) {
                        float a[4], b[4];
                        f4_to_arr(__ldg(reinterpret_cast<const float4 *>(p.pz + off + (size_t)k * p.pitch)), a);
                        f4_to_arr(__ldg(reinterpret_cast<const float4 *>(p.pz + off - (size_t)k * p.pitch)), b);
#pragma unroll
                        for (int j = 0; j < 4; ++j) dpz[j] = fmaf(p.cz1[k], a[j] - b[j], dpz[j]);
                    }
                    f4_to_arr(__ldg(reinterpret_cast<const float4 *>(p.py + off - 4)), pyv);
                    f4_to_arr(__ldg(reinterpret_cast<const float4 *>(p.py + off)), pyv + 4);
                    f4_to_arr(__ldg(reinterpret_cast<const float4 *>(p.py + off + 4)), pyv + 8);
                }
                float pzn[4], pyn[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float uc = v[4 + j];
#pragma unroll
                    for (int k = 1; k <= H; ++k) {
                        duz[j] = fmaf(p.cz1[k], W[r + H + k][j] - W[r + H - k][j], duz[j]);
                        dpy[j] = fmaf(p.cy1[k], pyv[4 + j + k] - pyv[4 + j - k], dpy[j]);
                        duy[j] = fmaf(p.cy1[k], v[4 + j + k] - v[4 + j - k], duy[j]);
                    }
                    const float q = (lap[j] + dpz[j]) + dpy[j];
                    const float hs = p.hdt * (sz + sy[j]);
                    const float a = 1.0f / (1.0f + hs);
                    const float cu = a * fmaf(-p.dt2, sz * sy[j], 2.0f);
                    const float cum = a * (hs - 1.0f);
                    const float aw = a * wv[r][j];
                    if (MODE == 2) Gacc[r][j] = fmaf(uc, qrv[r][j], Gacc[r][j]);
                    qv[r][j] = q;
                    un[r][j] = fmaf(aw, q, fmaf(cu, uc, cum * umv[r][j]));
                    cw[r][j] = p.inj_scale_a ? aw : wv[r][j];
                    pzn[j] = cm[j] ? fmaf(p.dt * (sy[j] - sz), duz[j], fmaf(-p.dt, sz, 1.0f) * pzc[j]) : 0.f;
                    pyn[j] = cm[j] ? fmaf(p.dt * (sz - sy[j]), duy[j], fmaf(-p.dt, sy[j], 1.0f) * pyv[4 + j]) : 0.f;
                }
                if (rowok[r]) {
                    *reinterpret_cast<float4 *>(p.pzn + off) = arr_to_f4(pzn);
                    *reinterpret_cast<float4 *>(p.pyn + off) = arr_to_f4(pyn);
                }
            }
        }

        // ---- injection: entries of (shot, tile), in their original order ----
        {
            const int *st = p.inj_start + (size_t)s * (p.n_tiles + 1) + tile;
            const int e0 = __ldg(st), e1 = __ldg(st + 1);
            for (int e = e0; e < e1; ++e) {
                const int cell = __ldg(p.inj_cell + e);
                const int z = cell / p.pitch, y = cell - z * p.pitch;
                const int rr = z - zr, jj = y - col;
                if (rr >= 0 && rr < RPT && jj >= 0 && jj < 4) {
                    const float amp = __ldg(p.inj_amp + (size_t)s * p.n_inj + __ldg(p.inj_col + e));
#pragma unroll
                    for (int r = 0; r < RPT; ++r)
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (r == rr && j == jj) un[r][j] = fmaf(cw[r][j], amp, un[r][j]);
                }
            }
        }

#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            if (rowok[r]) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (!cm[j]) { un[r][j] = 0.f; qv[r][j] = 0.f; }
                }
                const size_t off = sb + (size_t)(zr + r) * p.pitch + col;
                *reinterpret_cast<float4 *>(p.um + off) = arr_to_f4(un[r]);
                if (MODE == 1) __stcs(reinterpret_cast<float4 *>(p.qw + off), arr_to_f4(qv[r]));
            }
        }
        __syncthreads();  // every warp is done with `buf` before the next prefetch refills it
    }

    if (MODE == 2) {
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            if (rowok[r]) {
                float4 *gp = reinterpret_cast<float4 *>(p.G + (size_t)g * p.stride + (size_t)(zr + r) * p.pitch + col);
                float4 o = *gp;
                o.x += Gacc[r][0]; o.y += Gacc[r][1]; o.z += Gacc[r][2]; o.w += Gacc[r][3];
                *gp = o;
            }
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// Host side: layout, argument checks, the time loop.
// ------------------------------------------------------------------------------------------
namespace {

struct TileCfg { int nw, rpt; };

// Tile variants compiled in; FWI_B200_TILE=<index> picks one (tuning aid), default 0.
constexpr TileCfg kTileCfgs[] = {{8, 2}, {16, 2}, {8, 4}, {4, 2}};
constexpr int kNumTileCfgs = sizeof(kTileCfgs) / sizeof(kTileCfgs[0]);

int tile_cfg_index() {
    static int idx = [] {
        const char *e = getenv("FWI_B200_TILE");
        int i = e ? atoi(e) : 0;
        return (i < 0 || i >= kNumTileCfgs) ? 0 : i;
    }();
    return idx;
}

inline int round_up(int a, int b) { return (a + b - 1) / b * b; }

int check_desc(const fwi_acoustic_desc *d) {
    FWI_CHECK_ARG(d != nullptr, "descriptor is NULL");
    FWI_CHECK_ARG(d->order == 2 || d->order == 4 || d->order == 6 || d->order == 8,
                  "order must be 2, 4, 6 or 8 (got %d)", d->order);
    const int h = d->order / 2;
    FWI_CHECK_ARG(d->pml_width >= 0, "pml_width must be >= 0");
    FWI_CHECK_ARG(d->nzp > 2 * (h + d->pml_width) && d->nyp > 2 * (h + d->pml_width),
                  "padded grid %dx%d too small for order %d, pml_width %d", d->nzp, d->nyp,
                  d->order, d->pml_width);
    FWI_CHECK_ARG(d->n_shots >= 1, "n_shots must be >= 1");
    FWI_CHECK_ARG(d->nsteps >= 0, "nsteps must be >= 0");
    FWI_CHECK_ARG(d->n_inj >= 0 && d->n_smp >= 0, "negative point counts");
    FWI_CHECK_ARG(d->dz > 0 && d->dy > 0 && d->dt != 0, "dz, dy must be > 0 and dt != 0");
    FWI_CHECK_ARG(d->shots_per_group >= 0, "shots_per_group must be >= 0");
    return FWI_OK;
}

int make_layout(const fwi_acoustic_desc *d, fwi_acoustic_layout *L) {
    const int h = d->order / 2;
    const TileCfg tc = kTileCfgs[tile_cfg_index()];
    const int yend = kLM + d->nyp - 2 * h;
    L->col0 = kLM - h;
    L->pitch = round_up(round_up(yend, 4) + 4, 32);
    L->shot_stride = (int64_t)d->nzp * L->pitch;
    L->tile_z = tc.nw * tc.rpt;
    L->tile_y = kBY;
    L->tiles_z = (d->nzp - 2 * h + L->tile_z - 1) / L->tile_z;
    L->tiles_y = (d->nyp - 2 * h + kBY - 1) / kBY;
    int spg = d->shots_per_group;
    if (spg <= 0) {
        // auto: keep at least ~4 CTAs per SM in flight, otherwise walk as many shots as possible
        const int tiles = L->tiles_z * L->tiles_y;
        const int want = 148 * 4;
        int groups = (want + tiles - 1) / tiles;
        if (groups > d->n_shots) groups = d->n_shots;
        if (groups < 1) groups = 1;
        spg = (d->n_shots + groups - 1) / groups;
    }
    if (spg > d->n_shots) spg = d->n_shots;
    L->n_groups = (d->n_shots + spg - 1) / spg;
    L->wave_fields = d->pml_width > 0 ? 6 : 2;
    return spg;
}

template <int H, int MODE, int NW, int RPT>
cudaError_t launch_step(const StepParams &p, dim3 grid, cudaStream_t st) {
    constexpr int smem = 2 * (NW * RPT + 2 * H) * kSP * (int)sizeof(float);
    static bool configured = false;  // per instantiation
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(acoustic_step_kernel<H, MODE, NW, RPT>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    acoustic_step_kernel<H, MODE, NW, RPT><<<grid, NW * 32, smem, st>>>(p);
    return cudaGetLastError();
}

template <int H, int MODE>
cudaError_t launch_cfg(int cfg, const StepParams &p, dim3 grid, cudaStream_t st) {
    switch (cfg) {
    case 1: return launch_step<H, MODE, 16, 2>(p, grid, st);
    case 2: return launch_step<H, MODE, 8, 4>(p, grid, st);
    case 3: return launch_step<H, MODE, 4, 2>(p, grid, st);
    default: return launch_step<H, MODE, 8, 2>(p, grid, st);
    }
}

template <int MODE>
cudaError_t launch_h(int h, int cfg, const StepParams &p, dim3 grid, cudaStream_t st) {
    switch (h) {
    case 1: return launch_cfg<1, MODE>(cfg, p, grid, st);
    case 2: return launch_cfg<2, MODE>(cfg, p, grid, st);
    case 3: return launch_cfg<3, MODE>(cfg, p, grid, st);
    default: return launch_cfg<4, MODE>(cfg, p, grid, st);
    }
}

}  // namespace

extern "C" int fwi_acoustic_get_layout(const fwi_acoustic_desc *desc, fwi_acoustic_layout *out) {
    int rc = check_desc(desc);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(out != nullptr, "layout output is NULL");
    make_layout(desc, out);
    return FWI_OK;
}

extern "C" int fwi_acoustic_run(const fwi_acoustic_desc *d, const fwi_acoustic_buffers *b, void *stream) {
    int rc = check_desc(d);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(b != nullptr, "buffers struct is NULL");
    FWI_CHECK_ARG(b->w && b->wave, "w and wave must not be NULL");
    FWI_CHECK_ARG(d->pml_width == 0 || (b->sigz && b->sigy), "sigz/sigy required when pml_width > 0");
    FWI_CHECK_ARG(!d->smp_div_a || (b->sigz && b->sigy), "sigz/sigy required when smp_div_a is set");
    FWI_CHECK_ARG(d->n_smp == 0 || (b->smp_cell && b->smp_out), "smp_cell/smp_out required when n_smp > 0");
    FWI_CHECK_ARG(b->inj_tile_start != nullptr, "inj_tile_start required (all-zero offsets when n_inj == 0)");
    FWI_CHECK_ARG(d->n_inj == 0 || (b->inj_cell && b->inj_col && b->inj_amp), "injection lists required when n_inj > 0");
    FWI_CHECK_ARG(!(b->q_write && b->q_read), "q_write and q_read are mutually exclusive");
    FWI_CHECK_ARG(!b->q_read || b->image, "image required with q_read");
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    fwi_acoustic_layout L;
    const int spg = make_layout(d, &L);
    const int h = d->order / 2, pw = d->pml_width;
    const int cfg = tile_cfg_index();

    StepParams p;
    p.nzp = d->nzp; p.pitch = L.pitch;
    p.zend = d->nzp - h; p.yend = kLM + d->nyp - 2 * h;
    p.zlo = 2 * h + pw; p.zhi = d->nzp - 2 * h - pw;
    p.ylo = L.col0 + 2 * h + pw; p.yhi = L.col0 + d->nyp - 2 * h - pw;
    p.stride = L.shot_stride;
    p.S = d->n_shots; p.spg = spg;
    p.tiles_y = L.tiles_y; p.n_tiles = L.tiles_y * L.tiles_z;
    p.pml = pw > 0;
    p.w = b->w; p.sigz = b->sigz; p.sigy = b->sigy;
    p.inj_start = b->inj_tile_start; p.inj_cell = b->inj_cell; p.inj_col = b->inj_col;
    p.n_inj = d->n_inj; p.inj_scale_a = d->inj_scale_a;
    p.smp_cell = b->smp_cell; p.n_smp = d->n_smp; p.smp_div_a = d->smp_div_a;
    const double dz = d->dz, dy = d->dy;
    p.c0 = (float)(fwi::kD2[h][0] / (dz * dz) + fwi::kD2[h][0] / (dy * dy));
    for (int k = 0; k < 5; ++k) p.cz2[k] = p.cy2[k] = p.cz1[k] = p.cy1[k] = 0.f;
    for (int k = 1; k <= h; ++k) {
        p.cz2[k] = (float)(fwi::kD2[h][k] / (dz * dz));
        p.cy2[k] = (float)(fwi::kD2[h][k] / (dy * dy));
        p.cz1[k] = (float)(fwi::kD1[h][k - 1] / dz);
        p.cy1[k] = (float)(fwi::kD1[h][k - 1] / dy);
    }
    p.dt = (float)d->dt; p.dt2 = p.dt * p.dt; p.hdt = 0.5f * p.dt;

    const size_t fld = (size_t)d->n_shots * L.shot_stride;
    float *u = b->wave, *um = b->wave + fld;
    float *pz = nullptr, *pzn = nullptr, *py = nullptr, *pyn = nullptr;
    if (pw > 0) { pz = b->wave + 2 * fld; pzn = b->wave + 3 * fld; py = b->wave + 4 * fld; pyn = b->wave + 5 * fld; }

    const int mode = b->q_read ? 2 : (b->q_write ? 1 : 0);
    const dim3 grid(p.n_tiles, L.n_groups);
    for (int it = 0; it < d->nsteps; ++it) {
        const int n = d->adjoint ? d->nsteps - 1 - it : it;
        p.u = u; p.um = um; p.pz = pz; p.pzn = pzn; p.py = py; p.pyn = pyn;
        p.qw = b->q_write ? b->q_write + (size_t)n * fld : nullptr;
        p.qr = b->q_read ? b->q_read + (size_t)n * fld : nullptr;
        p.G = b->image;
        p.inj_amp = d->n_inj ? b->inj_amp + (size_t)n * d->n_shots * d->n_inj : nullptr;
        p.smp_out = d->n_smp ? b->smp_out + (size_t)n * d->n_shots * d->n_smp : nullptr;
        cudaError_t e;
        if (mode == 2) e = launch_h<2>(h, cfg, p, grid, st);
        else if (mode == 1) e = launch_h<1>(h, cfg, p, grid, st);
        else e = launch_h<0>(h, cfg, p, grid, st);
        if (e != cudaSuccess) {
            fwi::set_error("acoustic step launch failed at step %d: %s", n, cudaGetErrorString(e));
            return FWI_ERR_CUDA;
        }
        fwi::count_launch();
        float *t = u; u = um; um = t;
        t = pz; pz = pzn; pzn = t;
        t = py; py = pyn; pyn = t;
    }
    return FWI_OK;
}

extern "C" int fwi_acoustic_forward(const fwi_acoustic_desc *d, const fwi_acoustic_buffers *b, void *stream) {
    FWI_CHECK_ARG(d && b, "NULL argument");
    FWI_CHECK_ARG(d->adjoint == 0, "fwi_acoustic_forward needs desc->adjoint == 0");
    FWI_CHECK_ARG(b->q_read == nullptr, "fwi_acoustic_forward does not read snapshots");
    return fwi_acoustic_run(d, b, stream);
}

extern "C" int fwi_acoustic_backward(const fwi_acoustic_desc *d, const fwi_acoustic_buffers *b, void *stream) {
    FWI_CHECK_ARG(d && b, "NULL argument");
    FWI_CHECK_ARG(d->adjoint == 1, "fwi_acoustic_backward needs desc->adjoint == 1");
    FWI_CHECK_ARG(b->q_write == nullptr, "fwi_acoustic_backward does not write snapshots");
    return fwi_acoustic_run(d, b, stream);
}
