// 2-D constant-density acoustic propagator for sm_100a: one fused kernel launch per time step.
//
// Replaces the time loop of deepwave.scalar.Propagator.forward / backward as called at
// reference models/networks.py:5464 (forward) and :5491 (adjoint).  Not a port: deepwave's
// kernels are one-thread-per-cell global-memory stencils.  Here the grid of one shot is cut
// into WORK UNITS, each owned by ONE WARP that marches down z through a strip of columns:
//
//   * the warp keeps the 2H+RPT rows of u it needs for the z-stencil in REGISTERS (a sliding
//     window), so every u value is loaded from memory once per strip (no z-halo re-reads inside
//     a unit) with 16-byte coalesced loads;
//   * the y-neighbours of the centre row are exchanged through a per-warp shared-memory row
//     (no block-level barrier anywhere: warps never wait for one another);
//   * the operands of the next RPT rows (u, u-, v^2dt^2, snapshot) are prefetched into registers
//     while the current rows are computed, which keeps enough bytes in flight per SM to cover
//     HBM latency;
//   * the step fuses receiver sampling + stencil/PML update + source injection + snapshot write
//     (forward) or snapshot read + imaging accumulation (adjoint): each field crosses HBM once
//     per step.
//
// Three unit types share the kernel: interior strips (128 columns, float4 per lane, no PML
// arithmetic at all), top/bottom PML bands (128 columns, full PML update) and left/right PML
// bands (32 columns, one column per lane).  PML units are listed first so the heavier ones start
// early.  The layout puts the first interior column on a 128-byte boundary.
//
// Arithmetic order per cell follows oracle/cpu_ref_impl.h (explicit fma, no contraction).
#include <type_traits>

#include "common.h"

namespace {

constexpr int kNW = 4;        // warps per CTA (warps are independent; this is scheduling granularity)

struct StepParams {
    int nzp, pitch;
    int zs, ze, ys, ye;      // computational region (device coordinates, end exclusive)
    int zlo, zhi, ylo, yhi;  // interior: sigma = 0 and no PML memory variable within reach
    int xb;                  // first column of the top/bottom band strips (multiple of 4, <= ys)
    int xr;                  // first column of the right band strips (multiple of 4, <= yhi)
    long long stride;        // floats per shot per field
    int S;
    int pml;
    // unit bookkeeping (per shot)
    int tb_strips, tb_chunks, side_strips_l, side_strips_r, side_chunks, int_strips, int_chunks;
    int bz, bzs, bzt;        // rows per interior chunk / per side-band chunk / per top-bottom chunk
    int n_tb, n_side, n_int; // units per shot of each type
    const float *u;
    float *um;
    const float *pz, *py;
    float *pzn, *pyn;
    const float *w, *sigz, *sigy;
    float *qw;
    const float *qr;
    float *G;
    const int *inj_row_start, *inj_cell, *inj_col;
    const float *inj_amp;
    int n_inj, inj_scale_a;
    const int *smp_cell;
    float *smp_out;
    int n_smp, smp_div_a;
    float c0, cz2[5], cy2[5], cz1[5], cy1[5], dt, dt2, hdt;
};

// ---- float4 helpers ---------------------------------------------------------------------------
struct F4 { float a[4]; };
__device__ __forceinline__ F4 f4(const float4 t) { F4 r; r.a[0] = t.x; r.a[1] = t.y; r.a[2] = t.z; r.a[3] = t.w; return r; }
__device__ __forceinline__ float4 t4(const F4 &r) { return make_float4(r.a[0], r.a[1], r.a[2], r.a[3]); }
__device__ __forceinline__ F4 ld_nc(const float *p) { return f4(__ldg(reinterpret_cast<const float4 *>(p))); }
__device__ __forceinline__ F4 ld_plain(const float *p) { return f4(*reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ F4 ld_stream(const float *p) { return f4(__ldcs(reinterpret_cast<const float4 *>(p))); }
// masked store of 4 values starting at device column `col` of row pointer `row`
template <bool STREAM>
__device__ __forceinline__ void st_masked(float *row, int col, const F4 &v, int cmin, int cmax, bool rowok) {
    if (!rowok) return;
    if (col >= cmin && col + 3 < cmax) {
        if (STREAM) __stcs(reinterpret_cast<float4 *>(row + col), t4(v));
        else *reinterpret_cast<float4 *>(row + col) = t4(v);
        return;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
        if (col + j >= cmin && col + j < cmax) {
            if (STREAM) __stcs(row + col + j, v.a[j]);
            else row[col + j] = v.a[j];
        }
}

// ring of u rows when the next block is prefetched: >= 2H + 2*RPT rows, an even number of blocks
__host__ __device__ constexpr int ring_rows(int H, int RPT) { return (2 * H + 2 * RPT + 2 * RPT - 1) / (2 * RPT) * (2 * RPT); }

// One warp marches down z through a strip of columns of shot s.  The 32 lanes form 32/LW lane
// groups of LW lanes; group g owns columns [x0, x0 + 4*LW) and rows [zb + g*gstep, ... + nrows),
// clipped to zend; only columns in [cmin, cmax) are stored.  Every lane owns one float4 per row.
//   LW = 32: one strip of 128 columns (interior strips, top/bottom PML bands)
//   LW =  8: four row-chunks of a 32-column strip side by side (left/right PML bands)
// MODE 0: advance only.  MODE 1: advance and write q (forward with gradient).
// MODE 2: read q, accumulate the image, advance (adjoint with model gradient).
// ROT: the z loop is unrolled over one period of the register ring so that the sliding window of
// u rows never moves between registers (hot interior path); otherwise the window is shifted.
template <int H, int MODE, bool PML, int RPT, int LW, bool ROT>
__device__ __forceinline__ void march(const StepParams &p, const int s, const int x0, const int zb0,
                                      const int gstep, const int nrows, const int zend,
                                      const int cmin, const int cmax, float *sm) {
    constexpr int SW = 4 * LW + 8;              // shared row of one lane group: halo | own | halo
    constexpr int WR = ROT ? ring_rows(H, RPT) : 2 * H + RPT;
    constexpr int PER = ROT ? WR / RPT : 1;
    const int lane = threadIdx.x & 31;
    const int lx = lane % LW, grp = lane / LW;
    const int zb = zb0 + grp * gstep;
    const int ze = min(zend, zb + nrows);       // this lane group's end row (stores only)
    const int col = x0 + 4 * lx;                // true column (store mask)
    const int colL = min(col, p.pitch - 4);     // column used for loads (always in range)
    const int zmax = p.ze - 1;                  // last computational row: loads are clamped to it
    const size_t sb = (size_t)s * p.stride;
    const float *__restrict__ u = p.u + sb;
    float *__restrict__ um = p.um + sb;
    // halo float4 this lane fetches for the y-stencil (first lane: left of the strip, last: right)
    const bool hl = (lx == 0) || (lx == LW - 1);
    const int hcol = (lx == 0) ? x0 - 4 : min(x0 + 4 * LW, p.pitch - 4);
    float *srow = sm + grp * (2 * SW);

    F4 W[WR];
#pragma unroll
    for (int i = 0; i < 2 * H; ++i) W[i] = ld_nc(u + min(zb - H + i, zmax + H) * p.pitch + colL);

    F4 sy;
#pragma unroll
    for (int j = 0; j < 4; ++j) sy.a[j] = PML ? __ldg(p.sigy + colL + j) : 0.f;

    // operands of the centre rows of a block (two sets: the block being computed / being fetched)
    F4 oM[2][RPT], oW[2][RPT], oQ[2][RPT];
    float4 oHal[2][RPT];
    // issues the loads of the block starting at row z: its RPT new window rows go to ring slots
    // slot0 + r, its centre-row operands to set `par`
    auto fetch = [&](const int z, auto slot_tag, auto par_tag) {
        constexpr int slot0 = decltype(slot_tag)::value, par = decltype(par_tag)::value;
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            const int off = min(z + r, zmax) * p.pitch;
            W[(slot0 + r) % WR] = ld_nc(u + off + H * p.pitch + colL);
            oM[par][r] = ld_plain(um + off + colL);
            oW[par][r] = ld_nc(p.w + off + colL);
            if (MODE == 2) oQ[par][r] = ld_stream(p.qr + sb + off + colL);
            if (hl) oHal[par][r] = __ldg(reinterpret_cast<const float4 *>(u + off + hcol));
        }
    };
    const int zstop = zb + nrows;
    if constexpr (ROT) fetch(zb, std::integral_constant<int, 2 * H>{}, std::integral_constant<int, 0>{});

    // one block of RPT rows starting at row z; the ring slot of window row i is (base + i) % WR
    auto block = [&](const int z, auto base_tag, auto par_tag) {
        constexpr int base = decltype(base_tag)::value, par = decltype(par_tag)::value;
        if constexpr (ROT) {
            // prefetch the next block while this one is computed
            if (z + RPT < zstop)
                fetch(z + RPT, std::integral_constant<int, (base + 2 * H + RPT) % WR>{},
                      std::integral_constant<int, par ^ 1>{});
        } else {
            fetch(z, std::integral_constant<int, 2 * H>{}, std::integral_constant<int, 0>{});
        }
        F4 (&cM)[RPT] = oM[par];
        F4 (&cW)[RPT] = oW[par];
        F4 (&cQ)[RPT] = oQ[par];
        float4 (&cHal)[RPT] = oHal[par];
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            const int row = z + r;
            const bool rowok = row < ze;
            const F4 &C = W[(base + H + r) % WR];
            // ---- y-neighbours of the centre row through the lane group's shared row ----
            float *sbuf = srow + (r & 1) * SW;
            *reinterpret_cast<float4 *>(sbuf + 4 + 4 * lx) = t4(C);
            if (hl) *reinterpret_cast<float4 *>(sbuf + (lx == 0 ? 0 : 4 + 4 * LW)) = cHal[r];
            __syncwarp();
            float v[12];
            {
                const float4 L = *reinterpret_cast<const float4 *>(sbuf + 4 * lx);
                const float4 R = *reinterpret_cast<const float4 *>(sbuf + 4 * lx + 8);
                v[0] = L.x; v[1] = L.y; v[2] = L.z; v[3] = L.w;
                v[8] = R.x; v[9] = R.y; v[10] = R.z; v[11] = R.w;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) v[4 + j] = C.a[j];

            F4 lap;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float l = p.c0 * v[4 + j];
#pragma unroll
                for (int k = 1; k <= H; ++k) l = fmaf(p.cy2[k], v[4 + j - k] + v[4 + j + k], l);
#pragma unroll
                for (int k = 1; k <= H; ++k)
                    l = fmaf(p.cz2[k], W[(base + H + r - k) % WR].a[j] + W[(base + H + r + k) % WR].a[j], l);
                lap.a[j] = l;
            }

            F4 un, qv;
            const int off = min(row, zmax) * p.pitch;
            if (!PML) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    qv.a[j] = lap.a[j];
                    un.a[j] = fmaf(cW[r].a[j], lap.a[j], fmaf(2.0f, v[4 + j], -cM[r].a[j]));
                }
            } else {
                const float sz = __ldg(p.sigz + min(row, zmax));
                const float *__restrict__ pz = p.pz + sb + off;
                const float *__restrict__ py = p.py + sb + off;
                const F4 pzc = ld_nc(pz + colL);
                F4 dpz;
#pragma unroll
                for (int j = 0; j < 4; ++j) dpz.a[j] = 0.f;
#pragma unroll
                for (int k = 1; k <= H; ++k) {
                    const F4 a = ld_nc(pz + k * p.pitch + colL), b = ld_nc(pz - k * p.pitch + colL);
#pragma unroll
                    for (int j = 0; j < 4; ++j) dpz.a[j] = fmaf(p.cz1[k], a.a[j] - b.a[j], dpz.a[j]);
                }
                float pyv[12];
                {
                    const F4 t0 = ld_nc(py + colL - 4), t1 = ld_nc(py + colL);
                    const F4 t2 = ld_nc(py + min(colL + 4, p.pitch - 4));
#pragma unroll
                    for (int j = 0; j < 4; ++j) { pyv[j] = t0.a[j]; pyv[4 + j] = t1.a[j]; pyv[8 + j] = t2.a[j]; }
                }
                F4 pzn, pyn;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float uc = v[4 + j];
                    float duz = 0.f, dpy = 0.f, duy = 0.f;
#pragma unroll
                    for (int k = 1; k <= H; ++k) {
                        duz = fmaf(p.cz1[k], W[(base + H + r + k) % WR].a[j] - W[(base + H + r - k) % WR].a[j], duz);
                        dpy = fmaf(p.cy1[k], pyv[4 + j + k] - pyv[4 + j - k], dpy);
                        duy = fmaf(p.cy1[k], v[4 + j + k] - v[4 + j - k], duy);
                    }
                    const float q = (lap.a[j] + dpz.a[j]) + dpy;
                    const float hs = p.hdt * (sz + sy.a[j]);
                    const float a = 1.0f / (1.0f + hs);
                    const float cu = a * fmaf(-p.dt2, sz * sy.a[j], 2.0f);
                    const float cum = a * (hs - 1.0f);
                    const float aw = a * cW[r].a[j];
                    qv.a[j] = q;
                    un.a[j] = fmaf(aw, q, fmaf(cu, uc, cum * cM[r].a[j]));
                    pzn.a[j] = fmaf(p.dt * (sy.a[j] - sz), duz, fmaf(-p.dt, sz, 1.0f) * pzc.a[j]);
                    pyn.a[j] = fmaf(p.dt * (sz - sy.a[j]), duy, fmaf(-p.dt, sy.a[j], 1.0f) * pyv[4 + j]);
                }
                st_masked<false>(p.pzn + sb + off, col, pzn, cmin, cmax, rowok);
                st_masked<false>(p.pyn + sb + off, col, pyn, cmin, cmax, rowok);
            }
            st_masked<false>(um + off, col, un, cmin, cmax, rowok);
            if (MODE == 1) st_masked<true>(p.qw + sb + off, col, qv, cmin, cmax, rowok);
            if (MODE == 2) {
                // image of this shot: G += field * q_read (single owner per cell: deterministic)
                F4 g = ld_plain(p.G + sb + off + colL);
#pragma unroll
                for (int j = 0; j < 4; ++j) g.a[j] = fmaf(v[4 + j], cQ[r].a[j], g.a[j]);
                st_masked<false>(p.G + sb + off, col, g, cmin, cmax, rowok);
            }
        }
    };

    if constexpr (ROT) {
        int z = zb;
        while (z < zstop) {
            // one period of the ring: PER blocks, window slots are compile-time constants
            auto run = [&](auto self, auto ph_tag) -> void {
                constexpr int ph = decltype(ph_tag)::value;
                if constexpr (ph < PER) {
                    if (z < zstop) {
                        block(z, std::integral_constant<int, (ph * RPT) % WR>{}, std::integral_constant<int, ph & 1>{});
                        z += RPT;
                        self(self, std::integral_constant<int, ph + 1>{});
                    }
                }
            };
            run(run, std::integral_constant<int, 0>{});
        }
    } else {
        for (int z = zb; z < zstop; z += RPT) {
            block(z, std::integral_constant<int, 0>{}, std::integral_constant<int, 0>{});
#pragma unroll
            for (int i = 0; i < 2 * H; ++i) W[i] = W[i + RPT];
        }
    }

    // ---- injection into the freshly written field: points of this shot inside this lane group ----
    if (p.n_inj > 0) {
        __syncwarp();  // orders this warp's stores above before the read-modify-writes below
        const int *rs = p.inj_row_start + (size_t)s * (p.nzp + 1);
        const int base = s * p.n_inj;
        const int zq = min(zb, ze);
        const int e0 = __ldg(rs + zq), e1 = __ldg(rs + ze);
        const int c_lo = max(cmin, x0), c_hi = min(cmax, x0 + 4 * LW);
        for (int e = e0 + lx; e < e1; e += LW) {
            const int cell = __ldg(p.inj_cell + e);
            const int zc = cell / p.pitch, yc = cell - zc * p.pitch;
            if (yc < c_lo || yc >= c_hi) continue;
            if (e > base && __ldg(p.inj_cell + e - 1) == cell) continue;  // not the head of its run
            float coef = __ldg(p.w + cell);
            if (p.inj_scale_a) {
                const float hs = p.hdt * (__ldg(p.sigz + zc) + __ldg(p.sigy + yc));
                coef = (1.0f / (1.0f + hs)) * coef;
            }
            float val = um[cell];
            for (int k = e; k < base + p.n_inj && __ldg(p.inj_cell + k) == cell; ++k)
                val = fmaf(coef, __ldg(p.inj_amp + base + __ldg(p.inj_col + k)), val);
            um[cell] = val;
        }
    }
}

template <int H, int MODE, int OCC>
__global__ void __launch_bounds__(kNW * 32, OCC) acoustic_step_kernel(const __grid_constant__ StepParams p) {
    __shared__ __align__(16) float smem[kNW][2 * (32 * 4 + 8) + 64];
    const int tid = threadIdx.x, warp = tid >> 5;

    // ---- sampling duty: record the current field at the sampling cells (before the step) ----
    {
        const int total = p.S * p.n_smp;
        for (int item = blockIdx.x * (kNW * 32) + tid; item < total; item += gridDim.x * (kNW * 32)) {
            const int s = item / p.n_smp;
            const int cell = __ldg(p.smp_cell + item);
            float val = __ldg(p.u + (size_t)s * p.stride + cell);
            if (p.smp_div_a) {
                const int z = cell / p.pitch, y = cell - z * p.pitch;
                const float hs = p.hdt * (__ldg(p.sigz + z) + __ldg(p.sigy + y));
                val = val * (1.0f + hs);
            }
            p.smp_out[item] = val;
        }
    }

    // ---- unit decode: [top/bottom bands | side bands | interior], each shot-major ----
    int id = blockIdx.x * kNW + warp;
    float *sm = smem[warp];
    const int tb_total = p.S * p.n_tb, side_total = p.S * p.n_side, int_total = p.S * p.n_int;
    if (id < tb_total) {
        const int s = id / p.n_tb, t = id - s * p.n_tb;
        const int bc = t / p.tb_strips, strip = t - bc * p.tb_strips;
        const int band = bc / p.tb_chunks, chunk = bc - band * p.tb_chunks;
        const int zb = (band ? p.zhi : p.zs) + chunk * p.bzt;
        march<H, MODE, true, 1, 32, false>(p, s, p.xb + 128 * strip, zb, 0, p.bzt, band ? p.ze : p.zlo,
                                           p.ys, p.ye, sm);
        return;
    }
    id -= tb_total;
    if (id < side_total) {
        // four row-chunks of one 32-column strip per warp (8 lanes x float4 each)
        const int s = id / p.n_side, t = id - s * p.n_side;
        const int per_chunk = p.side_strips_l + p.side_strips_r;
        const int chunk = t / per_chunk, k = t - chunk * per_chunk;
        const int zb = p.zlo + chunk * (4 * p.bzs);
        if (k < p.side_strips_l)
            march<H, MODE, true, 1, 8, false>(p, s, p.xb + 32 * k, zb, p.bzs, p.bzs, p.zhi, p.ys, p.ylo, sm);
        else
            march<H, MODE, true, 1, 8, false>(p, s, p.xr + 32 * (k - p.side_strips_l), zb, p.bzs, p.bzs,
                                              p.zhi, p.yhi, p.ye, sm);
        return;
    }
    id -= side_total;
    if (id < int_total) {
        const int s = id / p.n_int, t = id - s * p.n_int;
        const int chunk = t / p.int_strips, strip = t - chunk * p.int_strips;
        const int zb = p.zlo + chunk * p.bz;
        march<H, MODE, false, 2, 32, true>(p, s, p.ylo + 128 * strip, zb, 0, p.bz, p.zhi, p.ylo, p.yhi, sm);
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// Host side: layout, argument checks, the time loop.
// ------------------------------------------------------------------------------------------
namespace {

inline int round_up(int a, int b) { return (a + b - 1) / b * b; }
inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

int env_int(const char *name, int dflt) {
    const char *e = getenv(name);
    return e ? atoi(e) : dflt;
}

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
    FWI_CHECK_ARG(d->chunk_rows >= 0, "chunk_rows must be >= 0");
    return FWI_OK;
}

// Geometry shared by get_layout and run.
struct Geom {
    int h, F, col0, pitch;
    int zs, ze, ys, ye, zlo, zhi, ylo, yhi, xb, xr;
    int tb_strips, tb_chunks, side_l, side_r, side_chunks, int_strips, int_chunks, bz, bzs, bzt;
    int n_tb, n_side, n_int;
};

Geom make_geom(const fwi_acoustic_desc *d) {
    Geom g;
    g.h = d->order / 2;
    g.F = d->pml_width > 0 ? d->pml_width + g.h : 0;  // frame: sigma != 0 or PML memory within reach
    // first interior column on a 128-byte boundary, at least 4 spare columns on the left
    const int lead = g.h + g.F;
    g.col0 = round_up(lead + 4, 32) - lead;
    g.pitch = round_up(g.col0 + d->nyp + 8, 32);
    g.zs = g.h; g.ze = d->nzp - g.h;
    g.ys = g.col0 + g.h; g.ye = g.col0 + d->nyp - g.h;
    g.zlo = g.zs + g.F; g.zhi = g.ze - g.F;
    g.ylo = g.ys + g.F; g.yhi = g.ye - g.F;
    g.xb = g.ylo - round_up(g.F, 4);
    g.xr = g.yhi / 4 * 4;
    const int rows = g.zhi - g.zlo, cols = g.yhi - g.ylo;
    g.int_strips = ceil_div(cols, 128);
    if (g.F > 0) {
        g.tb_strips = ceil_div(g.ye - g.xb, 128);
        g.side_l = ceil_div(g.ylo - g.xb, 32);
        g.side_r = ceil_div(g.ye - g.xr, 32);
    } else {
        g.tb_strips = g.side_l = g.side_r = 0;
    }
    // interior chunk height: enough units to fill the machine (148 SMs x 12 resident warps),
    // but not so short that the 2H-row window prologue dominates
    int bz = d->chunk_rows > 0 ? d->chunk_rows : env_int("FWI_B200_BZ", 0);
    if (bz <= 0) {
        const int target = env_int("FWI_B200_UNITS", 148 * 12 * 2);
        int chunks = ceil_div(target, d->n_shots * g.int_strips);
        if (chunks < 1) chunks = 1;
        bz = ceil_div(rows, chunks);
        if (bz < 24) bz = 24;
    }
    if (bz > rows) bz = rows;
    bz = round_up(bz, 2);
    g.bz = bz;
    g.int_chunks = ceil_div(rows, bz);
    g.bzs = env_int("FWI_B200_BZS", 8);
    g.bzt = env_int("FWI_B200_BZT", 8);
    g.side_chunks = g.F > 0 ? ceil_div(rows, 4 * g.bzs) : 0;  // a warp marches 4 row-chunks side by side
    g.tb_chunks = g.F > 0 ? ceil_div(g.F, g.bzt) : 0;
    g.n_tb = 2 * g.tb_chunks * g.tb_strips;
    g.n_side = g.side_chunks * (g.side_l + g.side_r);
    g.n_int = g.int_chunks * g.int_strips;
    return g;
}

template <int H, int MODE>
cudaError_t launch_step(const StepParams &p, int blocks, cudaStream_t st) {
    static const int occ = env_int("FWI_B200_OCC", 3);
    if (occ == 3) acoustic_step_kernel<H, MODE, 3><<<blocks, kNW * 32, 0, st>>>(p);
    else acoustic_step_kernel<H, MODE, 4><<<blocks, kNW * 32, 0, st>>>(p);
    return cudaGetLastError();
}

template <int MODE>
cudaError_t launch_h(int h, const StepParams &p, int blocks, cudaStream_t st) {
    switch (h) {
    case 1: return launch_step<1, MODE>(p, blocks, st);
    case 2: return launch_step<2, MODE>(p, blocks, st);
    case 3: return launch_step<3, MODE>(p, blocks, st);
    default: return launch_step<4, MODE>(p, blocks, st);
    }
}

}  // namespace

extern "C" int fwi_acoustic_get_layout(const fwi_acoustic_desc *desc, fwi_acoustic_layout *out) {
    int rc = check_desc(desc);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(out != nullptr, "layout output is NULL");
    const Geom g = make_geom(desc);
    out->pitch = g.pitch;
    out->col0 = g.col0;
    out->shot_stride = (int64_t)desc->nzp * g.pitch;
    out->n_images = desc->n_shots;
    out->wave_fields = desc->pml_width > 0 ? 6 : 2;
    out->chunk_rows = g.bz;
    out->units_per_shot = g.n_tb + g.n_side + g.n_int;
    return FWI_OK;
}

extern "C" int fwi_acoustic_run(const fwi_acoustic_desc *d, const fwi_acoustic_buffers *b, void *stream) {
    int rc = check_desc(d);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(b != nullptr, "buffers struct is NULL");
    FWI_CHECK_ARG(b->w && b->wave, "w and wave must not be NULL");
    FWI_CHECK_ARG(d->pml_width == 0 || (b->sigz && b->sigy), "sigz/sigy required when pml_width > 0");
    FWI_CHECK_ARG(!d->smp_div_a || (b->sigz && b->sigy), "sigz/sigy required when smp_div_a is set");
    FWI_CHECK_ARG(!d->inj_scale_a || (b->sigz && b->sigy), "sigz/sigy required when inj_scale_a is set");
    FWI_CHECK_ARG(d->n_smp == 0 || (b->smp_cell && b->smp_out), "smp_cell/smp_out required when n_smp > 0");
    FWI_CHECK_ARG(d->n_inj == 0 || (b->inj_row_start && b->inj_cell && b->inj_col && b->inj_amp),
                  "injection lists required when n_inj > 0");
    FWI_CHECK_ARG(!(b->q_write && b->q_read), "q_write and q_read are mutually exclusive");
    FWI_CHECK_ARG(!b->q_read || b->image, "image required with q_read");
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    const Geom g = make_geom(d);
    const int h = g.h;

    StepParams p;
    p.nzp = d->nzp; p.pitch = g.pitch;
    p.zs = g.zs; p.ze = g.ze; p.ys = g.ys; p.ye = g.ye;
    p.zlo = g.zlo; p.zhi = g.zhi; p.ylo = g.ylo; p.yhi = g.yhi;
    p.xb = g.xb; p.xr = g.xr;
    p.stride = (long long)d->nzp * g.pitch;
    p.S = d->n_shots;
    p.pml = d->pml_width > 0;
    p.tb_strips = g.tb_strips; p.tb_chunks = g.tb_chunks; p.bzt = g.bzt; p.side_strips_l = g.side_l; p.side_strips_r = g.side_r;
    p.side_chunks = g.side_chunks; p.int_strips = g.int_strips; p.int_chunks = g.int_chunks;
    p.bz = g.bz; p.bzs = g.bzs;
    p.n_tb = g.n_tb; p.n_side = g.n_side; p.n_int = g.n_int;
    p.w = b->w; p.sigz = b->sigz; p.sigy = b->sigy;
    p.inj_row_start = b->inj_row_start; p.inj_cell = b->inj_cell; p.inj_col = b->inj_col;
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

    const size_t fld = (size_t)d->n_shots * p.stride;
    float *u = b->wave, *um = b->wave + fld;
    float *pz = nullptr, *pzn = nullptr, *py = nullptr, *pyn = nullptr;
    if (p.pml) { pz = b->wave + 2 * fld; pzn = b->wave + 3 * fld; py = b->wave + 4 * fld; pyn = b->wave + 5 * fld; }

    const int mode = b->q_read ? 2 : (b->q_write ? 1 : 0);
    const long long units = (long long)d->n_shots * (g.n_tb + g.n_side + g.n_int);
    const int blocks = (int)((units + kNW - 1) / kNW);
    for (int it = 0; it < d->nsteps; ++it) {
        const int n = d->adjoint ? d->nsteps - 1 - it : it;
        p.u = u; p.um = um; p.pz = pz; p.pzn = pzn; p.py = py; p.pyn = pyn;
        p.qw = b->q_write ? b->q_write + (size_t)n * fld : nullptr;
        p.qr = b->q_read ? b->q_read + (size_t)n * fld : nullptr;
        p.G = b->image;
        p.inj_amp = d->n_inj ? b->inj_amp + (size_t)n * d->n_shots * d->n_inj : nullptr;
        p.smp_out = d->n_smp ? b->smp_out + (size_t)n * d->n_shots * d->n_smp : nullptr;
        cudaError_t e;
        if (mode == 2) e = launch_h<2>(h, p, blocks, st);
        else if (mode == 1) e = launch_h<1>(h, p, blocks, st);
        else e = launch_h<0>(h, p, blocks, st);
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
