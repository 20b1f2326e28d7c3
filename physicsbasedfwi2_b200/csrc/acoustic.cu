// 2-D constant-density acoustic propagator for sm_100a: one fused kernel launch per time step.
//
// Replaces the time loop of deepwave.scalar.Propagator.forward / backward as called at
// reference models/networks.py:5464 (forward) and :5491 (adjoint).  Not a port: deepwave's
// kernels are one-thread-per-cell global-memory stencils.  Here the grid of one shot is cut
// into column strips and every WARP owns a contiguous run of strip rows (its "segments"), which
// it marches down in z as its own software pipeline:
//
//   * rows are brought from HBM/L2 into a per-warp shared-memory ring by TMA bulk copies
//     (cp.async.bulk, completion on an mbarrier) issued DO rows ahead of the row being computed,
//     so the bytes in flight per SM are set by the ring depth, not by registers or occupancy;
//   * the 2H+1 rows of u needed by the z-stencil are kept in REGISTERS (a sliding window), the
//     y-neighbours of the centre row are read from the ring (the row lands with a 4-column halo),
//     so every operand crosses HBM once per step;
//   * warps never wait for one another: no block-level barrier anywhere, each warp waits only on
//     the mbarriers of its own ring;
//   * the step fuses receiver sampling + stencil/PML update + source injection + snapshot write
//     (forward) or snapshot read + imaging accumulation (adjoint);
//   * the host cuts the work into exactly one cost-balanced unit per resident warp (a table of
//     segments), so a launch is a single wave with no tail.
//
// Three segment types share the kernel: interior strips (128 columns, float4 per lane, no PML
// arithmetic at all, ring depth 4), top/bottom PML band strips (128 columns, full PML update, the
// z-window of the PML memory variable lives in a second ring) and left/right PML band strips
// (32 columns, direct loads).  The layout puts the first interior column on a 128-byte boundary.
//
// Arithmetic order per cell follows oracle/cpu_ref_impl.h (explicit fma, no contraction).
#include <cuda.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <thread>
#include <type_traits>
#include <vector>

#include "common.h"
#include "acoustic_small.h"

namespace {

// warps per CTA and resident CTAs per SM (warps are independent; this is scheduling granularity):
// forward / advance-only 2 x 6 warps with a 3-deep ring, imaging launches (more operands per row) 2 x 4 warps, 2-deep
// (the FWI_CFG_* macros exist so that scripts/dev_variants.sh can build tuning variants; the defaults are the product)
#ifndef FWI_CFG_W1
#define FWI_CFG_W1 6
#define FWI_CFG_C1 2
#define FWI_CFG_D1 3
#endif
#ifndef FWI_CFG_W2
#define FWI_CFG_W2 4      // 2 x 4 warps: a multiple of the 4 schedulers (3 x 3 measured 3.5 % slower: uneven warps per scheduler)
#define FWI_CFG_C2 2
#define FWI_CFG_D2 2
#endif
#ifndef FWI_CFG_W3
#define FWI_CFG_W3 4
#define FWI_CFG_C3 2
#define FWI_CFG_D3 2
#endif
// MODE 0: advance only; 1: advance + write the snapshot q; 2: read q_n, image, advance; 3: read q_n and q_{n+1},
// image both steps (the pairing of run_impl), advance
__host__ __device__ constexpr int warps_per_cta(int MODE) { return MODE == 3 ? FWI_CFG_W3 : (MODE == 2 ? FWI_CFG_W2 : FWI_CFG_W1); }
__host__ __device__ constexpr int ctas_per_sm(int MODE) { return MODE == 3 ? FWI_CFG_C3 : (MODE == 2 ? FWI_CFG_C2 : FWI_CFG_C1); }
constexpr int kSegInt = 0, kSegBand = 1, kSegSide = 2;
constexpr int kMaxUnits = 16384;   // bound on resident warps of any device this could run on

// one run of rows [zb, ze) of a column strip starting at device column x0 of shot s; only columns
// in [cmin, cmax) are stored
struct Seg {
    int s, x0, zb, ze;
    int type, cmin, cmax, pad;
};

struct StepParams {
    int nzp, pitch;
    int zs, ze, ys, ye;      // computational region (device coordinates, end exclusive)
    long long stride;        // floats per shot per field
    int S;
    const Seg *segs;
    const int *unit_start;   // [n_units + 1] segment ranges of the warp units
    int n_units;
    unsigned long long *trace;   // dev aid (FWI_B200_TRACE): [n_units][2] start / end %globaltimer of each unit
    // TMA descriptors (2-D, rows x pitch, boxes of kRB rows): the wave workspace with 136- and
    // 128-column boxes, w, the snapshots being read and the image; and the first row of each
    // wavefield of this step inside the wave workspace / of this step's snapshot
    CUtensorMap tm_wave_u, tm_wave_o, tm_w, tm_q, tm_g;
    CUtensorMap ts_wave_u, ts_wave_o, ts_w, ts_q, ts_g;   // the same tensors with the narrow boxes of side strips
    int row_u, row_um, row_pz, row_py, row_q, row_q2;   // row_q2: snapshot of the step imaged together with this one (MODE 3)
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
// plain (coherent) 128-bit load: wavefields are written by the previous grid, which may still be running when this one
// becomes resident under programmatic dependent launch - ld.global.nc is only for data that is read-only for the whole
// lifetime of the grid (w, sigma, the index lists)
__device__ __forceinline__ F4 ld_field(const float *p) { return f4(*reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ F4 lds4(const float *p) { return f4(*reinterpret_cast<const float4 *>(p)); }
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

// ---- mbarrier / TMA bulk copy (per-warp pipelines) -------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void *p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
// global -> shared TMA tile copy of one box of the 2-D tensor `tm` whose corner is (x, y)
__device__ __forceinline__ void tma_g2s(float *dst, const CUtensorMap *tm, int x, int y, unsigned bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            smem_u32(dst)),
        "l"(tm), "r"(x), "r"(y), "r"(bar)
        : "memory");
}

// ---- per-warp shared-memory budget ------------------------------------------------------------
constexpr int kRB = 4;        // rows per TMA box / ring block (>= H)
constexpr int kRowU = 136;    // a ring row of u / py: 4-column halo | 128 own columns | 4-column halo
constexpr int kRowO = 128;    // a ring row of everything else
constexpr int kBlkU = kRB * kRowU, kBlkO = kRB * kRowO;   // floats per ring block
__host__ __device__ constexpr int do_int(int MODE) { return MODE == 3 ? FWI_CFG_D3 : (MODE == 2 ? FWI_CFG_D2 : FWI_CFG_D1); }   // ring depth (blocks in flight) of interior segments
#ifndef FWI_CFG_DBAND
#define FWI_CFG_DBAND 1
#endif
constexpr int kDOBand = FWI_CFG_DBAND;    // ring depth of top/bottom band segments
constexpr int kBarFloats = 32;  // 8 mbarriers at the head of the warp's region (keeps blocks 128-byte aligned)

// operand blocks fetched per iteration by side strips: u-, w (+ snapshot, image in the adjoint)
__host__ __device__ constexpr int n_ops(int MODE) { return MODE == 3 ? 5 : (MODE == 2 ? 4 : 2); }
// ... and by 128-column strips, which read w (shared by all shots, L2 resident) with plain loads
__host__ __device__ constexpr int n_ops_ring(int MODE) { return MODE == 3 ? 4 : (MODE == 2 ? 3 : 1); }
__host__ __device__ constexpr int ring_floats(int MODE, bool PML, int DO) {
    return (DO + 1) * kBlkU + DO * n_ops_ring(MODE) * kBlkO + (PML ? (DO + 2) * kBlkO + DO * kBlkU : 0);
}
constexpr int kSideU = 40, kSideO = 32;          // ring row widths of side strips
constexpr int kSideBlkU = kRB * kSideU, kSideBlkO = kRB * kSideO;
constexpr int kDOSide = 3;
__host__ __device__ constexpr int side_ring_floats(int MODE) {
    return (kDOSide + 2) * (kSideBlkU + kSideBlkO) + kDOSide * (n_ops(MODE) * kSideBlkO + kSideBlkU);
}
__host__ __device__ constexpr int cmax_i(int a, int b) { return a > b ? a : b; }
constexpr int kWRing = 2 * kBlkO;   // w = v^2 dt^2 blocks (L2 resident: one block of lead is enough), own mbarriers 6, 7
__host__ __device__ constexpr int warp_smem_floats(int MODE) {
    return kBarFloats + kWRing + cmax_i(cmax_i(ring_floats(MODE, false, do_int(MODE)), ring_floats(MODE, true, kDOBand)),
                               side_ring_floats(MODE));
}

// ---- injection into the freshly written field: points of shot s inside rows [zb, ze), columns
// [c_lo, c_hi); lanes lx of LW take the list entries round robin ----
__device__ __forceinline__ void inject(const StepParams &p, const int s, float *__restrict__ um, const int zb,
                                       const int ze, const int c_lo, const int c_hi, const int lx, const int LW) {
    const int *rs = p.inj_row_start + (size_t)s * (p.nzp + 1);
    const int base = s * p.n_inj;
    const int zq = min(zb, ze);
    const int e0 = __ldg(rs + zq), e1 = __ldg(rs + ze);
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

// One warp marches down z through rows [zb, ze) of the 128-column strip at x0 of shot sg.s in blocks
// of kRB rows, fed by TMA tile copies into its shared-memory ring `sm`; bar0 + 8*i, i < DO, are the
// ring mbarriers, bar0 + 8*DO the prologue mbarrier, `phase` holds their parity bits.
// Ring blocks (k = block index, b = iteration block, rows relative to the segment start zb):
//   u   block k: rows [H - kRB + k*kRB, +kRB)      k = 0 prologue, k = b + 1 arrives with iteration b
//   pz  block k: rows [H - 2kRB + k*kRB, +kRB)     k = 0, 1 prologue, k = b + 2 arrives with iteration b
//   operands (u-, w, snapshot, image) and py block b: rows [b*kRB, +kRB)
// MODE 0: advance only.  MODE 1: advance and write q (forward with gradient).
// MODE 2: read q, accumulate the image, advance (adjoint with model gradient).
template <int H, int MODE, bool PML, int DO>
__device__ __forceinline__ void march_ring(const StepParams &p, const Seg &sg, float *sm, float *WR,
                                           const unsigned bar0, unsigned &phase) {
    static_assert(H <= kRB, "a ring block must hold the stencil half width");
    constexpr int RB = kRB;
    constexpr int DUB = DO + 1, DPB = DO + 2;
    constexpr int NOPS = n_ops_ring(MODE);
    float *const U = sm;
    float *const OPS = U + DUB * kBlkU;
    float *const PZ = OPS + DO * NOPS * kBlkO;
    float *const PY = PZ + DPB * kBlkO;
    const int lane = threadIdx.x & 31;
    const int s = sg.s, x0 = sg.x0, zb = sg.zb, ze = sg.ze;
    const int nblk = (ze - zb + RB - 1) / RB;
    const int col = x0 + 4 * lane;
    const int colL = min(col, p.pitch - 4);
    const size_t sb = (size_t)s * p.stride;
    const float *__restrict__ u = p.u + sb;
    float *__restrict__ um = p.um + sb;
    const int srow = s * p.nzp;          // first tensor row of this shot inside a field
    constexpr unsigned tx_blk = 4u * (kBlkU + NOPS * kBlkO + (PML ? kBlkO + kBlkU : 0));

    // lane 0: the copies of the iteration block whose first centre row is z0 into operand slot so
    // (which names the barrier too), u slot su and pz slot sp
    auto issue = [&](const int z0, const int so, const int su, const int sp) {
        const unsigned bar = bar0 + 8u * so;
        mbar_expect_tx(bar, tx_blk);
        tma_g2s(U + su * kBlkU, &p.tm_wave_u, x0 - 4, p.row_u + srow + z0 + H, bar);
        float *o = OPS + so * NOPS * kBlkO;
        tma_g2s(o, &p.tm_wave_o, x0, p.row_um + srow + z0, bar);
        if (MODE >= 2) {
            tma_g2s(o + kBlkO, &p.tm_q, x0, p.row_q + srow + z0, bar);
            tma_g2s(o + 2 * kBlkO, &p.tm_g, x0, srow + z0, bar);
        }
        if (MODE == 3) tma_g2s(o + 3 * kBlkO, &p.tm_q, x0, p.row_q2 + srow + z0, bar);
        if (PML) {
            tma_g2s(PZ + sp * kBlkO, &p.tm_wave_o, x0, p.row_pz + srow + z0 + H, bar);
            tma_g2s(PY + so * kBlkU, &p.tm_wave_u, x0 - 4, p.row_py + srow + z0, bar);
        }
    };

    // ---- prologue: u block 0, pz blocks 0 and 1, and the first DO iteration blocks ----
    const unsigned barP = bar0 + 8u * DO;
    if (lane == 0) {
        mbar_expect_tx(barP, 4u * (kBlkU + (PML ? 2 * kBlkO : 0)));
        tma_g2s(U, &p.tm_wave_u, x0 - 4, p.row_u + srow + zb + H - RB, barP);
        if (PML) {
            tma_g2s(PZ, &p.tm_wave_o, x0, p.row_pz + srow + zb + H - 2 * RB, barP);
            tma_g2s(PZ + kBlkO, &p.tm_wave_o, x0, p.row_pz + srow + zb + H - RB, barP);
        }
        for (int j = 0; j < DO && j < nblk; ++j) issue(zb + j * RB, j, (j + 1) % DUB, (j + 2) % DPB);
        mbar_expect_tx(bar0 + 8u * 6, 4u * kBlkO);
        tma_g2s(WR, &p.tm_w, x0, zb, bar0 + 8u * 6);
    }

    // sliding window: inside iteration block b (first centre row z0), W[i] = u row z0 - H + i
    F4 W[2 * H + RB];
#pragma unroll
    for (int k = 0; k < 2 * H - RB; ++k) W[RB + k] = ld_field(u + (size_t)(zb - H + k) * p.pitch + colL);
    F4 sy;
#pragma unroll
    for (int j = 0; j < 4; ++j) sy.a[j] = PML ? __ldg(p.sigy + colL + j) : 0.f;
    // sigma_y == 0 on all 128 columns of the strip (a top / bottom band strip away from the corners): the coefficient
    // fields are then the same for every cell of a row and are evaluated once per row - the general expressions with
    // sigma_y = 0, so the same bits
    const bool sy_zero = PML && __all_sync(0xffffffffu, sy.a[0] == 0.f && sy.a[1] == 0.f && sy.a[2] == 0.f && sy.a[3] == 0.f);
    mbar_wait(barP, (phase >> DO) & 1u);
    phase ^= 1u << DO;
#pragma unroll
    for (int k = (2 * H - RB > 0 ? 2 * H - RB : 0); k < 2 * H; ++k)
        W[RB + k] = lds4(U + (k - 2 * H + RB) * kRowU + 4 + 4 * lane);

    int so = 0;                  // operand / barrier slot of this iteration block
    int su_prev = 0, su = 1 % DUB;   // u slots of blocks b and b + 1
    int sp0 = 0, sp1 = 1, sp2 = 2 % DPB;   // pz slots of blocks b, b + 1, b + 2
    for (int b = 0; b < nblk; ++b) {
        const int z0 = zb + b * RB;
        // w of the next block goes to the other half of its two-block ring (its last reader was
        // iteration b - 1), on its own barrier
        const int wsl = b & 1;
        if (lane == 0 && b + 1 < nblk) {
            const unsigned wbar = bar0 + 8u * (6 + (wsl ^ 1));
            mbar_expect_tx(wbar, 4u * kBlkO);
            tma_g2s(WR + (wsl ^ 1) * kBlkO, &p.tm_w, x0, z0 + RB, wbar);
        }
        mbar_wait(bar0 + 8u * so, (phase >> so) & 1u);
        phase ^= 1u << so;
        mbar_wait(bar0 + 8u * (6 + wsl), (phase >> (6 + wsl)) & 1u);
        phase ^= 1u << (6 + wsl);
#pragma unroll
        for (int k = 0; k < 2 * H; ++k) W[k] = W[k + RB];
#pragma unroll
        for (int r = 0; r < RB; ++r) W[2 * H + r] = lds4(U + su * kBlkU + r * kRowU + 4 + 4 * lane);
        const float *const pzb[3] = {PZ + sp0 * kBlkO, PZ + sp1 * kBlkO, PZ + sp2 * kBlkO};

#pragma unroll
        for (int r = 0; r < RB; ++r) {
            const int row = z0 + r;
            if (row < ze) {
                const size_t off = (size_t)row * p.pitch;
                // the centre row sits H rows above the bottom of the window: u block b or b + 1
                const float *uc = (r < H) ? U + su_prev * kBlkU + (r - H + RB) * kRowU : U + su * kBlkU + (r - H) * kRowU;
                float v[12];
                {
                    const float4 L = *reinterpret_cast<const float4 *>(uc + 4 * lane);
                    const float4 R = *reinterpret_cast<const float4 *>(uc + 8 + 4 * lane);
                    v[0] = L.x; v[1] = L.y; v[2] = L.z; v[3] = L.w;
                    v[8] = R.x; v[9] = R.y; v[10] = R.z; v[11] = R.w;
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) v[4 + q] = W[r + H].a[q];
                const float *o = OPS + so * NOPS * kBlkO + r * kRowO + 4 * lane;
                const F4 cM = lds4(o), cW = lds4(WR + wsl * kBlkO + r * kRowO + 4 * lane);

                F4 lap;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float l = p.c0 * v[4 + q];
#pragma unroll
                    for (int k = 1; k <= H; ++k) l = fmaf(p.cy2[k], v[4 + q - k] + v[4 + q + k], l);
#pragma unroll
                    for (int k = 1; k <= H; ++k) l = fmaf(p.cz2[k], W[r + H - k].a[q] + W[r + H + k].a[q], l);
                    lap.a[q] = l;
                }

                F4 un, qv;
                if (!PML) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        qv.a[q] = lap.a[q];
                        un.a[q] = fmaf(cW.a[q], lap.a[q], fmaf(2.0f, v[4 + q], -cM.a[q]));
                    }
                } else {
                    const float sz = __ldg(p.sigz + row);
                    // pz row (row + d) sits at local index r + d - H + 2 RB of the three live pz blocks
                    auto pzrow = [&](const int d) -> const float * {
                        const int l = r + d - H + 2 * RB;
                        return pzb[l / RB] + (l % RB) * kRowO + 4 * lane;
                    };
                    const F4 pzc = lds4(pzrow(0));
                    F4 dpz;
#pragma unroll
                    for (int q = 0; q < 4; ++q) dpz.a[q] = 0.f;
#pragma unroll
                    for (int k = 1; k <= H; ++k) {
                        const F4 a = lds4(pzrow(k)), bb = lds4(pzrow(-k));
#pragma unroll
                        for (int q = 0; q < 4; ++q) dpz.a[q] = fmaf(p.cz1[k], a.a[q] - bb.a[q], dpz.a[q]);
                    }
                    float pyv[12];
                    {
                        const float *pr = PY + so * kBlkU + r * kRowU + 4 * lane;
                        const F4 t0 = lds4(pr), t1 = lds4(pr + 4), t2 = lds4(pr + 8);
#pragma unroll
                        for (int q = 0; q < 4; ++q) { pyv[q] = t0.a[q]; pyv[4 + q] = t1.a[q]; pyv[8 + q] = t2.a[q]; }
                    }
                    F4 pzn, pyn;
                    float duz[4], dpy[4], duy[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        duz[q] = 0.f; dpy[q] = 0.f; duy[q] = 0.f;
#pragma unroll
                        for (int k = 1; k <= H; ++k) {
                            duz[q] = fmaf(p.cz1[k], W[r + H + k].a[q] - W[r + H - k].a[q], duz[q]);
                            dpy[q] = fmaf(p.cy1[k], pyv[4 + q + k] - pyv[4 + q - k], dpy[q]);
                            duy[q] = fmaf(p.cy1[k], v[4 + q + k] - v[4 + q - k], duy[q]);
                        }
                        qv.a[q] = (lap.a[q] + dpz.a[q]) + dpy[q];
                    }
                    if (sy_zero) {       // warp-uniform: straight-line code on both sides, the four cells interleave
                        const float hs = p.hdt * sz;
                        const float a0 = 1.0f / (1.0f + hs);
                        const float cu0 = a0 * 2.0f;
                        const float cum0 = a0 * (hs - 1.0f);
                        const float cpz1 = p.dt * (0.0f - sz);
                        const float cpz2 = fmaf(-p.dt, sz, 1.0f);
                        const float cpy1 = p.dt * sz;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            un.a[q] = fmaf(a0 * cW.a[q], qv.a[q], fmaf(cu0, v[4 + q], cum0 * cM.a[q]));
                            pzn.a[q] = fmaf(cpz1, duz[q], cpz2 * pzc.a[q]);
                            pyn.a[q] = fmaf(cpy1, duy[q], pyv[4 + q]);
                        }
                    } else {
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const float hs = p.hdt * (sz + sy.a[q]);
                            const float a = 1.0f / (1.0f + hs);
                            const float cu = a * fmaf(-p.dt2, sz * sy.a[q], 2.0f);
                            const float cum = a * (hs - 1.0f);
                            const float aw = a * cW.a[q];
                            un.a[q] = fmaf(aw, qv.a[q], fmaf(cu, v[4 + q], cum * cM.a[q]));
                            pzn.a[q] = fmaf(p.dt * (sy.a[q] - sz), duz[q], fmaf(-p.dt, sz, 1.0f) * pzc.a[q]);
                            pyn.a[q] = fmaf(p.dt * (sz - sy.a[q]), duy[q], fmaf(-p.dt, sy.a[q], 1.0f) * pyv[4 + q]);
                        }
                    }
                    st_masked<false>(p.pzn + sb + off, col, pzn, sg.cmin, sg.cmax, true);
                    st_masked<false>(p.pyn + sb + off, col, pyn, sg.cmin, sg.cmax, true);
                }
                st_masked<false>(um + off, col, un, sg.cmin, sg.cmax, true);
                if (MODE == 1) st_masked<true>(p.qw + sb + off, col, qv, sg.cmin, sg.cmax, true);
                if (MODE >= 2) {
                    // image of this shot: G += field * q_read (single owner per cell: deterministic).  MODE 3 first
                    // adds the product the previous step left out: its field is this step's u- (cM), in the same
                    // order of additions as one imaging per step
                    const F4 cQ = lds4(o + kBlkO);
                    F4 g = lds4(o + 2 * kBlkO);
                    if (MODE == 3) {
                        const F4 cQ2 = lds4(o + 3 * kBlkO);
#pragma unroll
                        for (int q = 0; q < 4; ++q) g.a[q] = fmaf(cM.a[q], cQ2.a[q], g.a[q]);
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) g.a[q] = fmaf(v[4 + q], cQ.a[q], g.a[q]);
                    st_masked<false>(p.G + sb + off, col, g, sg.cmin, sg.cmax, true);
                }
            }
        }

        // ---- refill: the slots this iteration block has finished with take block b + DO ----
        __syncwarp();
        if (lane == 0 && b + DO < nblk) issue(z0 + DO * RB, so, su_prev, sp0);
        if (++so == DO) so = 0;
        su_prev = su;
        if (++su == DUB) su = 0;
        sp0 = sp1; sp1 = sp2;
        if (++sp2 == DPB) sp2 = 0;
    }

    if (p.n_inj > 0) {
        __syncwarp();  // orders this warp's stores above before the read-modify-writes below
        inject(p, s, um, zb, ze, max(sg.cmin, x0), min(sg.cmax, x0 + 128), lane, 32);
    }
    __syncwarp();
}

// Left/right PML band strips (32 columns), also TMA fed.  The 32 lanes form 4 lane groups of 8
// lanes; lane group g computes row z0 + g of every block of kRB = 4 rows, so a block iteration
// covers 4 rows x 32 columns.  Nothing is kept in registers between blocks: the z-windows of u and
// pz are read from rings that hold three live blocks (rows z0 - kRB .. z0 + 2 kRB) plus DO in
// flight.  Block k of the u / pz rings holds rows [(k - 1) kRB, k kRB) relative to zb.

template <int H, int MODE>
__device__ __forceinline__ void march_side(const StepParams &p, const Seg &sg, float *sm, const unsigned bar0,
                                           unsigned &phase) {
    constexpr int RB = kRB, DO = kDOSide, D = DO + 2;
    constexpr int NOPS = n_ops(MODE);
    float *const U = sm;
    float *const PZ = U + D * kSideBlkU;
    float *const OPS = PZ + D * kSideBlkO;
    float *const PY = OPS + DO * NOPS * kSideBlkO;
    const int lane = threadIdx.x & 31;
    const int lx = lane & 7, grp = lane >> 3;
    const int s = sg.s, x0 = sg.x0, zb = sg.zb, ze = sg.ze, cmin = sg.cmin, cmax = sg.cmax;
    const int nblk = (ze - zb + RB - 1) / RB;
    const int col = x0 + 4 * lx;
    const int colL = min(col, p.pitch - 4);
    const size_t sb = (size_t)s * p.stride;
    float *__restrict__ um = p.um + sb;
    const int srow = s * p.nzp;
    constexpr unsigned tx_blk = 4u * (kSideBlkU + kSideBlkO + NOPS * kSideBlkO + kSideBlkU);

    auto issue = [&](const int z0, const int so, const int sk) {
        const unsigned bar = bar0 + 8u * so;
        mbar_expect_tx(bar, tx_blk);
        tma_g2s(U + sk * kSideBlkU, &p.ts_wave_u, x0 - 4, p.row_u + srow + z0 + RB, bar);
        tma_g2s(PZ + sk * kSideBlkO, &p.ts_wave_o, x0, p.row_pz + srow + z0 + RB, bar);
        float *o = OPS + so * NOPS * kSideBlkO;
        tma_g2s(o, &p.ts_wave_o, x0, p.row_um + srow + z0, bar);
        tma_g2s(o + kSideBlkO, &p.ts_w, x0, z0, bar);
        if (MODE >= 2) {
            tma_g2s(o + 2 * kSideBlkO, &p.ts_q, x0, p.row_q + srow + z0, bar);
            tma_g2s(o + 3 * kSideBlkO, &p.ts_g, x0, srow + z0, bar);
        }
        if (MODE == 3) tma_g2s(o + 4 * kSideBlkO, &p.ts_q, x0, p.row_q2 + srow + z0, bar);
        tma_g2s(PY + so * kSideBlkU, &p.ts_wave_u, x0 - 4, p.row_py + srow + z0, bar);
    };

    const unsigned barP = bar0 + 8u * 4;
    if (lane == 0) {
        mbar_expect_tx(barP, 4u * 2 * (kSideBlkU + kSideBlkO));
        tma_g2s(U, &p.ts_wave_u, x0 - 4, p.row_u + srow + zb - RB, barP);
        tma_g2s(U + kSideBlkU, &p.ts_wave_u, x0 - 4, p.row_u + srow + zb, barP);
        tma_g2s(PZ, &p.ts_wave_o, x0, p.row_pz + srow + zb - RB, barP);
        tma_g2s(PZ + kSideBlkO, &p.ts_wave_o, x0, p.row_pz + srow + zb, barP);
        for (int j = 0; j < DO && j < nblk; ++j) issue(zb + j * RB, j, (j + 2) % D);
    }
    F4 sy;
#pragma unroll
    for (int j = 0; j < 4; ++j) sy.a[j] = __ldg(p.sigy + colL + j);
    // coefficients of a row with sigma_z == 0 (all rows of a side segment), per column: a, cu = 2a, cum,
    // dt (sy - 0), dt (0 - sy), 1 - dt sy - the general expressions evaluated with sigma_z = 0
    F4 ca, ccu, ccum, cdts, cpy1, cpy2;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float hs = p.hdt * sy.a[j];
        ca.a[j] = 1.0f / (1.0f + hs);
        ccu.a[j] = ca.a[j] * 2.0f;
        ccum.a[j] = ca.a[j] * (hs - 1.0f);
        cdts.a[j] = p.dt * sy.a[j];
        cpy1.a[j] = p.dt * (0.0f - sy.a[j]);
        cpy2.a[j] = fmaf(-p.dt, sy.a[j], 1.0f);
    }
    mbar_wait(barP, (phase >> 4) & 1u);
    phase ^= 1u << 4;

    int so = 0, s0 = 0;          // operand slot; ring slot of block b (rows z0 - RB ..)
    for (int b = 0; b < nblk; ++b) {
        const int z0 = zb + b * RB;
        mbar_wait(bar0 + 8u * so, (phase >> so) & 1u);
        phase ^= 1u << so;
        const int row = z0 + grp;
        // ring row of (row + d): local index l = grp + d + RB inside the three live blocks
        auto urow = [&](const int d) -> const float * {
            const int l = grp + d + RB;
            int sl = s0 + (l >> 2);
            if (sl >= D) sl -= D;
            return U + sl * kSideBlkU + (l & 3) * kSideU + 4 * lx;
        };
        auto pzrow = [&](const int d) -> const float * {
            const int l = grp + d + RB;
            int sl = s0 + (l >> 2);
            if (sl >= D) sl -= D;
            return PZ + sl * kSideBlkO + (l & 3) * kSideO + 4 * lx;
        };
        float v[12];
        {
            const float *uc = urow(0);
            const F4 t0 = lds4(uc), t1 = lds4(uc + 4), t2 = lds4(uc + 8);
#pragma unroll
            for (int q = 0; q < 4; ++q) { v[q] = t0.a[q]; v[4 + q] = t1.a[q]; v[8 + q] = t2.a[q]; }
        }
        const float *o = OPS + so * NOPS * kSideBlkO + grp * kSideO + 4 * lx;
        const F4 cM = lds4(o), cW = lds4(o + kSideBlkO);
        F4 lap, duz4, dpz;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float l = p.c0 * v[4 + q];
#pragma unroll
            for (int k = 1; k <= H; ++k) l = fmaf(p.cy2[k], v[4 + q - k] + v[4 + q + k], l);
            lap.a[q] = l;
            duz4.a[q] = 0.f;
            dpz.a[q] = 0.f;
        }
#pragma unroll
        for (int k = 1; k <= H; ++k) {
            const F4 up = lds4(urow(k) + 4), dn = lds4(urow(-k) + 4);
            const F4 a = lds4(pzrow(k)), bb = lds4(pzrow(-k));
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                lap.a[q] = fmaf(p.cz2[k], dn.a[q] + up.a[q], lap.a[q]);
                duz4.a[q] = fmaf(p.cz1[k], up.a[q] - dn.a[q], duz4.a[q]);
                dpz.a[q] = fmaf(p.cz1[k], a.a[q] - bb.a[q], dpz.a[q]);
            }
        }
        const F4 pzc = lds4(pzrow(0));
        float pyv[12];
        {
            const float *pr = PY + so * kSideBlkU + grp * kSideU + 4 * lx;
            const F4 t0 = lds4(pr), t1 = lds4(pr + 4), t2 = lds4(pr + 8);
#pragma unroll
            for (int q = 0; q < 4; ++q) { pyv[q] = t0.a[q]; pyv[4 + q] = t1.a[q]; pyv[8 + q] = t2.a[q]; }
        }
        const bool rowok = row < ze;
        const float sz = __ldg(p.sigz + min(row, p.ze - 1));
        const size_t off = (size_t)min(row, p.ze - 1) * p.pitch;
        F4 un, qv, pzn, pyn;
        float duy4[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float dpy = 0.f, duy = 0.f;
#pragma unroll
            for (int k = 1; k <= H; ++k) {
                dpy = fmaf(p.cy1[k], pyv[4 + q + k] - pyv[4 + q - k], dpy);
                duy = fmaf(p.cy1[k], v[4 + q + k] - v[4 + q - k], duy);
            }
            qv.a[q] = (lap.a[q] + dpz.a[q]) + dpy;
            duy4[q] = duy;
        }
        if (__all_sync(0xffffffffu, sz == 0.f)) {
            // (every row of a side segment: the column coefficients of the segment prologue, same bits)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                un.a[q] = fmaf(ca.a[q] * cW.a[q], qv.a[q], fmaf(ccu.a[q], v[4 + q], ccum.a[q] * cM.a[q]));
                pzn.a[q] = fmaf(cdts.a[q], duz4.a[q], pzc.a[q]);
                pyn.a[q] = fmaf(cpy1.a[q], duy4[q], cpy2.a[q] * pyv[4 + q]);
            }
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float hs = p.hdt * (sz + sy.a[q]);
                const float a = 1.0f / (1.0f + hs);
                const float cu = a * fmaf(-p.dt2, sz * sy.a[q], 2.0f);
                const float cum = a * (hs - 1.0f);
                const float aw = a * cW.a[q];
                un.a[q] = fmaf(aw, qv.a[q], fmaf(cu, v[4 + q], cum * cM.a[q]));
                pzn.a[q] = fmaf(p.dt * (sy.a[q] - sz), duz4.a[q], fmaf(-p.dt, sz, 1.0f) * pzc.a[q]);
                pyn.a[q] = fmaf(p.dt * (sz - sy.a[q]), duy4[q], fmaf(-p.dt, sy.a[q], 1.0f) * pyv[4 + q]);
            }
        }
        st_masked<false>(p.pzn + sb + off, col, pzn, cmin, cmax, rowok);
        st_masked<false>(p.pyn + sb + off, col, pyn, cmin, cmax, rowok);
        st_masked<false>(um + off, col, un, cmin, cmax, rowok);
        if (MODE == 1) st_masked<true>(p.qw + sb + off, col, qv, cmin, cmax, rowok);
        if (MODE >= 2) {
            const F4 cQ = lds4(o + 2 * kSideBlkO);
            F4 g = lds4(o + 3 * kSideBlkO);
            if (MODE == 3) {
                const F4 cQ2 = lds4(o + 4 * kSideBlkO);
#pragma unroll
                for (int q = 0; q < 4; ++q) g.a[q] = fmaf(cM.a[q], cQ2.a[q], g.a[q]);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) g.a[q] = fmaf(v[4 + q], cQ.a[q], g.a[q]);
            st_masked<false>(p.G + sb + off, col, g, cmin, cmax, rowok);
        }

        __syncwarp();
        if (lane == 0 && b + DO < nblk) issue(z0 + DO * RB, so, s0);
        if (++so == DO) so = 0;
        if (++s0 == D) s0 = 0;
    }

    if (p.n_inj > 0) {
        __syncwarp();
        inject(p, s, um, zb, ze, max(cmin, x0), min(cmax, x0 + 32), lane, 32);
    }
    __syncwarp();
}

template <int H, int MODE>
__global__ void __launch_bounds__(warps_per_cta(MODE) * 32, ctas_per_sm(MODE)) acoustic_step_kernel(const __grid_constant__ StepParams p) {
    extern __shared__ __align__(128) float smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    // Programmatic dependent launch: this grid may become resident while the previous step is
    // still draining.  Let the next step do the same, set up the warp's barriers, then wait for
    // the previous step's stores before reading any wavefield.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    float *wsm = smem + (size_t)warp * warp_smem_floats(MODE);
    const unsigned bar0 = smem_u32(wsm);
    if (lane == 0) {
#pragma unroll
        for (int b = 0; b < 8; ++b) mbar_init(bar0 + 8u * b, 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    // everything that does not depend on the previous step is fetched before the wait: the warp's
    // unit, its first segment, the cell of its first sampling item
    // consecutive units share a CTA: PML units (listed first) end up grouped on the same SMs, which
    // measured faster than mixing them with interior units (their shallow rings lose the competition)
    const int unit = blockIdx.x * warps_per_cta(MODE) + warp;
    int g0 = 0, g1 = 0;
    if (unit < p.n_units) { g0 = __ldg(p.unit_start + unit); g1 = __ldg(p.unit_start + unit + 1); }
    int4 sa = make_int4(0, 0, 0, 0), sb4 = sa;
    if (g0 < g1) {
        sa = __ldg(reinterpret_cast<const int4 *>(p.segs + g0));
        sb4 = __ldg(reinterpret_cast<const int4 *>(p.segs + g0) + 1);
    }
    const int total_smp = p.S * p.n_smp;
    const int item0 = blockIdx.x * blockDim.x + tid;
    int cell0 = 0;
    if (item0 < total_smp) cell0 = __ldg(p.smp_cell + item0);
    __syncwarp();
    asm volatile("griddepcontrol.wait;" ::: "memory");

    // ---- sampling duty: record the current field at the sampling cells (before the step) ----
    for (int item = item0; item < total_smp; item += gridDim.x * blockDim.x) {
        const int s = item / p.n_smp;
        const int cell = item == item0 ? cell0 : __ldg(p.smp_cell + item);
        float val = p.u[(size_t)s * p.stride + cell];      // (plain load: written by the previous grid)
        if (p.smp_div_a) {
            const int z = cell / p.pitch, y = cell - z * p.pitch;
            const float hs = p.hdt * (__ldg(p.sigz + z) + __ldg(p.sigy + y));
            val = val * (1.0f + hs);
        }
        p.smp_out[item] = val;
    }
    if (g0 >= g1) return;

    unsigned long long t_start = 0;
    if (p.trace) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
    unsigned phase = 0;
    float *const wring = wsm + kBarFloats;
    float *const ring = wring + kWRing;
    for (int g = g0; g < g1; ++g) {
        if (g > g0) {
            sa = __ldg(reinterpret_cast<const int4 *>(p.segs + g));
            sb4 = __ldg(reinterpret_cast<const int4 *>(p.segs + g) + 1);
        }
        Seg sg;
        sg.s = sa.x; sg.x0 = sa.y; sg.zb = sa.z; sg.ze = sa.w;
        sg.type = sb4.x; sg.cmin = sb4.y; sg.cmax = sb4.z; sg.pad = 0;
        if (sg.type == kSegInt) march_ring<H, MODE, false, do_int(MODE)>(p, sg, ring, wring, bar0, phase);
        else if (sg.type == kSegBand) march_ring<H, MODE, true, kDOBand>(p, sg, ring, wring, bar0, phase);
        else march_side<H, MODE>(p, sg, ring, bar0, phase);
    }
    if (p.trace && lane == 0) {
        unsigned long long t_end;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        p.trace[2 * unit] = t_start;
        p.trace[2 * unit + 1] = t_end;
        p.trace[2 * kMaxUnits + unit] = smid;                       // the SM the unit ran on
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// Host side: layout, argument checks, the work table, the time loop.
// ------------------------------------------------------------------------------------------
namespace {

inline int round_up(int a, int b) { return (a + b - 1) / b * b; }
inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

int env_int(const char *name, int dflt) {
    const char *e = getenv(name);
    return e ? atoi(e) : dflt;
}
double env_dbl(const char *name, double dflt) {
    const char *e = getenv(name);
    return e ? atof(e) : dflt;
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
    FWI_CHECK_ARG(d->save_every >= 0, "save_every must be >= 0");
    return FWI_OK;
}

// A column of work: rows [z0, z1) of one strip of one shot.
struct Col {
    int s, type, x0, z0, z1, cmin, cmax;
};

// Geometry shared by get_layout and run.
struct Geom {
    int h, F, col0, pitch;
    int zs, ze, ys, ye, zlo, zhi, ylo, yhi, xb, xr;
    int tb_strips, side_l, side_r, int_strips;
    int max_rows;     // cap on the rows of one segment (chunk_rows; 0 = none)
};

// (kMaxUnits, the bound on resident warps of any device this could run on, is defined at the top of the file)

Geom make_geom(const fwi_acoustic_desc *d) {
    Geom g;
    g.h = d->order / 2;
    g.F = d->pml_width > 0 ? d->pml_width + g.h : 0;  // frame: sigma != 0 or PML memory within reach
    // first interior column on a 128-byte boundary; band strips start a whole number of 128-byte
    // lines to its left and at least 4 columns into the row
    const int lead = g.h + g.F;
    const int back = round_up(g.F, 32);
    int ylo = round_up(lead + 4, 32);
    while (ylo - back < 4) ylo += 32;
    g.col0 = ylo - lead;
    g.pitch = round_up(g.col0 + d->nyp + 8, 32);
    g.zs = g.h; g.ze = d->nzp - g.h;
    g.ys = g.col0 + g.h; g.ye = g.col0 + d->nyp - g.h;
    g.zlo = g.zs + g.F; g.zhi = g.ze - g.F;
    g.ylo = g.ys + g.F; g.yhi = g.ye - g.F;
    if (g.zhi < g.zlo) g.zlo = g.zhi = (g.zs + g.ze) / 2;      // model thinner than the frame: bands meet
    if (g.yhi < g.ylo) g.ylo = g.yhi = (g.ys + g.ye) / 2 / 4 * 4;
    g.xb = ylo - back;
    g.xr = g.yhi / 4 * 4;
    g.int_strips = ceil_div(g.yhi - g.ylo, 128);
    if (g.F > 0) {
        g.tb_strips = ceil_div(g.ye - g.xb, 128);
        g.side_l = ceil_div(g.ylo - g.xb, 32);
        g.side_r = ceil_div(g.ye - g.xr, 32);
    } else {
        g.tb_strips = g.side_l = g.side_r = 0;
    }
    g.max_rows = d->chunk_rows > 0 ? d->chunk_rows : env_int("FWI_B200_BZ", 0);
    return g;
}

// every column of work of one launch: side strips, then band strips, then interior strips
// Every column of work of one launch: side strips, then band strips, then interior strips.  With
// piece_rows > 0 the side and interior columns are cut into pieces of that many rows and listed
// row-chunk major (all strips of a shot at the same rows are neighbours in the list), so that the
// warps of a CTA, which take consecutive list entries, stream adjacent strips of the same rows:
// their 512-byte row reads then merge into multi-KB contiguous runs in DRAM pages and L2 lines.
void list_columns(const fwi_acoustic_desc *d, const Geom &g, std::vector<Col> &cols, int piece_rows = 0) {
    cols.clear();
    const int rows = g.zhi - g.zlo;
    const int pr = piece_rows > 0 ? piece_rows : std::max(rows, 1);
    const int n_pieces = rows > 0 ? ceil_div(rows, pr) : 0;
    if (g.F > 0)
        for (int s = 0; s < d->n_shots; ++s)
            for (int c = 0; c < n_pieces; ++c) {
                const int z0 = g.zlo + c * pr, z1 = std::min(g.zhi, z0 + pr);
                for (int k = 0; k < g.side_l; ++k) cols.push_back({s, kSegSide, g.xb + 32 * k, z0, z1, g.ys, g.ylo});
                for (int k = 0; k < g.side_r; ++k) cols.push_back({s, kSegSide, g.xr + 32 * k, z0, z1, g.yhi, g.ye});
            }
    if (g.F > 0)
        for (int s = 0; s < d->n_shots; ++s) {
            if (g.zlo > g.zs)
                for (int k = 0; k < g.tb_strips; ++k) cols.push_back({s, kSegBand, g.xb + 128 * k, g.zs, g.zlo, g.ys, g.ye});
            if (g.ze > g.zhi)
                for (int k = 0; k < g.tb_strips; ++k) cols.push_back({s, kSegBand, g.xb + 128 * k, g.zhi, g.ze, g.ys, g.ye});
        }
    for (int s = 0; s < d->n_shots; ++s)
        for (int c = 0; c < n_pieces; ++c) {
            const int z0 = g.zlo + c * pr, z1 = std::min(g.zhi, z0 + pr);
            for (int k = 0; k < g.int_strips; ++k) cols.push_back({s, kSegInt, g.ylo + 128 * k, z0, z1, g.ylo, g.yhi});
        }
}

// bound on the segments of any work table for these (uncut) columns
int64_t max_segments(const std::vector<Col> &cols, const Geom &g) {
    int64_t n = 0;
    for (const Col &c : cols) n += g.max_rows > 0 ? ceil_div(c.z1 - c.z0, g.max_rows) : 1;
    return 2 * n + 3 * (int64_t)kMaxUnits + 8;
}

// bytes of one work table: segments, then the unit offsets (16-byte aligned)
inline int64_t table_bytes(int64_t max_segs) {
    return max_segs * (int64_t)sizeof(Seg) + ((kMaxUnits + 1) * (int64_t)sizeof(int) + 15) / 16 * 16;
}

struct CostModel {
    double band, side, seg;
    int min_rows;
    double rate(int type) const { return type == kSegInt ? 1.0 : (type == kSegBand ? band : side); }
};

// Relative costs measured on B200 with the per-unit trace (FWI_B200_TRACE): a band row takes
// 3.0x (forward) / 3.4x (adjoint) the time of an interior row, a side-strip row about 1.0x.
CostModel cost_model(const Geom &g, int mode = 1) {
    CostModel c;
    // defaults per kernel mode (0 advance only, 1 + snapshot write, 2 imaging, 3 paired imaging); the environment
    // overrides are tuning aids: FWI_B200_COST_BAND[_M<mode>] etc.
    // (r02 sweep on C2, profiles/r02b_cost_sweep.txt: the advance-only and forward launches are bandwidth-bound on
    // their interior rows, so a latency-bound band row weighs 5 interior rows there)
    static const double band_d[4] = {5.0, 5.0, 3.4, 3.4}, side_d[4] = {1.0, 1.0, 1.0, 1.0}, seg_d[4] = {6.0, 6.0, 9.0, 9.0};
    auto env_mode = [&](const char *name, double dflt) {
        char buf[64];
        snprintf(buf, sizeof(buf), "%s_M%d", name, mode);
        return env_dbl(buf, env_dbl(name, dflt));
    };
    const int m = mode < 0 ? 0 : (mode > 3 ? 3 : mode);
    c.band = env_mode("FWI_B200_COST_BAND", band_d[m]);   // cost of a band row relative to an interior row
    c.side = env_mode("FWI_B200_COST_SIDE", side_d[m]);   // cost of one row of a side segment
    c.seg = env_mode("FWI_B200_COST_SEG", seg_d[m]);      // prologue of a segment, in interior rows
    c.min_rows = env_int("FWI_B200_MINROWS", 4);
    if (g.max_rows > 0) c.min_rows = std::min(c.min_rows, g.max_rows);
    if (c.min_rows < 1) c.min_rows = 1;
    return c;
}

// Walks the columns in order, filling units up to capacity T; returns the number of units used.
int fill_units(const std::vector<Col> &cols, const Geom &g, const CostModel &cm, double T, std::vector<Seg> *segs,
               std::vector<int> *unit_start) {
    if (segs) segs->clear();
    if (unit_start) unit_start->assign(1, 0);
    double cur = 0;
    int n_segs = 0, units = 0;
    for (const Col &c : cols) {
        const double r = cm.rate(c.type);
        int z = c.z0;
        while (z < c.z1) {
            const int left = c.z1 - z;
            int n = (int)((T - cur - cm.seg) / r + 1e-9);
            if (n < std::min(cm.min_rows, left)) {
                if (cur > 0) {      // close the unit and retry with an empty one
                    ++units;
                    if (unit_start) unit_start->push_back(n_segs);
                    cur = 0;
                    continue;
                }
                n = std::min(cm.min_rows, left);
            }
            n = std::min(n, left);
            if (g.max_rows > 0) n = std::min(n, g.max_rows);
            if (segs) segs->push_back({c.s, c.x0, z, z + n, c.type, c.cmin, c.cmax, 0});
            ++n_segs;
            cur += n * r + cm.seg;
            z += n;
        }
    }
    if (cur > 0) {
        ++units;
        if (unit_start) unit_start->push_back(n_segs);
    }
    return units;
}

// Cuts the work of one launch into segments and deals them to at most `n_units` warp units: the
// smallest unit capacity for which a contiguous fill of the (row-chunk major) column list needs no
// more than n_units units (bisection).
void build_units(const fwi_acoustic_desc *d, const Geom &g, int n_units, std::vector<Col> &cols,
                 std::vector<Seg> &segs, std::vector<int> &unit_start, int mode = 1) {
    CostModel cm = cost_model(g, mode);
    list_columns(d, g, cols);
    double total = 0;
    for (const Col &c : cols) total += (c.z1 - c.z0) * cm.rate(c.type) + cm.seg;
    // small problems (less than a few rows of work per resident warp, e.g. the reference's own
    // 100x300 / 151x200 grids): a launch is latency-bound, so cut the work into the shortest
    // segments the kernel supports and spread them over as many warps as there are segments
    // (measured on the 100x300 x 10-shot case: 6.5 / 8.4 us per forward / adjoint launch with
    // units of >= 6 row-costs and 2-row segments, 11.3 / 11.0 us with 24 and 4)
    static const double min_unit = env_dbl("FWI_B200_MINUNIT", 6.0);
    if (total / n_units < 24.0 && g.max_rows == 0 && env_int("FWI_B200_MINROWS", 0) == 0) cm.min_rows = 2;
    if (total / n_units < min_unit) n_units = std::max(1, (int)(total / min_unit));
    if (env_int("FWI_B200_PIECES", 0)) {
        // pieces of about one unit's worth of interior rows
        const int piece = std::max(cm.min_rows, (int)(total / n_units - cm.seg + 0.5));
        list_columns(d, g, cols, piece);
        total = 0;
        for (const Col &c : cols) total += (c.z1 - c.z0) * cm.rate(c.type) + cm.seg;
    }
    double lo = total / n_units, hi = 2.0 * lo + cm.seg + cm.min_rows * std::max(cm.band, std::max(cm.side, 1.0)) + 64.0;
    while (fill_units(cols, g, cm, hi, nullptr, nullptr) > n_units) hi *= 2.0;
    for (int it = 0; it < 24; ++it) {
        const double mid = 0.5 * (lo + hi);
        if (fill_units(cols, g, cm, mid, nullptr, nullptr) <= n_units) hi = mid;
        else lo = mid;
    }
    fill_units(cols, g, cm, hi, &segs, &unit_start);
    if (unit_start.size() < 2) unit_start.push_back((int)segs.size());
}

// ---- TMA descriptors ----------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}

// 2-D fp32 tensor [rows][pitch] at `base`, boxes of kRB rows x `box_cols` columns, zero fill outside
int make_map(CUtensorMap *tm, const float *base, int64_t rows, int pitch, int box_cols) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) {
        fwi::set_error("cuTensorMapEncodeTiled is not available from this driver");
        return FWI_ERR_CUDA;
    }
    const cuuint64_t gdim[2] = {(cuuint64_t)pitch, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)pitch * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)kRB};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        fwi::set_error("cuTensorMapEncodeTiled failed (%d) for a [%lld][%d] tensor", (int)r, (long long)rows, pitch);
        return FWI_ERR_CUDA;
    }
    return FWI_OK;
}

template <int H, int MODE>
struct Launcher {
    static constexpr int kNW = warps_per_cta(MODE);
    static constexpr size_t smem_bytes() { return (size_t)kNW * warp_smem_floats(MODE) * sizeof(float); }
    // resident CTAs per SM of this kernel on the current device (cached per device)
    static int resident_ctas(int *sms) {
        static std::mutex mu;                       // (fwi_acoustic_run_multi calls in from several host threads)
        std::lock_guard<std::mutex> lock(mu);
        static int cached_dev = -1, cached_occ = 0, cached_sms = 0;
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev != cached_dev) {
            auto kern = acoustic_step_kernel<H, MODE>;
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes());
            int occ = 0;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kNW * 32, smem_bytes());
            cudaDeviceGetAttribute(&cached_sms, cudaDevAttrMultiProcessorCount, dev);
            cached_occ = occ;
            cached_dev = dev;
        }
        *sms = cached_sms;
        return cached_occ;
    }
    static cudaError_t launch(const StepParams &p, int blocks, cudaStream_t st, bool allow_pdl) {
        static const int pdl_env = env_int("FWI_B200_PDL", 1);
        const int pdl = pdl_env && allow_pdl;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(blocks);
        cfg.blockDim = dim3(kNW * 32);
        cfg.dynamicSmemBytes = smem_bytes();
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = pdl ? 1 : 0;
        return cudaLaunchKernelEx(&cfg, acoustic_step_kernel<H, MODE>, p);
    }
};

template <int MODE>
int resident_h(int h, int *sms) {
    switch (h) {
    case 1: return Launcher<1, MODE>::resident_ctas(sms);
    case 2: return Launcher<2, MODE>::resident_ctas(sms);
    case 3: return Launcher<3, MODE>::resident_ctas(sms);
    default: return Launcher<4, MODE>::resident_ctas(sms);
    }
}

template <int MODE>
cudaError_t launch_h(int h, const StepParams &p, int blocks, cudaStream_t st, bool pdl) {
    switch (h) {
    case 1: return Launcher<1, MODE>::launch(p, blocks, st, pdl);
    case 2: return Launcher<2, MODE>::launch(p, blocks, st, pdl);
    case 3: return Launcher<3, MODE>::launch(p, blocks, st, pdl);
    default: return Launcher<4, MODE>::launch(p, blocks, st, pdl);
    }
}

// ---- x-slab ghost exchange (include/fwi_b200.h, fwi_slab_link) ----------------------------------
struct SlabXfer {
    const float *src[4];      // push: the live fields of this rank (u, u-, pz, py)
    float *dst[4];            // unpack: the same fields
    int n_live, S, nzp, pitch, G;
    long long stride;         // floats per shot per field
    int col[2];               // push: first sent column per side; unpack: first ghost column per side
    int has[2];
    float *stage[2];          // push: the neighbour's staging slot per side; unpack: this rank's slot per side
    unsigned *flag[2];        // push: the neighbour's flag to raise per side; unpack: this rank's flag per side
    unsigned *counter;        // push: this rank's block counters [2]
    unsigned *status;
    unsigned epoch;
    unsigned long long timeout_ns;
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// blockIdx.y = side.  Packs [live][S][nzp][G] from the owned edge columns into the neighbour's
// staging slot with peer stores; the last block to finish raises the neighbour's flag.
__global__ void __launch_bounds__(256) slab_push_kernel(const SlabXfer x) {
    const int side = blockIdx.y;
    if (!x.has[side]) return;
    const long long per_field = (long long)x.S * x.nzp * x.G;
    const long long total = per_field * x.n_live;
    float *out = x.stage[side];
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const int f = (int)(e / per_field);
        long long r = e - f * per_field;
        const int c = (int)(r % x.G);
        r /= x.G;
        const int z = (int)(r % x.nzp), s = (int)(r / x.nzp);
        out[e] = x.src[f][s * x.stride + (long long)z * x.pitch + x.col[side] + c];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned done = atomicAdd(&x.counter[side], 1u);
        if (done == gridDim.x - 1) {
            x.counter[side] = 0;
            __threadfence_system();
            st_release_sys(x.flag[side], x.epoch);
        }
    }
}

// blockIdx.y = side.  Waits until the neighbour has delivered exchange `epoch`, then copies the
// staging slot into the ghost columns.  A wait that exceeds the timeout sets status and gives up
// (the data is then wrong, and the host reports it) instead of hanging the device.
__global__ void __launch_bounds__(256) slab_unpack_kernel(const SlabXfer x) {
    const int side = blockIdx.y;
    if (!x.has[side]) return;
    if (threadIdx.x == 0) {
        const unsigned long long t0 = global_ns();
        while ((int)(ld_acquire_sys(x.flag[side]) - x.epoch) < 0) {
            if (*(volatile unsigned *)x.status != 0) break;
            if (global_ns() - t0 > x.timeout_ns) { atomicExch(x.status, 1u); break; }
            __nanosleep(200);
        }
    }
    __syncthreads();
    const long long per_field = (long long)x.S * x.nzp * x.G;
    const long long total = per_field * x.n_live;
    const float *in = x.stage[side];
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const int f = (int)(e / per_field);
        long long r = e - f * per_field;
        const int c = (int)(r % x.G);
        r /= x.G;
        const int z = (int)(r % x.nzp), s = (int)(r / x.nzp);
        x.dst[f][s * x.stride + (long long)z * x.pitch + x.col[side] + c] = __ldcv(in + e);
    }
}

}  // namespace

extern "C" int fwi_acoustic_get_layout(const fwi_acoustic_desc *desc, fwi_acoustic_layout *out) {
    int rc = check_desc(desc);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(out != nullptr, "layout output is NULL");
    const Geom g = make_geom(desc);
    std::vector<Col> cols;
    list_columns(desc, g, cols);
    out->pitch = g.pitch;
    out->col0 = g.col0;
    out->shot_stride = (int64_t)desc->nzp * g.pitch;
    out->n_images = desc->n_shots;
    out->wave_fields = desc->pml_width > 0 ? 6 : 2;
    out->chunk_rows = g.max_rows;
    out->units_per_shot = (int32_t)(cols.size() / desc->n_shots);
    // up to three work tables per call (an adjoint loop alternates advance-only, paired-imaging and single-imaging launches)
    out->unit_table_bytes = 3 * table_bytes(max_segments(cols, g));
    return FWI_OK;
}

extern "C" int fwi_acoustic_plan_stats(const fwi_acoustic_desc *desc, int n_units, int64_t *n_segments,
                                       int32_t *n_units_used, double *max_cost, double *mean_cost) {
    int rc = check_desc(desc);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(n_units >= 1 && n_units <= kMaxUnits, "n_units out of range");
    const Geom g = make_geom(desc);
    std::vector<Col> cols;
    std::vector<Seg> segs;
    std::vector<int> unit_start;
    build_units(desc, g, n_units, cols, segs, unit_start);
    // every computational cell must be covered exactly once: check by area
    int64_t area = 0;
    for (const Seg &sg : segs) {
        const int w = sg.type == kSegSide ? 32 : 128;
        const int c0 = std::max(sg.cmin, sg.x0), c1 = std::min(sg.cmax, sg.x0 + w);
        if (c1 > c0) area += (int64_t)(c1 - c0) * (sg.ze - sg.zb);
    }
    const int64_t want = (int64_t)desc->n_shots * (g.ze - g.zs) * (g.ye - g.ys);
    FWI_CHECK_ARG(area == want, "internal: segments cover %lld cells, computational region has %lld",
                  (long long)area, (long long)want);
    const CostModel cm = cost_model(g);
    double mx = 0, sum = 0;
    const int nu = (int)unit_start.size() - 1;
    for (int u = 0; u < nu; ++u) {
        double c = 0;
        for (int k = unit_start[u]; k < unit_start[u + 1]; ++k)
            c += (segs[k].ze - segs[k].zb) * cm.rate(segs[k].type) + cm.seg;
        mx = std::max(mx, c);
        sum += c;
    }
    if (n_segments) *n_segments = (int64_t)segs.size();
    if (n_units_used) *n_units_used = nu;
    if (max_cost) *max_cost = mx;
    if (mean_cost) *mean_cost = nu ? sum / nu : 0;
    return FWI_OK;
}

static int run_impl(const fwi_acoustic_desc *d, const fwi_acoustic_buffers *b, const fwi_slab_link *link, void *stream) {
    int rc = check_desc(d);
    if (rc != FWI_OK) return rc;
    FWI_CHECK_ARG(b != nullptr, "buffers struct is NULL");
    FWI_CHECK_ARG(b->w && b->wave, "w and wave must not be NULL");
    FWI_CHECK_ARG(b->unit_table, "unit_table workspace must not be NULL");
    FWI_CHECK_ARG(d->pml_width == 0 || (b->sigz && b->sigy), "sigz/sigy required when pml_width > 0");
    FWI_CHECK_ARG(!d->smp_div_a || (b->sigz && b->sigy), "sigz/sigy required when smp_div_a is set");
    FWI_CHECK_ARG(!d->inj_scale_a || (b->sigz && b->sigy), "sigz/sigy required when inj_scale_a is set");
    FWI_CHECK_ARG(d->n_smp == 0 || (b->smp_cell && b->smp_out), "smp_cell/smp_out required when n_smp > 0");
    FWI_CHECK_ARG(d->n_inj == 0 || (b->inj_row_start && b->inj_cell && b->inj_col && b->inj_amp),
                  "injection lists required when n_inj > 0");
    FWI_CHECK_ARG(!(b->q_write && b->q_read), "q_write and q_read are mutually exclusive");
    FWI_CHECK_ARG(b->q_step_stride == 0 || ((b->q_write || b->q_read) && b->q_step_stride % make_geom(d).pitch == 0 &&
                                            b->q_step_stride >= (int64_t)d->n_shots * d->nzp * make_geom(d).pitch),
                  "q_step_stride must be 0 or a whole number of rows, at least one snapshot of this call's shots, with q_write / q_read");
    FWI_CHECK_ARG(!b->q_read || b->image, "image required with q_read");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (d->nsteps == 0) return FWI_OK;

    const Geom g = make_geom(d);
    const int h = g.h;
    const int mode = b->q_read ? 2 : (b->q_write ? 1 : 0);

    // ---- small grids: the whole time loop as one cluster-resident kernel (acoustic_small.cu) ----
    const int small_on = env_int("FWI_B200_SMALL", 1);      // read per call: tests switch paths in-process
    static const int small_cells = env_int("FWI_B200_SMALL_CELLS", 1500000);
    if (small_on && !link && !getenv("FWI_B200_TRACE") && d->chunk_rows == 0 && (mode != 2 || d->adjoint) && b->q_step_stride == 0 && d->save_every <= 1 &&
        (int64_t)d->n_shots * d->nzp * d->nyp <= small_cells) {
        fwi::SmallParams sp = {};
        sp.nzp = d->nzp; sp.nyp = d->nyp; sp.pitch = g.pitch; sp.col0 = g.col0; sp.h = h; sp.pw = d->pml_width;
        sp.S = d->n_shots; sp.nsteps = d->nsteps; sp.adjoint = d->adjoint; sp.mode = mode;
        sp.n_inj = d->n_inj; sp.inj_scale_a = d->inj_scale_a; sp.n_smp = d->n_smp; sp.smp_div_a = d->smp_div_a;
        sp.wave_pml = d->pml_width > 0;
        sp.stride = (long long)d->nzp * g.pitch;
        const double dz = d->dz, dy = d->dy;
        sp.c0 = (float)(fwi::kD2[h][0] / (dz * dz) + fwi::kD2[h][0] / (dy * dy));
        for (int k = 1; k <= h; ++k) {
            sp.cz2[k] = (float)(fwi::kD2[h][k] / (dz * dz));
            sp.cy2[k] = (float)(fwi::kD2[h][k] / (dy * dy));
            sp.cz1[k] = (float)(fwi::kD1[h][k - 1] / dz);
            sp.cy1[k] = (float)(fwi::kD1[h][k - 1] / dy);
        }
        sp.dt = (float)d->dt; sp.dt2 = sp.dt * sp.dt; sp.hdt = 0.5f * sp.dt;
        sp.w = b->w; sp.sigz = b->sigz; sp.sigy = b->sigy; sp.wave = b->wave;
        sp.inj_row_start = b->inj_row_start; sp.inj_cell = b->inj_cell; sp.inj_col = b->inj_col; sp.inj_amp = b->inj_amp;
        sp.smp_cell = b->smp_cell; sp.smp_out = b->smp_out;
        sp.q_write = b->q_write; sp.q_read = b->q_read; sp.image = b->image;
        bool launched = false;
        if ((rc = fwi::small_run(sp, stream, &launched)) != FWI_OK) return rc;
        if (launched) return FWI_OK;
    }

    // ---- work tables: one cost-balanced unit per resident warp, one table per kernel mode the loop launches ----
    // Adjoint with imaging: G += field * q costs an 8-byte read-modify-write of the image per cell per step.  The field
    // of the previous step is still on hand as this step's u-, so two steps are imaged by ONE pass over G: iterations
    // alternate between an advance-only launch (MODE 0, 16 B per cell) and a launch that images both steps (MODE 3),
    // the additions to G staying in the order one imaging per step makes them - results are bit-identical.  An odd
    // last iteration images alone (MODE 2).  FWI_B200_PAIR=0 keeps one imaging per step.
    const int kimg = d->save_every > 1 ? d->save_every : 1;          // snapshot decimation: image / snapshot every kimg-th step
    const bool pair = mode == 2 && kimg == 1 && env_int("FWI_B200_PAIR", 1) != 0 && d->nsteps >= 2;
    int sms = 0;
    struct Table { int mode, n_units, blocks; Seg *d_segs; int *d_units; };
    Table tables[3];
    int n_tables = 0;
    const int want[3] = {pair ? 0 : mode, pair ? 3 : (kimg > 1 && mode != 0 ? 0 : -1), pair && (d->nsteps & 1) ? 2 : -1};
    thread_local std::vector<Col> cols;
    thread_local std::vector<Seg> segs;
    thread_local std::vector<int> unit_start;
    list_columns(d, g, cols);
    const int64_t max_segs = max_segments(cols, g);
    for (int t = 0; t < 3; ++t) {
        const int m = want[t];
        if (m < 0) continue;
        const int occ = m == 3 ? resident_h<3>(h, &sms) : (m == 2 ? resident_h<2>(h, &sms) : (m == 1 ? resident_h<1>(h, &sms) : resident_h<0>(h, &sms)));
        if (occ <= 0 || sms <= 0) {
            fwi::set_error("acoustic step kernel (mode %d) cannot be resident on this device (occupancy %d): %s", m, occ,
                           cudaGetErrorString(cudaGetLastError()));
            return FWI_ERR_CUDA;
        }
        const int nw = warps_per_cta(m);
        int n_units = std::min(kMaxUnits, sms * occ * nw);
        const int units_env = env_int("FWI_B200_UNITS", 0);
        if (units_env > 0) n_units = std::min(kMaxUnits, units_env);
        build_units(d, g, n_units, cols, segs, unit_start, m);
        list_columns(d, g, cols);
        n_units = (int)unit_start.size() - 1;
        FWI_CHECK_ARG((int64_t)segs.size() <= max_segs, "internal: work table overflow");
        Table &T = tables[n_tables];
        T.mode = m; T.n_units = n_units; T.blocks = ceil_div(n_units, nw);
        char *base = static_cast<char *>(b->unit_table) + (size_t)n_tables * table_bytes(max_segs);
        T.d_segs = reinterpret_cast<Seg *>(base);
        T.d_units = reinterpret_cast<int *>(base + max_segs * sizeof(Seg));
        FWI_CHECK_CUDA(cudaMemcpyAsync(T.d_segs, segs.data(), segs.size() * sizeof(Seg), cudaMemcpyHostToDevice, st));
        FWI_CHECK_CUDA(cudaMemcpyAsync(T.d_units, unit_start.data(), unit_start.size() * sizeof(int),
                                       cudaMemcpyHostToDevice, st));
        // (pageable source: cudaMemcpyAsync returns once the host data has been staged, so the vectors
        // may be reused for the next table and by the next call)
        ++n_tables;
    }
    auto table_of = [&](int m) -> const Table & {
        for (int t = 0; t < n_tables; ++t)
            if (tables[t].mode == m) return tables[t];
        return tables[0];
    };

    StepParams p;
    p.nzp = d->nzp; p.pitch = g.pitch;
    p.zs = g.zs; p.ze = g.ze; p.ys = g.ys; p.ye = g.ye;
    p.stride = (long long)d->nzp * g.pitch;
    p.S = d->n_shots;
    p.segs = tables[0].d_segs; p.unit_start = tables[0].d_units; p.n_units = tables[0].n_units;
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
    const int fld_rows = d->n_shots * d->nzp;
    const int n_fields = d->pml_width > 0 ? 6 : 2;
    // rows between the snapshots of consecutive steps (a call on a subset of the shots of a snapshot array: q_step_stride)
    const int64_t q_rows = b->q_step_stride > 0 ? b->q_step_stride / g.pitch : fld_rows;
    const int64_t q_map_rows = (int64_t)((d->nsteps - 1) / kimg) * q_rows + fld_rows;
    FWI_CHECK_ARG(q_map_rows < (int64_t)1 << 31, "snapshot tensor has too many rows for one TMA descriptor");
    if ((rc = make_map(&p.tm_wave_u, b->wave, (int64_t)n_fields * fld_rows, g.pitch, kRowU)) != FWI_OK) return rc;
    if ((rc = make_map(&p.tm_wave_o, b->wave, (int64_t)n_fields * fld_rows, g.pitch, kRowO)) != FWI_OK) return rc;
    if ((rc = make_map(&p.tm_w, b->w, d->nzp, g.pitch, kRowO)) != FWI_OK) return rc;
    if (mode >= 2) {
        if ((rc = make_map(&p.tm_q, b->q_read, q_map_rows, g.pitch, kRowO)) != FWI_OK) return rc;
        if ((rc = make_map(&p.tm_g, b->image, fld_rows, g.pitch, kRowO)) != FWI_OK) return rc;
    } else {
        p.tm_q = p.tm_w;
        p.tm_g = p.tm_w;
    }
    if (d->pml_width > 0) {
        if ((rc = make_map(&p.ts_wave_u, b->wave, (int64_t)n_fields * fld_rows, g.pitch, kSideU)) != FWI_OK) return rc;
        if ((rc = make_map(&p.ts_wave_o, b->wave, (int64_t)n_fields * fld_rows, g.pitch, kSideO)) != FWI_OK) return rc;
        if ((rc = make_map(&p.ts_w, b->w, d->nzp, g.pitch, kSideO)) != FWI_OK) return rc;
        if (mode >= 2) {
            if ((rc = make_map(&p.ts_q, b->q_read, q_map_rows, g.pitch, kSideO)) != FWI_OK) return rc;
            if ((rc = make_map(&p.ts_g, b->image, fld_rows, g.pitch, kSideO)) != FWI_OK) return rc;
        } else {
            p.ts_q = p.ts_w;
            p.ts_g = p.ts_w;
        }
    } else {
        p.ts_wave_u = p.ts_wave_o = p.ts_w = p.ts_q = p.ts_g = p.tm_w;
    }
    float *u = b->wave, *um = b->wave + fld;
    float *pz = nullptr, *pzn = nullptr, *py = nullptr, *pyn = nullptr;
    if (d->pml_width > 0) { pz = b->wave + 2 * fld; pzn = b->wave + 3 * fld; py = b->wave + 4 * fld; pyn = b->wave + 5 * fld; }

    // dev aid: per-unit start / end times of the last launch, dumped to the file named by FWI_B200_TRACE
    static const char *trace_path = getenv("FWI_B200_TRACE");
    p.trace = nullptr;
    if (trace_path) {
        static unsigned long long *d_trace = nullptr;
        if (!d_trace) cudaMalloc(&d_trace, 3 * sizeof(unsigned long long) * kMaxUnits);
        p.trace = d_trace;
    }
    SlabXfer push = {}, unpack = {};
    unsigned exchanges = 0;
    if (link) {
        char *me = static_cast<char *>(link->comm);
        char *peer[2] = {static_cast<char *>(link->peer_left), static_cast<char *>(link->peer_right)};
        push.n_live = unpack.n_live = d->pml_width > 0 ? 4 : 2;
        push.S = unpack.S = d->n_shots; push.nzp = unpack.nzp = d->nzp; push.pitch = unpack.pitch = g.pitch;
        push.G = unpack.G = link->ghost_cols; push.stride = unpack.stride = p.stride;
        push.has[0] = unpack.has[0] = link->has_left; push.has[1] = unpack.has[1] = link->has_right;
        push.col[0] = link->send_left_col; push.col[1] = link->send_right_col;
        unpack.col[0] = link->ghost_left_col; unpack.col[1] = link->ghost_right_col;
        unsigned *hdr = reinterpret_cast<unsigned *>(me);
        push.counter = hdr + 2;
        push.status = unpack.status = hdr + 4;
        unpack.flag[0] = hdr + 0; unpack.flag[1] = hdr + 1;
        // what I send to my left neighbour arrives "from its right" (its side 1), and vice versa
        push.flag[0] = peer[0] ? reinterpret_cast<unsigned *>(peer[0]) + 1 : nullptr;
        push.flag[1] = peer[1] ? reinterpret_cast<unsigned *>(peer[1]) + 0 : nullptr;
        push.timeout_ns = unpack.timeout_ns = (unsigned long long)(link->timeout_ms ? link->timeout_ms : 20000u) * 1000000ull;
    }
    for (int it = 0; it < d->nsteps; ++it) {
        const int n = d->adjoint ? d->nsteps - 1 - it : it;
        p.u = u; p.um = um; p.pz = pz; p.pzn = pzn; p.py = py; p.pyn = pyn;
        p.row_u = (int)((u - b->wave) / g.pitch);
        p.row_um = (int)((um - b->wave) / g.pitch);
        p.row_pz = pz ? (int)((pz - b->wave) / g.pitch) : 0;
        p.row_py = py ? (int)((py - b->wave) / g.pitch) : 0;
        const int nq = n / kimg;                          // snapshot index of this step
        p.row_q = (int)(nq * q_rows);
        p.row_q2 = (int)((nq + 1) * q_rows);
        // which launch this iteration is: see the comment on `pair` above; with decimation only every kimg-th step
        // writes / images a snapshot, the others advance only
        int m = mode;
        if (pair) m = (it & 1) ? 3 : (it + 1 < d->nsteps ? 0 : 2);
        else if (kimg > 1 && n % kimg != 0) m = 0;
        const Table &T = table_of(m);
        p.segs = T.d_segs; p.unit_start = T.d_units; p.n_units = T.n_units;
        const int blocks = T.blocks;
        p.qw = b->q_write ? b->q_write + (size_t)nq * (b->q_step_stride > 0 ? (size_t)b->q_step_stride : fld) : nullptr;
        p.qr = b->q_read ? b->q_read + (size_t)nq * (b->q_step_stride > 0 ? (size_t)b->q_step_stride : fld) : nullptr;
        p.G = b->image;
        p.inj_amp = d->n_inj ? b->inj_amp + (size_t)n * d->n_shots * d->n_inj : nullptr;
        p.smp_out = d->n_smp ? b->smp_out + (size_t)n * d->n_shots * d->n_smp : nullptr;
        cudaError_t e;
        const bool pdl = !(d->flags & FWI_FLAG_NO_PDL);
        if (m == 3) e = launch_h<3>(h, p, blocks, st, pdl);
        else if (m == 2) e = launch_h<2>(h, p, blocks, st, pdl);
        else if (m == 1) e = launch_h<1>(h, p, blocks, st, pdl);
        else e = launch_h<0>(h, p, blocks, st, pdl);
        if (e != cudaSuccess) {
            fwi::set_error("acoustic step launch failed at step %d: %s", n, cudaGetErrorString(e));
            return FWI_ERR_CUDA;
        }
        fwi::count_launch();
        float *t = u; u = um; um = t;
        t = pz; pz = pzn; pzn = t;
        t = py; py = pyn; pyn = t;
        if (link && it + 1 < d->nsteps && (it + 1) % link->halo_steps == 0) {
            // ghost exchange number epoch + k of the live fields, which are now u, um, pz, py
            const unsigned epoch = link->epoch + ++exchanges;
            const size_t slot = (size_t)push.n_live * d->n_shots * d->nzp * link->ghost_cols;
            char *me = static_cast<char *>(link->comm);
            char *peer[2] = {static_cast<char *>(link->peer_left), static_cast<char *>(link->peer_right)};
            const float *live[4] = {u, um, pz, py};
            for (int f = 0; f < 4; ++f) { push.src[f] = live[f]; unpack.dst[f] = const_cast<float *>(live[f]); }
            push.epoch = unpack.epoch = epoch;
            for (int side = 0; side < 2; ++side) {
                // stage[parity][side]: I write the neighbour's slot of the OPPOSITE side, and read my own
                float *mine = reinterpret_cast<float *>(me + FWI_SLAB_HEADER_BYTES) + ((size_t)(epoch & 1) * 2 + side) * slot;
                unpack.stage[side] = mine;
                push.stage[side] = peer[side] ? reinterpret_cast<float *>(peer[side] + FWI_SLAB_HEADER_BYTES)
                                                     + ((size_t)(epoch & 1) * 2 + (1 - side)) * slot
                                              : nullptr;
            }
            const int xb = (int)std::min<size_t>(64, (slot + 255) / 256);
            slab_push_kernel<<<dim3(xb, 2), 256, 0, st>>>(push);
            slab_unpack_kernel<<<dim3(xb, 2), 256, 0, st>>>(unpack);
            cudaError_t e2 = cudaGetLastError();
            if (e2 != cudaSuccess) {
                fwi::set_error("slab exchange launch failed at step %d: %s", n, cudaGetErrorString(e2));
                return FWI_ERR_CUDA;
            }
            fwi::count_launch(2);
        }
    }
    if (p.trace) {
        // (dev aid: the trace is of the last launch; segs / unit_start are those of the last table built - run with
        // FWI_B200_PAIR=0 or an even step count so that the two coincide)
        const int n_units = (int)unit_start.size() - 1;
        std::vector<unsigned long long> tr(2 * (size_t)n_units);
        cudaStreamSynchronize(st);
        cudaMemcpy(tr.data(), p.trace, tr.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
        static int trace_call = 0;
        char path[512];
        snprintf(path, sizeof(path), "%s.%d", trace_path, trace_call++);     // one file per call
        if (FILE *f = fopen(path, "w")) {
            unsigned long long t0 = ~0ull;
            std::vector<int> smid(n_units);
            {
                std::vector<unsigned long long> sm(n_units);
                cudaMemcpy(sm.data(), p.trace + 2 * kMaxUnits, sm.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
                for (int k = 0; k < n_units; ++k) smid[k] = (int)sm[k];
            }
            for (int k = 0; k < n_units; ++k) t0 = std::min(t0, tr[2 * k]);
            const CostModel cm = cost_model(g, mode);
            for (int k = 0; k < n_units; ++k) {
                int ni = 0, nb = 0, ns = 0;
                double cost = 0;
                for (int q = unit_start[k]; q < unit_start[k + 1]; ++q) {
                    const int rows = segs[q].ze - segs[q].zb;
                    (segs[q].type == kSegInt ? ni : segs[q].type == kSegBand ? nb : ns) += rows;
                    cost += rows * cm.rate(segs[q].type) + cm.seg;
                }
                fprintf(f, "%d %d %llu %llu %d %d %d %d %.1f %d\n", mode, k, tr[2 * k] - t0, tr[2 * k + 1] - t0,
                        unit_start[k + 1] - unit_start[k], ni, nb, ns, cost, smid[k]);
            }
            fclose(f);
        }
    }
    return FWI_OK;
}

extern "C" int fwi_acoustic_run(const fwi_acoustic_desc *d, const fwi_acoustic_buffers *b, void *stream) {
    return run_impl(d, b, nullptr, stream);
}

extern "C" int fwi_acoustic_run_multi(int32_t n_jobs, const fwi_acoustic_desc *const *descs,
                                      const fwi_acoustic_buffers *const *bufs, void *const *streams) {
    FWI_CHECK_ARG(n_jobs >= 1 && n_jobs <= 16 && descs && bufs && streams, "bad job list");
    if (n_jobs == 1) return fwi_acoustic_run(descs[0], bufs[0], streams[0]);
    // one host thread per job: their launches interleave on the device, each job on its own stream
    int dev = 0;
    FWI_CHECK_CUDA(cudaGetDevice(&dev));
    std::vector<int> rc(n_jobs, FWI_OK);
    std::vector<std::string> err(n_jobs);
    std::vector<std::thread> th;
    for (int j = 0; j < n_jobs; ++j)
        th.emplace_back([&, j]() {
            cudaSetDevice(dev);
            rc[j] = fwi_acoustic_run(descs[j], bufs[j], streams[j]);
            if (rc[j] != FWI_OK) err[j] = fwi_last_error();
        });
    for (auto &t : th) t.join();
    for (int j = 0; j < n_jobs; ++j)
        if (rc[j] != FWI_OK) {
            fwi::set_error("job %d: %s", j, err[j].c_str());
            return rc[j];
        }
    return FWI_OK;
}

extern "C" int32_t fwi_slab_exchanges(int32_t nsteps, int32_t halo_steps) {
    if (nsteps <= 0 || halo_steps <= 0) return 0;
    return (nsteps - 1) / halo_steps;
}

extern "C" int fwi_acoustic_run_slab(const fwi_acoustic_desc *d, const fwi_acoustic_buffers *b,
                                     const fwi_slab_link *link, void *stream) {
    FWI_CHECK_ARG(d && b && link, "NULL argument");
    FWI_CHECK_ARG(link->halo_steps >= 1 && link->ghost_cols == link->halo_steps * (d->order / 2),
                  "ghost_cols must equal halo_steps * order/2");
    FWI_CHECK_ARG(link->comm != nullptr, "comm block is NULL");
    FWI_CHECK_ARG(!link->has_left || link->peer_left, "peer_left is NULL");
    FWI_CHECK_ARG(!link->has_right || link->peer_right, "peer_right is NULL");
    return run_impl(d, b, link, stream);
}

extern "C" int fwi_slab_status(const void *comm, void *stream, uint32_t *status) {
    FWI_CHECK_ARG(comm && status, "NULL argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    FWI_CHECK_CUDA(cudaMemcpyAsync(status, static_cast<const char *>(comm) + 16, 4, cudaMemcpyDeviceToHost, st));
    FWI_CHECK_CUDA(cudaStreamSynchronize(st));
    return FWI_OK;
}

extern "C" int fwi_peer_alloc(int64_t bytes, void **ptr, unsigned char handle[64]) {
    FWI_CHECK_ARG(bytes > 0 && ptr && handle, "bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    FWI_CHECK_CUDA(cudaMalloc(ptr, (size_t)bytes));
    FWI_CHECK_CUDA(cudaMemset(*ptr, 0, (size_t)bytes));
    cudaIpcMemHandle_t h;
    FWI_CHECK_CUDA(cudaIpcGetMemHandle(&h, *ptr));
    memcpy(handle, &h, 64);
    return FWI_OK;
}

extern "C" int fwi_peer_open(const unsigned char handle[64], void **ptr) {
    FWI_CHECK_ARG(handle && ptr, "bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    FWI_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return FWI_OK;
}

extern "C" int fwi_peer_close(void *ptr) {
    FWI_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
    return FWI_OK;
}

extern "C" int fwi_peer_free(void *ptr) {
    FWI_CHECK_CUDA(cudaFree(ptr));
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
