/* CPU oracle (2/2): plain C + OpenMP restatement of the 2-D acoustic propagator,
 * forward and hand-derived exact discrete adjoint.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library, and only as the checker /
 * the timed CPU baseline.  The product (physicsbasedfwi2_b200/) never links or calls it.
 *
 * PARITY UNPINNED: the reference (/root/reference) contains no propagator arithmetic; the
 * path is the un-vendored, un-pinned third-party package deepwave (models/networks.py:10,
 * call sites :5408-5411, :5449, :5464, backward :5491).  This file restates the scheme of
 * SURVEY.md section 7.0 exactly as oracle/torch_ref.py does (same recurrences), but with
 * the adjoint written out by hand; tests require the two to agree to round-off in fp64.
 *
 * Scheme (per shot, padded grid nzp x nyp, halo h = order/2, never-updated outer ring):
 *   q    = Lap(u) + Dz(pz) + Dy(py)
 *   u+   = cu*u + cum*u- + a*w*q        cu = a(2 - dt^2 sz sy), cum = -a(1 - dt(sz+sy)/2),
 *                                        a = 1/(1 + dt(sz+sy)/2),  w = v^2 dt^2
 *   pz+  = (1 - dt sz) pz - dt (sz - sy) Dz(u)
 *   py+  = (1 - dt sy) py - dt (sy - sz) Dy(u)
 *   u+[inj cells] += coef * amp(t),      coef = w (forward sources) or a*w (adjoint receivers)
 * The adjoint runs the SAME recurrence on mu = a*w*lambda backwards in time (the scheme is
 * self-adjoint under that scaling; derivation in DESIGN.md), accumulating the imaging sum
 *   G += mu_{n+1} * q_n       so that   dJ/dv_pad = (2/v_pad) * (G + source-cell term).
 *
 * The per-cell operation order (explicit fma) is the one the CUDA kernels use, so the fp32
 * build of this file and the GPU agree far below the 1e-5 / 1e-4 parity tolerances.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static const double D2C[5][5] = {
    {0, 0, 0, 0, 0},
    {-2.0, 1.0, 0, 0, 0},
    {-5.0 / 2.0, 4.0 / 3.0, -1.0 / 12.0, 0, 0},
    {-49.0 / 18.0, 3.0 / 2.0, -3.0 / 20.0, 1.0 / 90.0, 0},
    {-205.0 / 72.0, 8.0 / 5.0, -1.0 / 5.0, 8.0 / 315.0, -1.0 / 560.0}};
static const double D1C[5][4] = {
    {0, 0, 0, 0},
    {1.0 / 2.0, 0, 0, 0},
    {2.0 / 3.0, -1.0 / 12.0, 0, 0},
    {3.0 / 4.0, -3.0 / 20.0, 1.0 / 60.0, 0},
    {4.0 / 5.0, -1.0 / 5.0, 4.0 / 105.0, -1.0 / 280.0}};

int fwi_oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* bench.py's CPU legs set the thread count explicitly: torchrun exports OMP_NUM_THREADS=1 */
void fwi_oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

#define REAL float
#define FMA fmaf
#define NAME(x) x##_f32
#include "cpu_ref_impl.h"
#undef REAL
#undef FMA
#undef NAME

#define REAL double
#define FMA fma
#define NAME(x) x##_f64
#include "cpu_ref_impl.h"
#undef REAL
#undef FMA
#undef NAME
