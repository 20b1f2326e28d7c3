// Cluster-resident small-grid time loop (acoustic_small.cu): parameters and entry point, shared with
// acoustic.cu which decides when to use it.  Not part of the C ABI.
#pragma once

namespace fwi {

struct SmallParams {
    int nzp, nyp, pitch, col0, h, pw, S, C;     // C = CTAs per cluster (= per shot)
    int rows_per;                               // most computational rows any CTA owns (sizes the shared-memory bands)
    int row_start[9];                           // CTA r owns the rows [row_start[r], row_start[r+1]) (cost-balanced split)
    int nsteps, adjoint, mode;                  // mode 0: plain, 1: write q, 2: read q + image
    int n_inj, inj_scale_a, n_smp, smp_div_a;
    int wave_pml;                               // 1: pz / py exist
    int async_halo;                             // 1: halo rows by st.async + mbarrier (set by small_run), 0: plain remote stores + cluster.sync
    long long stride;                           // floats per shot per field
    float c0, cz2[5], cy2[5], cz1[5], cy1[5], dt, dt2, hdt;
    const float *w, *sigz, *sigy;
    float *wave;                                // [fields][S][nzp][pitch]
    const int *inj_row_start, *inj_cell, *inj_col;
    const float *inj_amp;
    const int *smp_cell;
    float *smp_out;
    float *q_write;
    const float *q_read;
    float *image;
};

// Runs the whole time loop in one kernel when a shot fits a cluster's shared memory and all shots'
// clusters can be resident at once; *launched says whether it did (otherwise nothing was enqueued).
int small_run(const SmallParams &p, void *stream, bool *launched);

}  // namespace fwi
