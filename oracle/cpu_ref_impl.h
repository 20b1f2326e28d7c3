/* Included twice by cpu_ref.c (REAL = float / double).  See the header comment there.
 * Reference call sites restated: forward models/networks.py:5464, backward :5491. */

typedef struct {
    int nzp, nyp, h, pw;
    REAL dt, dt2, hdt;
    REAL c0, cz2[5], cy2[5], cz1[5], cy1[5];
    const REAL *w, *sigz, *sigy;
} NAME(ctx_t);

/* one row segment [y0, y1) with the full PML formula */
static void NAME(row_pml)(const NAME(ctx_t) * c, int z, int y0, int y1, const REAL *u, REAL *um,
                          const REAL *pz, const REAL *py, REAL *pzn, REAL *pyn,
                          REAL *qw, const REAL *qr, REAL *G) {
    const int nyp = c->nyp, h = c->h;
    const REAL sz = c->sigz[z];
    for (int y = y0; y < y1; ++y) {
        const size_t i = (size_t)z * nyp + y;
        const REAL sy = c->sigy[y];
        REAL lap = c->c0 * u[i];
        for (int k = 1; k <= h; ++k) lap = FMA(c->cy2[k], u[i - k] + u[i + k], lap);
        for (int k = 1; k <= h; ++k)
            lap = FMA(c->cz2[k], u[i - (size_t)k * nyp] + u[i + (size_t)k * nyp], lap);
        REAL dpz = 0, dpy = 0, duz = 0, duy = 0;
        for (int k = 1; k <= h; ++k) {
            dpz = FMA(c->cz1[k], pz[i + (size_t)k * nyp] - pz[i - (size_t)k * nyp], dpz);
            duz = FMA(c->cz1[k], u[i + (size_t)k * nyp] - u[i - (size_t)k * nyp], duz);
        }
        for (int k = 1; k <= h; ++k) {
            dpy = FMA(c->cy1[k], py[i + k] - py[i - k], dpy);
            duy = FMA(c->cy1[k], u[i + k] - u[i - k], duy);
        }
        const REAL q = (lap + dpz) + dpy;
        const REAL hs = c->hdt * (sz + sy);
        const REAL a = (REAL)1 / ((REAL)1 + hs);
        const REAL cu = a * FMA(-c->dt2, sz * sy, (REAL)2);
        const REAL cum = a * (hs - (REAL)1);
        const REAL aw = a * c->w[i];
        if (qr) G[i] = FMA(u[i], qr[i], G[i]);
        if (qw) qw[i] = q;
        um[i] = FMA(aw, q, FMA(cu, u[i], cum * um[i]));
        pzn[i] = FMA(c->dt * (sy - sz), duz, FMA(-c->dt, sz, (REAL)1) * pz[i]);
        pyn[i] = FMA(c->dt * (sz - sy), duy, FMA(-c->dt, sy, (REAL)1) * py[i]);
    }
}

/* one row segment of interior cells: sigma = 0 and no PML memory within reach */
static inline __attribute__((always_inline)) void NAME(row_int_h)(
    const NAME(ctx_t) * c, const int h, int z, int y0, int y1, const REAL *restrict u,
    REAL *restrict um, REAL *restrict qw, const REAL *restrict qr, REAL *restrict G) {
    const int nyp = c->nyp;
    const REAL *restrict w = c->w;
    for (int y = y0; y < y1; ++y) {
        const size_t i = (size_t)z * nyp + y;
        REAL lap = c->c0 * u[i];
        for (int k = 1; k <= h; ++k) lap = FMA(c->cy2[k], u[i - k] + u[i + k], lap);
        for (int k = 1; k <= h; ++k)
            lap = FMA(c->cz2[k], u[i - (size_t)k * nyp] + u[i + (size_t)k * nyp], lap);
        if (qr) G[i] = FMA(u[i], qr[i], G[i]);
        if (qw) qw[i] = lap;
        um[i] = FMA(w[i], lap, FMA((REAL)2, u[i], -um[i]));
    }
}

/* h is a literal in each call so the k-loops unroll and the y-loop vectorises */
static void NAME(row_int)(const NAME(ctx_t) * c, int z, int y0, int y1, const REAL *u, REAL *um,
                          REAL *qw, const REAL *qr, REAL *G) {
    switch (c->h) {
    case 1: NAME(row_int_h)(c, 1, z, y0, y1, u, um, qw, qr, G); break;
    case 2: NAME(row_int_h)(c, 2, z, y0, y1, u, um, qw, qr, G); break;
    case 3: NAME(row_int_h)(c, 3, z, y0, y1, u, um, qw, qr, G); break;
    default: NAME(row_int_h)(c, 4, z, y0, y1, u, um, qw, qr, G); break;
    }
}

/* Runs `nsteps` steps of the recurrence for S shots.
 *  adjoint == 0: n = 0 .. nsteps-1;  adjoint != 0: n = nsteps-1 .. 0 (time-reversed driver).
 *  per step n:  smp_out[n] = field[smp_cell] (/a if smp_div_a);   [G += field * q_read[n]];
 *               step (writing q_write[n] if given);   field+[inj_cell] += coef * inj_amp[n].
 *  All arrays use the padded pitch-free layout [.., nzp*nyp].  Returns 0, or 1 on bad args. */
static int NAME(acoustic_run_core)(int nzp, int nyp, int order, int pml_width, int S, double dz,
                                  double dy, double dt, const REAL *w, const REAL *sigz,
                                  const REAL *sigy, int nsteps, int adjoint, int n_inj,
                                  const int *inj_cell, const REAL *inj_amp, int inj_scale_a,
                                  int n_smp, const int *smp_cell, REAL *smp_out, int smp_div_a,
                                  REAL *q_write, const REAL *q_read, REAL *G, REAL *u_final,
                                  REAL *state, int save_every) {
    /* save_every = k > 1 (snapshot decimation): q is written / read and the image accumulated only at
     * steps n with n % k == 0, snapshot index n / k; the caller scales the image by k (rectangle rule) */
    const int k_img = save_every > 1 ? save_every : 1;
    if (order != 2 && order != 4 && order != 6 && order != 8) return 1;
    const int h = order / 2;
    if (nzp <= 2 * h || nyp <= 2 * h || S < 1 || nsteps < 0) return 1;
    NAME(ctx_t) c;
    c.nzp = nzp; c.nyp = nyp; c.h = h; c.pw = pml_width;
    c.dt = (REAL)dt; c.dt2 = c.dt * c.dt; c.hdt = (REAL)0.5 * c.dt;
    c.c0 = (REAL)(D2C[h][0] / (dz * dz) + D2C[h][0] / (dy * dy));
    for (int k = 1; k <= h; ++k) {
        c.cz2[k] = (REAL)(D2C[h][k] / (dz * dz));
        c.cy2[k] = (REAL)(D2C[h][k] / (dy * dy));
        c.cz1[k] = (REAL)(D1C[h][k - 1] / dz);
        c.cy1[k] = (REAL)(D1C[h][k - 1] / dy);
    }
    c.w = w; c.sigz = sigz; c.sigy = sigy;
    const size_t cells = (size_t)nzp * nyp;
    REAL *ua = (REAL *)calloc(S * cells, sizeof(REAL));
    REAL *ub = (REAL *)calloc(S * cells, sizeof(REAL));
    REAL *pza = (REAL *)calloc(S * cells, sizeof(REAL));
    REAL *pzb = (REAL *)calloc(S * cells, sizeof(REAL));
    REAL *pya = (REAL *)calloc(S * cells, sizeof(REAL));
    REAL *pyb = (REAL *)calloc(S * cells, sizeof(REAL));
    if (!ua || !ub || !pza || !pzb || !pya || !pyb) return 2;
    if (state) { /* carried state, fields in the order u, u-, pz, (scratch), py, (scratch) */
        memcpy(ua, state + 0 * S * cells, S * cells * sizeof(REAL));
        memcpy(ub, state + 1 * S * cells, S * cells * sizeof(REAL));
        memcpy(pza, state + 2 * S * cells, S * cells * sizeof(REAL));
        memcpy(pya, state + 4 * S * cells, S * cells * sizeof(REAL));
    }
    /* rows / columns that can see PML memory variables: frame widened by h */
    const int zlo = h + pml_width + h, zhi = nzp - h - pml_width - h;
    const int ylo = h + pml_width + h, yhi = nyp - h - pml_width - h;
    REAL *u = ua, *um = ub, *pz = pza, *pzn = pzb, *py = pya, *pyn = pyb;
    for (int it = 0; it < nsteps; ++it) {
        const int n = adjoint ? nsteps - 1 - it : it;
        if (n_smp > 0 && smp_out) {
            for (int s = 0; s < S; ++s)
                for (int r = 0; r < n_smp; ++r) {
                    const int cell = smp_cell[s * n_smp + r];
                    REAL val = u[s * cells + cell];
                    if (smp_div_a) {
                        const REAL hs = c.hdt * (sigz[cell / nyp] + sigy[cell % nyp]);
                        val = val * ((REAL)1 + hs); /* 1/a = 1 + hs */
                    }
                    smp_out[((size_t)n * S + s) * n_smp + r] = val;
                }
        }
#pragma omp parallel for collapse(2) schedule(static)
        for (int s = 0; s < S; ++s)
            for (int z = h; z < nzp - h; ++z) {
                const REAL *us = u + s * cells;
                REAL *ums = um + s * cells;
                const int img = (n % k_img) == 0;
                REAL *qw = (q_write && img) ? q_write + ((size_t)(n / k_img) * S + s) * cells : NULL;
                const REAL *qr = (q_read && img) ? q_read + ((size_t)(n / k_img) * S + s) * cells : NULL;
                REAL *Gs = G ? G + s * cells : NULL;
                if (!Gs) qr = NULL;
                const int pml_row = (pml_width > 0) && (z < zlo || z >= zhi);
                if (pml_row || yhi <= ylo) {
                    if (pml_width > 0)
                        NAME(row_pml)(&c, z, h, nyp - h, us, ums, pz + s * cells, py + s * cells,
                                      pzn + s * cells, pyn + s * cells, qw, qr, Gs);
                    else
                        NAME(row_int)(&c, z, h, nyp - h, us, ums, qw, qr, Gs);
                } else if (pml_width > 0) {
                    NAME(row_pml)(&c, z, h, ylo, us, ums, pz + s * cells, py + s * cells,
                                  pzn + s * cells, pyn + s * cells, qw, qr, Gs);
                    NAME(row_int)(&c, z, ylo, yhi, us, ums, qw, qr, Gs);
                    NAME(row_pml)(&c, z, yhi, nyp - h, us, ums, pz + s * cells, py + s * cells,
                                  pzn + s * cells, pyn + s * cells, qw, qr, Gs);
                } else {
                    NAME(row_int)(&c, z, h, nyp - h, us, ums, qw, qr, Gs);
                }
            }
        /* um now holds the new field; inject, then rotate */
        if (n_inj > 0 && inj_amp) {
            for (int s = 0; s < S; ++s)
                for (int r = 0; r < n_inj; ++r) {
                    const int cell = inj_cell[s * n_inj + r];
                    REAL coef = w[cell];
                    if (inj_scale_a) {
                        const REAL hs = c.hdt * (sigz[cell / nyp] + sigy[cell % nyp]);
                        coef = ((REAL)1 / ((REAL)1 + hs)) * coef;
                    }
                    um[s * cells + cell] =
                        FMA(coef, inj_amp[((size_t)n * S + s) * n_inj + r], um[s * cells + cell]);
                }
        }
        { REAL *t = u; u = um; um = t; }
        { REAL *t = pz; pz = pzn; pzn = t; }
        { REAL *t = py; py = pyn; pyn = t; }
    }
    if (u_final) memcpy(u_final, u, S * cells * sizeof(REAL));
    if (state) {
        memcpy(state + 0 * S * cells, u, S * cells * sizeof(REAL));
        memcpy(state + 1 * S * cells, um, S * cells * sizeof(REAL));
        memcpy(state + 2 * S * cells, pz, S * cells * sizeof(REAL));
        memcpy(state + 4 * S * cells, py, S * cells * sizeof(REAL));
    }
    free(ua); free(ub); free(pza); free(pzb); free(pya); free(pyb);
    return 0;
}

int NAME(fwi_oracle_acoustic_run)(int nzp, int nyp, int order, int pml_width, int S, double dz,
                                  double dy, double dt, const REAL *w, const REAL *sigz,
                                  const REAL *sigy, int nsteps, int adjoint, int n_inj,
                                  const int *inj_cell, const REAL *inj_amp, int inj_scale_a,
                                  int n_smp, const int *smp_cell, REAL *smp_out, int smp_div_a,
                                  REAL *q_write, const REAL *q_read, REAL *G, REAL *u_final) {
    return NAME(acoustic_run_core)(nzp, nyp, order, pml_width, S, dz, dy, dt, w, sigz, sigy, nsteps,
                                   adjoint, n_inj, inj_cell, inj_amp, inj_scale_a, n_smp, smp_cell,
                                   smp_out, smp_div_a, q_write, q_read, G, u_final, NULL, 1);
}

/* Same with snapshot decimation: q_write / q_read hold ceil(nsteps / save_every) snapshots. */
int NAME(fwi_oracle_acoustic_run_k)(int nzp, int nyp, int order, int pml_width, int S, double dz,
                                  double dy, double dt, const REAL *w, const REAL *sigz,
                                  const REAL *sigy, int nsteps, int adjoint, int n_inj,
                                  const int *inj_cell, const REAL *inj_amp, int inj_scale_a,
                                  int n_smp, const int *smp_cell, REAL *smp_out, int smp_div_a,
                                  REAL *q_write, const REAL *q_read, REAL *G, int save_every) {
    return NAME(acoustic_run_core)(nzp, nyp, order, pml_width, S, dz, dy, dt, w, sigz, sigy, nsteps,
                                   adjoint, n_inj, inj_cell, inj_amp, inj_scale_a, n_smp, smp_cell,
                                   smp_out, smp_div_a, q_write, q_read, G, NULL, NULL, save_every);
}

/* Same, continuing from (and leaving) the state [6][S][nzp*nyp] = u, u-, pz, -, py, - : lets a
 * caller cut the time loop into segments (tests of the checkpointed and x-slab drivers). */
int NAME(fwi_oracle_acoustic_run_state)(int nzp, int nyp, int order, int pml_width, int S, double dz,
                                  double dy, double dt, const REAL *w, const REAL *sigz,
                                  const REAL *sigy, int nsteps, int adjoint, int n_inj,
                                  const int *inj_cell, const REAL *inj_amp, int inj_scale_a,
                                  int n_smp, const int *smp_cell, REAL *smp_out, int smp_div_a,
                                  REAL *q_write, const REAL *q_read, REAL *G, REAL *state) {
    if (!state) return 1;
    return NAME(acoustic_run_core)(nzp, nyp, order, pml_width, S, dz, dy, dt, w, sigz, sigy, nsteps,
                                   adjoint, n_inj, inj_cell, inj_amp, inj_scale_a, n_smp, smp_cell,
                                   smp_out, smp_div_a, q_write, q_read, G, NULL, state, 1);
}
