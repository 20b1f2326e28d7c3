/* fwi_b200.h — C ABI of the B200-native finite-difference wave propagator.
 *
 * This is the drop-in boundary of the hot path: everything the reference reaches through
 *     deepwave.scalar.Propagator({'vp': v}, dx)(source_amplitudes, x_s, x_r, dt)
 * (reference call sites models/networks.py:5408-5411, :5449, :5464) and through the autograd
 * backward of that call (models/networks.py:5491).  The reference itself has no native
 * interface for this path (deepwave is an external Python/C/CUDA package), so the entry points
 * below are what a binding for the path has to offer: plain pointers, sizes and a CUDA stream.
 * No torch types appear here; physicsbasedfwi2_b200/scalar.py binds them with ctypes.
 *
 * All pointers are DEVICE pointers unless marked host.  All calls are asynchronous on `stream`
 * (a cudaStream_t passed as void*), never allocate, and return 0 on success or a negative
 * fwi_status; fwi_last_error() gives the thread-local message of the last failure.
 *
 * Scheme (derivation of the adjoint in DESIGN.md):
 *   q    = Lap(u) + Dz(pz) + Dy(py)
 *   u+   = cu*u + cum*u- + a*w*q ,  w = v^2 dt^2, a = 1/(1 + dt(sz+sy)/2)
 *   pz+  = (1 - dt sz) pz - dt (sz - sy) Dz(u),   py+ likewise
 *   u+[inj cells] += coef * amp(t)
 * The adjoint pass is the same recurrence run backwards in time on mu = a*w*lambda, so one
 * kernel family serves forward (models/networks.py:5464) and backward (:5491).
 */
#ifndef FWI_B200_H
#define FWI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum fwi_status {
    FWI_OK = 0,
    FWI_ERR_ARG = -1,     /* bad descriptor / null pointer / unsupported order */
    FWI_ERR_CUDA = -2,    /* a CUDA runtime call or launch failed */
    FWI_ERR_NODEVICE = -3 /* no sm_100 device */
};

/* ---- acoustic (constant density, 2-D) ------------------------------------------------- */

/* Problem descriptor (host POD).  Padded grid = model + 2*(pml_width + order/2) per axis; the
 * outermost order/2 ring is never updated and stays zero. */
typedef struct fwi_acoustic_desc {
    int32_t nzp, nyp;         /* padded grid extents (z slow, y fast) */
    int32_t order;            /* spatial accuracy: 2, 4, 6 or 8 */
    int32_t pml_width;        /* cells of PML on each of the four sides (0 = none) */
    int32_t n_shots;          /* shots advanced together by one call */
    int32_t chunk_rows;       /* cap on the rows of one work segment (tuning / tests; 0 = none) */
    int32_t nsteps;           /* inner time steps */
    int32_t adjoint;          /* 0: n = 0..nsteps-1;  1: n = nsteps-1..0 (time-reversed driver) */
    int32_t n_inj;            /* injection points per shot (sources fwd / receivers adj) */
    int32_t inj_scale_a;      /* 1: injection coefficient a*w (adjoint), 0: w (forward) */
    int32_t n_smp;            /* sampling points per shot (receivers fwd / sources adj) */
    int32_t smp_div_a;        /* 1: sampled value divided by a (adjoint source gradient) */
    double dz, dy, dt;        /* grid spacings (m) and inner time step (s) */
    int32_t flags;            /* FWI_FLAG_* */
    int32_t save_every;       /* k > 1: snapshot decimation - q_write / q_read hold ceil(nsteps / k) snapshots, those
                                 of the steps n with n % k == 0, and only those steps are imaged (the caller weighs the image by k);
                                 0 or 1: every step (the parity-graded mode) */
} fwi_acoustic_desc;

#define FWI_FLAG_NO_PDL 1     /* plain launches instead of programmatic dependent launches: for jobs that share the device with
                                 other jobs (fwi_acoustic_run_multi) - a dependent launch made resident early would sit on SM
                                 slots the other job's ready CTAs could use */

/* Device memory layout chosen by the library for a descriptor (host POD, output). */
typedef struct fwi_acoustic_layout {
    int32_t pitch;            /* floats per grid row in device arrays (multiple of 32) */
    int32_t col0;             /* device column that holds padded column 0 (first interior column is
                                 128-byte aligned) */
    int64_t shot_stride;      /* floats per shot per field = nzp * pitch */
    int32_t n_images;         /* image partial buffers (one per shot: single-owner accumulation) */
    int32_t wave_fields;      /* number of [n_shots][nzp][pitch] fields in the workspace */
    int32_t chunk_rows;       /* cap on segment rows actually used (0 = none) */
    int32_t units_per_shot;   /* strip columns of work per shot per step */
    int64_t unit_table_bytes; /* size of the work-table workspace fwi_acoustic_run needs */
} fwi_acoustic_layout;

/* Buffers of one propagation (host struct of device pointers). */
typedef struct fwi_acoustic_buffers {
    const float *w;           /* [nzp][pitch]   v^2 dt^2 on the padded grid */
    const float *sigz;        /* [nzp]          PML profile along z */
    const float *sigy;        /* [pitch]        PML profile along y, device columns */
    float *wave;              /* [wave_fields][n_shots][nzp][pitch], zero on entry: u, u-, pz, pz+, py, py+ */
    const int32_t *inj_row_start; /* [n_shots][nzp + 1] absolute offsets into the two lists below:
                                 entries of shot s in rows [z0, z1) are [start[s][z0], start[s][z1]) */
    const int32_t *inj_cell;  /* [n_shots*n_inj] device cell (z*pitch + col), sorted by (shot, cell), stable */
    const int32_t *inj_col;   /* [n_shots*n_inj] column of inj_amp each entry reads */
    const float *inj_amp;     /* [nsteps][n_shots][n_inj] */
    const int32_t *smp_cell;  /* [n_shots][n_smp] device cell */
    float *smp_out;           /* [nsteps][n_shots][n_smp] */
    float *q_write;           /* [nsteps][n_shots][nzp][pitch] or NULL: snapshot of q per step */
    const float *q_read;      /* same layout or NULL: snapshots consumed by the imaging sum */
    float *image;             /* [n_images][nzp][pitch] or NULL: image[s] += field * q_read */
    void *unit_table;         /* [unit_table_bytes] scratch: the library writes its work table here */
    int64_t q_step_stride;    /* floats between the snapshots of consecutive steps in q_write; 0 = n_shots * shot_stride.
                                 Lets a forward call that advances a SUBSET of the shots of a gradient (forward launches
                                 are fastest when their wavefields fit L2) write into the snapshot array of all of them */
} fwi_acoustic_buffers;

/* Fills `out` for `desc`.  Replaces nothing in the reference (layout is private to deepwave
 * there); needed so a host can size and fill the arrays above. */
int fwi_acoustic_get_layout(const fwi_acoustic_desc *desc, fwi_acoustic_layout *out);

/* Host-only: cuts the work of one launch into `n_units` cost-balanced warp units exactly as
 * fwi_acoustic_run does for a device with that many resident warps, checks that the segments tile
 * the computational region, and reports the balance (estimated cost of the heaviest unit vs the
 * mean).  Any output pointer may be NULL. */
int fwi_acoustic_plan_stats(const fwi_acoustic_desc *desc, int n_units, int64_t *n_segments,
                            int32_t *n_units_used, double *max_cost, double *mean_cost);

/* Runs desc->nsteps steps.  Per step n:  smp_out[n] = field[smp_cell] (/a);  [image += field *
 * q_read[n]];  advance (writing q_write[n]);  field+[inj_cell] += coef * inj_amp[n].
 * adjoint=0 replaces the time loop of deepwave's forward (models/networks.py:5464);
 * adjoint=1 with q_read/image replaces its backward (models/networks.py:5491). */
int fwi_acoustic_run(const fwi_acoustic_desc *desc, const fwi_acoustic_buffers *buf, void *stream);

/* Several independent time loops at once (e.g. the shot groups of one gradient), job j on streams[j]: their launches
 * interleave on the device, so the ramp and tail of one job's step are covered by the other job's step.  Each job is what
 * fwi_acoustic_run takes; the caller orders the streams against its own work. */
int fwi_acoustic_run_multi(int32_t n_jobs, const fwi_acoustic_desc *const *descs, const fwi_acoustic_buffers *const *bufs,
                           void *const *streams);

/* ---- x-slab domain decomposition: ghost-zone exchange by peer stores over NVLink ---------- */
/* One shot spread over the GPUs of a box (BASELINE config 4 variant; no reference
 * counterpart - the reference runs every shot whole on one GPU, models/networks.py:5316).  Each
 * rank advances a column window of the grid with ghost_cols = halo_steps * order/2 extra columns
 * per inner edge.  Every halo_steps steps the native time loop (a) packs its outermost owned
 * columns of the live fields (u, u-, pz, py) and stores them straight into the neighbour's staging
 * buffer through a peer mapping, then raises the neighbour's flag (release, system scope), and
 * (b) waits for its own flags and unpacks its staging buffer into its ghost columns.  Staging is
 * double-buffered by exchange parity, so a rank can run at most one exchange ahead of a neighbour
 * without overwriting data still in use.  No host synchronisation inside the loop.
 *
 * Exchange memory of one rank ("comm block", allocated with fwi_peer_alloc, mapped by the two
 * neighbours with fwi_peer_open): 256 header bytes = uint32 flags_in[2] (side 0: raised by the left
 * neighbour, side 1: by the right), uint32 counter[2], uint32 status; then
 * float stage[2 parity][2 side][slot_floats], slot_floats = live * n_shots * nzp * ghost_cols. */
typedef struct fwi_slab_link {
    int32_t halo_steps;       /* steps between exchanges (>= 1) */
    int32_t ghost_cols;       /* halo_steps * order/2 */
    int32_t has_left, has_right;
    int32_t send_left_col, send_right_col;   /* first DEVICE column of the owned columns sent to each side */
    int32_t ghost_left_col, ghost_right_col; /* first DEVICE column of each ghost zone */
    void *comm;               /* this rank's comm block */
    void *peer_left, *peer_right; /* the neighbours' comm blocks (peer mappings), NULL where none */
    uint32_t epoch;           /* exchanges completed before this call; the call's k-th exchange is epoch + k */
    uint32_t timeout_ms;      /* a flag wait longer than this sets status = 1 and stops waiting (0 = 20000) */
} fwi_slab_link;

#define FWI_SLAB_HEADER_BYTES 256

/* Peer-visible device memory for comm blocks (cudaMalloc + CUDA IPC; zero-filled).  `handle` is
 * 64 opaque bytes another process of the same box passes to fwi_peer_open. */
int fwi_peer_alloc(int64_t bytes, void **ptr, unsigned char handle[64]);
int fwi_peer_open(const unsigned char handle[64], void **ptr);
int fwi_peer_close(void *ptr);
int fwi_peer_free(void *ptr);
/* Reads the status word of a comm block (synchronises `stream`): 0 ok, 1 a flag wait timed out. */
int fwi_slab_status(const void *comm, void *stream, uint32_t *status);
/* Number of exchanges fwi_acoustic_run_slab performs for `nsteps` steps (host arithmetic). */
int32_t fwi_slab_exchanges(int32_t nsteps, int32_t halo_steps);

/* fwi_acoustic_run on a window of the grid with in-loop ghost exchange.  Forward: exchange after
 * step n when (n + 1) % halo_steps == 0 and n + 1 < nsteps.  Adjoint: after step n when
 * (nsteps - n) % halo_steps == 0 and n > 0.  All ranks must call with the same nsteps. */
int fwi_acoustic_run_slab(const fwi_acoustic_desc *desc, const fwi_acoustic_buffers *buf,
                          const fwi_slab_link *link, void *stream);

/* Named wrappers with the direction fixed (argument checks differ). */
int fwi_acoustic_forward(const fwi_acoustic_desc *desc, const fwi_acoustic_buffers *buf, void *stream);
int fwi_acoustic_backward(const fwi_acoustic_desc *desc, const fwi_acoustic_buffers *buf, void *stream);

/* ---- host-side preparation as device kernels ------------------------------------------------ */
/* What deepwave's Python layer does around its native time loop for the call at models/networks.py:5449,5464 and
 * its backward (:5491): pad the model (replicate), scale it to w = v^2 dt^2, build the PML profiles; turn metre
 * coordinates into cells and the sorted injection lists of fwi_acoustic_buffers; fold the image of the adjoint pass
 * back through the padding into dJ/dv.  Three launches instead of ~125 eager torch micro-kernels per gradient. */
typedef struct fwi_model_desc {
    int32_t nz, ny;           /* model extents (z slow, y fast) */
    int32_t order;            /* spatial accuracy: 2, 4, 6 or 8 */
    int32_t pml_width;
    int32_t c0, c1;           /* window [c0, c1) of padded columns this rank works on (x-slabs); c1 <= 0: the whole grid */
    double dz, dy, dt;        /* grid spacings and INNER time step */
    double vmax;              /* velocity the inner step and the PML profiles are built for */
} fwi_model_desc;

/* v [nz][ny] -> w [nzp][pitch] (layout of fwi_acoustic_get_layout for the padded window), sigz [nzp], sigy [pitch]. */
int fwi_acoustic_prepare_model(const fwi_model_desc *m, const float *v, float *w, float *sigz, float *sigy, void *stream);

/* x [n_shots][n][2] metres (z, y) -> cell [n_shots][n] (device cell z * pitch + col, clamped into the model) and the
 * lists fwi_acoustic_run takes: row_start [n_shots][nzp + 1], sorted_cell / sorted_col [n_shots * n] (stable sort by
 * (shot, cell)).  loc_round 0: cell = floor(x / d + 1e-6); 1: round half to even.  *bad (device int32, caller zeroes it)
 * is set to 1 when a point lies outside the model (the launch stays safe: points are clamped). */
int fwi_acoustic_plan_points(const fwi_model_desc *m, const float *x, int32_t n_shots, int32_t n, int32_t loc_round,
                             int32_t *cell, int32_t *row_start, int32_t *sorted_cell, int32_t *sorted_col, int32_t *bad,
                             void *stream);

/* dJ/dv [nz][ny] = fold( (2 / v_pad) * (sum_s image[s] + source-cell term) ): image [n_shots][nzp][pitch] from
 * fwi_acoustic_backward, gs [nsteps][n_shots][n_src] its sampled field (smp_out), amp the forward injection amplitudes,
 * src_cell [n_shots][n_src] device cells; gs may be NULL (no source term).  workspace: fwi_acoustic_fold_workspace_floats. */
int64_t fwi_acoustic_fold_workspace_floats(const fwi_model_desc *m, int32_t n_shots, int32_t n_src);
int fwi_acoustic_fold_gradient(const fwi_model_desc *m, int32_t n_shots, const float *image, const float *gs,
                               const float *amp, const int32_t *src_cell, int32_t n_src, int32_t nsteps, const float *v,
                               float *workspace, float *grad_v, void *stream);

/* ---- elastic (2-D P-SV, velocity-stress staggered grid, C-PML) -------------------------- */
/* Replaces what the reference reaches through pyapi_denise: Denise.grad(model, src, rec)
 * (models/networks.py:9173) = elastic forward + adjoint + correlation, gradients read back at
 * :9185-9188.  Scheme: oracle/elastic_ref.py.  Arrays are plain row-major [nzp][nyp] (no pitch). */

typedef struct fwi_elastic_desc {
    int32_t nzp, nyp;         /* padded grid extents (z slow, y fast) */
    int32_t order;            /* staggered-grid accuracy: 2, 4, 6 or 8 */
    int32_t pml_width;        /* cells of C-PML on each side (0 = none) */
    int32_t n_shots;
    int32_t nsteps;           /* inner time steps */
    int32_t n_src;            /* explosive sources per shot */
    int32_t n_rec;            /* two-component receivers per shot */
    double dz, dy, dt;
    int32_t source_kind;      /* 0: explosive (sxx, szz += dt s; DENISE QUELLTYPB 1), 1: vertical force (vz += dt bz s;
                                 QUELLTYPB 2, models/networks.py:10452), 2: horizontal force (vy += dt by s) */
    int32_t free_surface;     /* 1: stress-imaging free surface at the first model row (row pml_width + order/2 of the padded
                                 grid) instead of a top C-PML (DENISE FREE_SURF = 1, models/networks.py:9811): szz = 0 there,
                                 szz / syz antisymmetric about it, nothing above it is updated */
    int32_t adj_parity;       /* backward in segments (wavefield checkpointing): number of adjoint steps already taken
                                 on this wave / scratch pair, modulo 2 (the memory variables ping-pong between them) */
} fwi_elastic_desc;

typedef struct fwi_elastic_buffers {
    const float *l2m, *lam, *mu, *by, *bz; /* [nzp][nyp] lambda+2mu, lambda at (i,j); mu at (i+1/2,j+1/2);
                                 buoyancy at (i,j+1/2) and (i+1/2,j) */
    const float *prof_z;      /* [4][nzp] C-PML a, b at integer nodes, a, b at half nodes, along z */
    const float *prof_y;      /* [4][nyp] likewise along y */
    float *wave;              /* [22][n_shots][nzp][nyp] (fwi_elastic_wave_floats), zero on entry: vy vz sxx szz syz psi0..7
                                 and a second copy of vy vz sxx szz syz psi0..3 (the fused forward step ping-pongs)
                                 (backward: their adjoint variables) */
    const int32_t *src_cell;  /* [n_shots][n_src] cell = z*nyp + y */
    const int32_t *src_link;  /* [n_shots][n_src] chains of points sharing a cell: bits 0..29 = index of the
                                 next point of the chain + 1 (0 = last), bit 30 set = not the first of its chain */
    const float *src_amp;     /* [nsteps][n_shots][n_src] (forward; backward too when source_kind != 0: the force
                                 sources depend on the buoyancy, whose image gets a source-cell term) */
    const int32_t *rec_cell;  /* [n_shots][n_rec] */
    const int32_t *rec_link;  /* [n_shots][n_rec] as src_link (used by the adjoint injection) */
    float *rec_out;           /* [2][nsteps][n_shots][n_rec]: vy, vz recorded before each step (forward) */
    float *snap;              /* [nsteps][5][n_shots][nzp][nyp] fy fz e1 e2 e3: written by forward when not
                                 NULL, read by backward */
    float *scratch;           /* [8][n_shots][nzp][nyp] zero on entry (backward) */
    float *grad;              /* [5][n_shots][nzp][nyp] zero on entry: images for l2m lam mu by bz (backward) */
    const float *grad_rec;    /* [2][nsteps][n_shots][n_rec] dJ/d(rec_out) (backward) */
    float *grad_src;          /* [nsteps][n_shots][n_src] dJ/d(src_amp) (backward) */
} fwi_elastic_buffers;

int64_t fwi_elastic_wave_floats(const fwi_elastic_desc *desc);   /* floats in `wave`; -1 on a bad descriptor */
int fwi_elastic_forward(const fwi_elastic_desc *desc, const fwi_elastic_buffers *buf, void *stream);
int fwi_elastic_backward(const fwi_elastic_desc *desc, const fwi_elastic_buffers *buf, void *stream);

/* ---- trace misfit fused with its backward ------------------------------------------------ */
/* Replaces the torch glue between the propagator's output and loss.backward() in every `prop`
 * method of the reference (models/networks.py:5467-5476, backward :5491):
 *   d = pred - direct;  m = max_t |d| per trace;  n = d / (m + eps);
 *   loss = mean |n - obs| (kind 0, torch.nn.L1Loss) or mean (n - obs)^2 (kind 1, MSELoss);
 *   grad_pred = grad_scale * dloss/dpred.
 * Arrays are [nt][n_traces] (n_traces = shots * receivers, contiguous).  `direct` may be NULL
 * (no direct-wave subtraction).  fwi_trace_misfit writes the scalar loss and leaves the per-trace
 * maxima and sums in `workspace`; fwi_trace_misfit_backward (same desc, arrays and workspace) writes
 * grad_pred = (*grad_loss) * dloss/dpred, grad_loss a device scalar or NULL (= 1).  Deterministic. */
typedef struct fwi_misfit_desc {
    int32_t nt, n_traces;
    int32_t kind;             /* 0 = L1, 1 = MSE */
    int32_t normalize;        /* 1: divide every trace by its max |.| + eps (models/networks.py:5468-5470) */
    float eps;                /* 1e-10 in the reference */
} fwi_misfit_desc;

int64_t fwi_trace_misfit_workspace_bytes(const fwi_misfit_desc *desc);
int fwi_trace_misfit(const fwi_misfit_desc *desc, const float *pred, const float *direct, const float *obs,
                     float *loss, void *workspace, void *stream);
int fwi_trace_misfit_backward(const fwi_misfit_desc *desc, const float *pred, const float *direct,
                              const float *obs, const float *grad_loss, float *grad_pred, void *workspace,
                              void *stream);

/* Allocation-free variant of fwi_trace_misfit.  It has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  Same arguments and result as fwi_trace_misfit without `workspace`: the
 * partials live in a file-scope `static __device__` array of FWI_TRACE_MISFIT_STATIC_MAX_BYTES that every GPU loading
 * the library holds until exit.  Returns FWI_ERR_ARG (with a message naming both sizes) when
 * fwi_trace_misfit_workspace_bytes(desc) exceeds that maximum.  Calls no cudaMalloc / cudaFree / allocating routine.
 * All calls share the array (it keeps the previous call's contents; nothing needs to start at zero): calls that overlap
 * on one GPU - two streams, two threads - race.  fwi_trace_misfit_backward needs the partials of a forward call in a
 * workspace of the caller's, so it pairs with fwi_trace_misfit, not with this entry point. */
#define FWI_TRACE_MISFIT_STATIC_MAX_BYTES (4 * 1024 * 1024)
int fwi_trace_misfit_noalloc(const fwi_misfit_desc *desc, const float *pred, const float *direct, const float *obs,
                             float *loss, void *stream);

/* ---- the steps either side of the elastic gradient (DENISE glue of the reference's elastic `prop`) ----------- */
/* Frequency stage of d.add_fwi_stage(fc_low=0.0, fc_high=10.0, ...) (models/networks.py:9147): Butterworth low-pass of
 * even `order` on every trace of in [nt][n_traces] (traces fast), zero initial state; zero_phase: a second pass
 * backwards in time (the operator is then symmetric: it is its own adjoint).  in == out is allowed. */
typedef struct fwi_butter_desc {
    int32_t nt, n_traces;
    int32_t order;            /* even, 2..16 */
    int32_t zero_phase;
    double dt, fc_low, fc_high; /* fc_low must be 0 (low-pass) */
} fwi_butter_desc;
int fwi_butterworth_lowpass(const fwi_butter_desc *desc, const float *in, float *out, void *stream);

/* Gradient post-processing of models/networks.py:9194-9240, in place on three [nz][ny] gradients: rho gradient times
 * (row index)^depth_pow, rows [0, water_rows) of all three zeroed, each gradient scaled by max|model| / max|gradient|,
 * rho additionally times rho_factor plus rho_offset.  stats: 6 floats of device scratch (the maxima, left for inspection). */
typedef struct fwi_gradpost_desc {
    int32_t nz, ny, water_rows;
    double depth_pow, rho_factor, rho_offset;   /* 1.5, 0.5, 1e-10 in the reference */
} fwi_gradpost_desc;
int fwi_elastic_gradient_post(const fwi_gradpost_desc *desc, float *grad_vp, float *grad_vs, float *grad_rho,
                              const float *vp, const float *vs, const float *rho, float *stats, void *stream);

/* ---- common ---------------------------------------------------------------------------- */

const char *fwi_last_error(void);     /* thread-local, never NULL */
int fwi_version(void);                /* ABI version, bumps on any struct change */
int64_t fwi_kernel_launches(void);    /* kernels launched by this library in this process */
int fwi_device_ok(int device);        /* FWI_OK if `device` is sm_100 */

#ifdef __cplusplus
}
#endif
#endif /* FWI_B200_H */
