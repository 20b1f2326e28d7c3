// TORCH_LIBRARY(b200fwi, ...): the host side of the scalar propagator as PyTorch C++ operators over the C ABI of
// libfwi_b200.so (include/fwi_b200.h).  This is what a maintainer binds where the reference reaches deepwave's C++/CUDA
// extension from models/networks.py:5449,5464 (forward) and :5491 (backward): argument checks (TORCH_CHECK ->
// RuntimeError, as deepwave raises), workspaces from the caching allocator, launches on the current CUDA stream.
// physicsbasedfwi2_b200/scalar.py wraps the operators in a torch.autograd.Function.  No process-global state: every
// tensor an operator needs is an argument or allocated in the call.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <cuda_runtime.h>
#include <torch/library.h>
#include <torch/torch.h>

#include <cmath>
#include <cstdlib>
#include <tuple>
#include <vector>

#include "../../include/fwi_b200.h"

namespace {

using at::Tensor;

void check_rc(int rc, const char *what) {
    TORCH_CHECK(rc == FWI_OK, what, " failed (status ", rc, "): ", fwi_last_error());
}

void *stream_of(const Tensor &t) { return static_cast<void *>(c10::cuda::getCurrentCUDAStream(t.device().index()).stream()); }

const float *fptr(const Tensor &t) { return t.defined() && t.numel() ? t.data_ptr<float>() : nullptr; }
float *fptr_mut(Tensor &t) { return t.defined() && t.numel() ? t.data_ptr<float>() : nullptr; }
const int32_t *iptr(const Tensor &t) { return t.defined() && t.numel() ? t.data_ptr<int32_t>() : nullptr; }

void check_model(const Tensor &v) {
    TORCH_CHECK(v.dim() == 2, "model['vp'] must be 2-D [nz, ny], got ", v.dim(), " dimensions");
    TORCH_CHECK(v.is_cuda(), "model['vp'] must be a CUDA tensor: this propagator has no CPU path");
    TORCH_CHECK(v.scalar_type() == at::kFloat, "model['vp'] must be float32, got ", v.scalar_type());
}

fwi_model_desc model_desc(const Tensor &v, int64_t order, int64_t pml, double dz, double dy, double dti, double vmax,
                          int64_t c0, int64_t c1) {
    fwi_model_desc m;
    m.nz = (int32_t)v.size(0); m.ny = (int32_t)v.size(1);
    m.order = (int32_t)order; m.pml_width = (int32_t)pml;
    m.c0 = (int32_t)c0; m.c1 = (int32_t)c1;
    m.dz = dz; m.dy = dy; m.dt = dti; m.vmax = vmax;
    return m;
}

// ---- plan: inner time step from the CFL limit and the device layout ------------------------------------------
// returns a CPU float64 tensor [vmax, ratio, dt_inner, pitch, col0, nzp, nyp, wave_fields, shot_stride, unit_table_bytes]
Tensor scalar_plan(const Tensor &v, double dt, double dz, double dy, int64_t order, int64_t pml, double vpmax,
                   int64_t n_shots, int64_t chunk_rows, int64_t c0, int64_t c1) {
    check_model(v);
    TORCH_CHECK(order == 2 || order == 4 || order == 6 || order == 8, "accuracy must be 2, 4, 6 or 8, got ", order);
    TORCH_CHECK(pml >= 0, "pml_width must be >= 0");
    TORCH_CHECK(dz > 0 && dy > 0, "grid spacings must be > 0");
    double vmax = vpmax;
    if (!(vpmax > 0)) vmax = v.detach().abs().max().item<double>();     // (deepwave does the same: a host sync)
    TORCH_CHECK(vmax > 0 && std::isfinite(vmax), "maximum velocity must be positive and finite, got ", vmax);
    const double limit = 0.6 / (std::sqrt(1.0 / (dz * dz) + 1.0 / (dy * dy)) * vmax);
    int64_t ratio = (int64_t)std::ceil(std::fabs(dt) / limit - 1e-12);
    if (ratio < 1) ratio = 1;
    const int64_t P = pml + order / 2;
    const int64_t nzp = v.size(0) + 2 * P, nyp_global = v.size(1) + 2 * P;
    const int64_t w1 = c1 > 0 ? c1 : nyp_global;
    TORCH_CHECK(0 <= c0 && c0 < w1 && w1 <= nyp_global && w1 - c0 > 2 * P, "bad column window [", c0, ", ", w1, ")");
    fwi_acoustic_desc d = {};
    d.nzp = (int32_t)nzp; d.nyp = (int32_t)(w1 - c0); d.order = (int32_t)order; d.pml_width = (int32_t)pml;
    d.n_shots = (int32_t)n_shots; d.chunk_rows = (int32_t)chunk_rows; d.dz = dz; d.dy = dy; d.dt = dt / ratio;
    fwi_acoustic_layout lay;
    check_rc(fwi_acoustic_get_layout(&d, &lay), "fwi_acoustic_get_layout");
    Tensor meta = at::empty({10}, at::TensorOptions().dtype(at::kDouble));
    double *m = meta.data_ptr<double>();
    m[0] = vmax; m[1] = (double)ratio; m[2] = dt / ratio; m[3] = lay.pitch; m[4] = lay.col0; m[5] = (double)nzp;
    m[6] = (double)(w1 - c0); m[7] = lay.wave_fields; m[8] = (double)lay.shot_stride; m[9] = (double)lay.unit_table_bytes;
    return meta;
}

// ---- model -> (w, sigz, sigy) ----------------------------------------------------------------------------------
std::tuple<Tensor, Tensor, Tensor> scalar_prepare(const Tensor &v, const Tensor &meta, int64_t order, int64_t pml, double dz,
                                                  double dy, int64_t c0, int64_t c1) {
    check_model(v);
    TORCH_CHECK(meta.device().is_cpu() && meta.scalar_type() == at::kDouble && meta.numel() == 10, "bad plan tensor");
    const double *m = meta.data_ptr<double>();
    c10::cuda::CUDAGuard guard(v.device());
    const Tensor vc = v.detach().contiguous();
    const auto opt = vc.options();
    Tensor w = at::empty({(int64_t)m[5], (int64_t)m[3]}, opt);
    Tensor sigz = at::empty({(int64_t)m[5]}, opt), sigy = at::empty({(int64_t)m[3]}, opt);
    const fwi_model_desc md = model_desc(vc, order, pml, dz, dy, m[2], m[0], c0, c1);
    check_rc(fwi_acoustic_prepare_model(&md, vc.data_ptr<float>(), w.data_ptr<float>(), sigz.data_ptr<float>(),
                                        sigy.data_ptr<float>(), stream_of(vc)), "fwi_acoustic_prepare_model");
    return {w, sigz, sigy};
}

// ---- metres -> cells + sorted injection lists --------------------------------------------------------------------
// returns (cell [S, n] int32, row_start [S, nzp + 1], sorted_cell [S n], sorted_col [S n]); ORs the out-of-model flag
// into `bad` (int32 [1] on the device)
std::tuple<Tensor, Tensor, Tensor, Tensor> scalar_plan_points(const Tensor &x, const Tensor &v, const Tensor &meta, Tensor bad,
                                                              int64_t order, int64_t pml, double dz, double dy, bool loc_round,
                                                              int64_t c0, int64_t c1) {
    check_model(v);
    TORCH_CHECK(x.dim() == 3 && x.size(2) == 2, "locations must be [n_shots, n, 2], got ", x.sizes());
    TORCH_CHECK(bad.is_cuda() && bad.scalar_type() == at::kInt && bad.numel() == 1, "bad flag must be a CUDA int32 scalar");
    const double *m = meta.data_ptr<double>();
    c10::cuda::CUDAGuard guard(v.device());
    const Tensor xc = x.detach().to(v.device(), at::kFloat).contiguous();
    const int64_t S = xc.size(0), n = xc.size(1), nzp = (int64_t)m[5];
    const auto iopt = v.options().dtype(at::kInt);
    Tensor cell = at::empty({S, n}, iopt), start = at::empty({S, nzp + 1}, iopt);
    Tensor sc = at::empty({std::max<int64_t>(S * n, 1)}, iopt), so = at::empty({std::max<int64_t>(S * n, 1)}, iopt);
    const fwi_model_desc md = model_desc(v, order, pml, dz, dy, m[2], m[0], c0, c1);
    check_rc(fwi_acoustic_plan_points(&md, n ? xc.data_ptr<float>() : nullptr, (int32_t)S, (int32_t)n, loc_round ? 1 : 0,
                                      n ? cell.data_ptr<int32_t>() : nullptr, start.data_ptr<int32_t>(), sc.data_ptr<int32_t>(),
                                      so.data_ptr<int32_t>(), bad.data_ptr<int32_t>(), stream_of(v)),
             "fwi_acoustic_plan_points");
    return {cell, start, sc, so};
}

// ---- one native time loop ---------------------------------------------------------------------------------------
// wave [wave_fields * S * shot_stride] (state in / out, zero = at rest); returns smp_out [nsteps, S, n_smp]
Tensor scalar_run(const Tensor &w, const Tensor &sigz, const Tensor &sigy, Tensor wave, const Tensor &inj_start,
                  const Tensor &inj_cell, const Tensor &inj_col, const Tensor &inj_amp, const Tensor &smp_cell,
                  const c10::optional<Tensor> &q_write, const c10::optional<Tensor> &q_read, const c10::optional<Tensor> &image,
                  const Tensor &meta, int64_t n_shots, int64_t nsteps, bool adjoint, int64_t order, int64_t pml, double dz,
                  double dy, int64_t chunk_rows, int64_t q_step_stride, int64_t save_every) {
    TORCH_CHECK(w.is_cuda() && wave.is_cuda(), "scalar_run needs CUDA tensors: there is no CPU path");
    const double *m = meta.data_ptr<double>();
    c10::cuda::CUDAGuard guard(w.device());
    fwi_acoustic_desc d = {};
    d.nzp = (int32_t)m[5]; d.nyp = (int32_t)m[6]; d.order = (int32_t)order; d.pml_width = (int32_t)pml;
    d.n_shots = (int32_t)n_shots; d.chunk_rows = (int32_t)chunk_rows; d.nsteps = (int32_t)nsteps; d.adjoint = adjoint ? 1 : 0;
    d.n_inj = inj_amp.defined() && inj_amp.numel() ? (int32_t)inj_amp.size(2) : 0;
    d.inj_scale_a = adjoint ? 1 : 0;
    d.n_smp = smp_cell.defined() && smp_cell.numel() ? (int32_t)smp_cell.size(1) : 0;
    d.smp_div_a = adjoint ? 1 : 0;
    d.dz = dz; d.dy = dy; d.dt = m[2];
    d.save_every = (int32_t)save_every;
    TORCH_CHECK(wave.numel() == (int64_t)m[7] * n_shots * (int64_t)m[8], "wave workspace has the wrong size");
    TORCH_CHECK(d.n_inj == 0 || (inj_amp.is_contiguous() && inj_amp.size(0) == nsteps && inj_amp.size(1) == n_shots),
                "injection amplitudes must be contiguous [nsteps, n_shots, n]");
    Tensor smp_out = at::empty({nsteps, n_shots, (int64_t)d.n_smp}, w.options());
    Tensor table = at::empty({(int64_t)m[9]}, w.options().dtype(at::kByte));
    fwi_acoustic_buffers b = {};
    b.w = w.data_ptr<float>(); b.sigz = fptr(sigz); b.sigy = fptr(sigy); b.wave = wave.data_ptr<float>();
    b.inj_row_start = iptr(inj_start); b.inj_cell = iptr(inj_cell); b.inj_col = iptr(inj_col); b.inj_amp = fptr(inj_amp);
    b.smp_cell = iptr(smp_cell); b.smp_out = smp_out.numel() ? smp_out.data_ptr<float>() : nullptr;
    Tensor qw = q_write.has_value() ? *q_write : Tensor(), im = image.has_value() ? *image : Tensor();
    b.q_write = fptr_mut(qw);
    b.q_read = q_read.has_value() ? fptr(*q_read) : nullptr;
    b.image = fptr_mut(im);
    b.unit_table = table.data_ptr();
    b.q_step_stride = q_step_stride;
    check_rc(adjoint ? fwi_acoustic_backward(&d, &b, stream_of(w)) : fwi_acoustic_forward(&d, &b, stream_of(w)),
             adjoint ? "fwi_acoustic_backward" : "fwi_acoustic_forward");
    return smp_out;
}

// ---- several shot groups of one gradient at once, one stream per group ---------------------------------------
// Group j advances the shots [shot0[j], shot0[j] + n_j) of a gradient whose snapshot array q (and image) holds all
// n_total shots; the groups' launches interleave on the device (fwi_acoustic_run_multi), so one group's step covers the
// ramp and tail of the other's.  Group 0 runs on the current stream, the others on pool streams that wait for it and
// that it waits for: to the caller this is one asynchronous call on the current stream.
std::vector<Tensor> scalar_run_multi(const Tensor &w, const Tensor &sigz, const Tensor &sigy, at::TensorList wave,
                                     at::TensorList inj_start, at::TensorList inj_cell, at::TensorList inj_col,
                                     at::TensorList inj_amp, at::TensorList smp_cell, const c10::optional<Tensor> &q_write,
                                     const c10::optional<Tensor> &q_read, const c10::optional<Tensor> &image,
                                     at::IntArrayRef shot0, const Tensor &meta, int64_t n_total, int64_t nsteps, bool adjoint,
                                     int64_t order, int64_t pml, double dz, double dy, int64_t chunk_rows, int64_t save_every) {
    const size_t J = wave.size();
    TORCH_CHECK(J >= 1 && J <= 16 && inj_start.size() == J && inj_cell.size() == J && inj_col.size() == J &&
                inj_amp.size() == J && smp_cell.size() == J && shot0.size() == J, "one entry per shot group expected");
    TORCH_CHECK(w.is_cuda(), "scalar_run_multi needs CUDA tensors: there is no CPU path");
    const double *m = meta.data_ptr<double>();
    const int64_t stride = (int64_t)m[8];
    c10::cuda::CUDAGuard guard(w.device());
    auto main = c10::cuda::getCurrentCUDAStream(w.device().index());
    std::vector<c10::cuda::CUDAStream> streams;
    streams.push_back(main);
    for (size_t j = 1; j < J; ++j) streams.push_back(c10::cuda::getStreamFromPool(false, w.device().index()));
    cudaEvent_t ready;
    C10_CUDA_CHECK(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
    C10_CUDA_CHECK(cudaEventRecord(ready, main.stream()));
    std::vector<fwi_acoustic_desc> descs(J);
    std::vector<fwi_acoustic_buffers> bufs(J);
    std::vector<Tensor> out(J), tables(J);
    std::vector<const fwi_acoustic_desc *> dp(J);
    std::vector<const fwi_acoustic_buffers *> bp(J);
    std::vector<void *> sp(J);
    Tensor qw = q_write.has_value() ? *q_write : Tensor(), im = image.has_value() ? *image : Tensor();
    for (size_t j = 0; j < J; ++j) {
        const int64_t n_j = wave[j].numel() / ((int64_t)m[7] * stride);
        TORCH_CHECK(n_j >= 1 && wave[j].numel() == (int64_t)m[7] * n_j * stride, "wave workspace of group ", j, " has the wrong size");
        TORCH_CHECK(shot0[j] >= 0 && shot0[j] + n_j <= n_total, "shot range of group ", j, " outside the gradient's shots");
        fwi_acoustic_desc &d = descs[j];
        d = fwi_acoustic_desc{};
        d.nzp = (int32_t)m[5]; d.nyp = (int32_t)m[6]; d.order = (int32_t)order; d.pml_width = (int32_t)pml;
        d.n_shots = (int32_t)n_j; d.chunk_rows = (int32_t)chunk_rows; d.nsteps = (int32_t)nsteps; d.adjoint = adjoint ? 1 : 0;
        d.n_inj = inj_amp[j].numel() ? (int32_t)inj_amp[j].size(2) : 0;
        d.inj_scale_a = d.smp_div_a = adjoint ? 1 : 0;
        d.n_smp = smp_cell[j].numel() ? (int32_t)smp_cell[j].size(1) : 0;
        d.dz = dz; d.dy = dy; d.dt = m[2];
        d.save_every = (int32_t)save_every;
        // programmatic dependent launches stay on (measured on C2, two adjoint groups: 46.9 us per step with, 48.6 without;
        // FWI_B200_MULTI_PDL=0 switches them off for the groups)
        static const bool multi_pdl = !(getenv("FWI_B200_MULTI_PDL") && atoi(getenv("FWI_B200_MULTI_PDL")) == 0);
        d.flags = multi_pdl ? 0 : FWI_FLAG_NO_PDL;
        TORCH_CHECK(d.n_inj == 0 || (inj_amp[j].is_contiguous() && inj_amp[j].size(0) == nsteps && inj_amp[j].size(1) == n_j),
                    "injection amplitudes of group ", j, " must be contiguous [nsteps, n_shots, n]");
        out[j] = at::empty({nsteps, n_j, (int64_t)d.n_smp}, w.options());
        tables[j] = at::empty({(int64_t)m[9]}, w.options().dtype(at::kByte));
        fwi_acoustic_buffers &b = bufs[j];
        b = fwi_acoustic_buffers{};
        b.w = w.data_ptr<float>(); b.sigz = fptr(sigz); b.sigy = fptr(sigy); b.wave = wave[j].data_ptr<float>();
        b.inj_row_start = iptr(inj_start[j]); b.inj_cell = iptr(inj_cell[j]); b.inj_col = iptr(inj_col[j]);
        b.inj_amp = fptr(inj_amp[j]); b.smp_cell = iptr(smp_cell[j]);
        b.smp_out = out[j].numel() ? out[j].data_ptr<float>() : nullptr;
        b.q_write = qw.defined() && qw.numel() ? qw.data_ptr<float>() + shot0[j] * stride : nullptr;
        b.q_read = q_read.has_value() && q_read->numel() ? q_read->data_ptr<float>() + shot0[j] * stride : nullptr;
        b.image = im.defined() && im.numel() ? im.data_ptr<float>() + shot0[j] * stride : nullptr;
        b.q_step_stride = (b.q_write || b.q_read) ? n_total * stride : 0;
        b.unit_table = tables[j].data_ptr();
        dp[j] = &d; bp[j] = &b; sp[j] = static_cast<void *>(streams[j].stream());
        if (j > 0) C10_CUDA_CHECK(cudaStreamWaitEvent(streams[j].stream(), ready, 0));
    }
    const int rc = fwi_acoustic_run_multi((int32_t)J, dp.data(), bp.data(), sp.data());
    for (size_t j = 1; j < J; ++j) {                 // the current stream continues when every group is done
        C10_CUDA_CHECK(cudaEventRecord(ready, streams[j].stream()));
        C10_CUDA_CHECK(cudaStreamWaitEvent(main.stream(), ready, 0));
    }
    C10_CUDA_CHECK(cudaEventDestroy(ready));
    check_rc(rc, "fwi_acoustic_run_multi");
    return out;
}

// ---- image -> dJ/dv -----------------------------------------------------------------------------------------------
Tensor scalar_fold(const Tensor &image, const c10::optional<Tensor> &gs, const c10::optional<Tensor> &amp, const Tensor &src_cell,
                   const Tensor &v, const Tensor &meta, int64_t n_shots, int64_t order, int64_t pml, double dz, double dy) {
    check_model(v);
    const double *m = meta.data_ptr<double>();
    c10::cuda::CUDAGuard guard(v.device());
    const Tensor vc = v.detach().contiguous();
    const fwi_model_desc md = model_desc(vc, order, pml, dz, dy, m[2], m[0], 0, 0);
    const int32_t n_src = src_cell.defined() && src_cell.numel() ? (int32_t)src_cell.size(1) : 0;
    const int32_t nsteps = gs.has_value() && gs->numel() ? (int32_t)gs->size(0) : 0;
    Tensor ws = at::empty({fwi_acoustic_fold_workspace_floats(&md, (int32_t)n_shots, n_src)}, vc.options());
    Tensor out = at::empty_like(vc);
    check_rc(fwi_acoustic_fold_gradient(&md, (int32_t)n_shots, image.data_ptr<float>(), nsteps ? fptr(*gs) : nullptr,
                                        nsteps ? fptr(*amp) : nullptr, iptr(src_cell), n_src, nsteps, vc.data_ptr<float>(),
                                        ws.data_ptr<float>(), out.data_ptr<float>(), stream_of(vc)),
             "fwi_acoustic_fold_gradient");
    return out;
}

// ---- out-of-model flag: raise like deepwave does, without waiting for the time loop queued after the planning -------
// begin: the flag travels to pinned memory behind the planning kernels (an event marks the point); end: wait for that
// event only and raise.  Both are no-ops while the stream is being captured into a CUDA graph (a sync is illegal there;
// the caller validates the geometry in its eager warm-up iteration).
Tensor flag_begin(const Tensor &bad) {
    cudaStream_t st = static_cast<cudaStream_t>(stream_of(bad));
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(st, &cap);
    Tensor host = at::zeros({3}, at::TensorOptions().dtype(at::kLong).pinned_memory(cap == cudaStreamCaptureStatusNone));
    if (cap != cudaStreamCaptureStatusNone) return host;
    c10::cuda::CUDAGuard guard(bad.device());
    int64_t *h = host.data_ptr<int64_t>();
    cudaEvent_t ev;
    C10_CUDA_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    C10_CUDA_CHECK(cudaMemcpyAsync(h, bad.data_ptr<int32_t>(), sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    C10_CUDA_CHECK(cudaEventRecord(ev, st));
    h[1] = reinterpret_cast<int64_t>(ev);
    h[2] = 1;
    return host;
}

void flag_end(Tensor host) {
    int64_t *h = host.data_ptr<int64_t>();
    if (h[2] != 1) return;
    cudaEvent_t ev = reinterpret_cast<cudaEvent_t>(h[1]);
    h[2] = 0;
    C10_CUDA_CHECK(cudaEventSynchronize(ev));
    C10_CUDA_CHECK(cudaEventDestroy(ev));
    const int32_t flag = *reinterpret_cast<const int32_t *>(h);
    TORCH_CHECK(flag == 0, "source/receiver location outside the model");
}

int64_t abi_version() { return fwi_version(); }

}  // namespace

TORCH_LIBRARY(b200fwi, m) {
    m.def("scalar_plan(Tensor v, float dt, float dz, float dy, int order, int pml, float vpmax, int n_shots, int chunk_rows, "
          "int c0, int c1) -> Tensor");
    m.def("scalar_prepare(Tensor v, Tensor meta, int order, int pml, float dz, float dy, int c0, int c1) -> (Tensor, Tensor, Tensor)");
    m.def("scalar_plan_points(Tensor x, Tensor v, Tensor meta, Tensor(a!) bad, int order, int pml, float dz, float dy, "
          "bool loc_round, int c0, int c1) -> (Tensor, Tensor, Tensor, Tensor)");
    m.def("scalar_run(Tensor w, Tensor sigz, Tensor sigy, Tensor(a!) wave, Tensor inj_start, Tensor inj_cell, Tensor inj_col, "
          "Tensor inj_amp, Tensor smp_cell, Tensor(b!)? q_write, Tensor? q_read, Tensor(c!)? image, Tensor meta, int n_shots, "
          "int nsteps, bool adjoint, int order, int pml, float dz, float dy, int chunk_rows, int q_step_stride=0, int save_every=1) -> Tensor");
    m.def("scalar_run_multi(Tensor w, Tensor sigz, Tensor sigy, Tensor(a!)[] wave, Tensor[] inj_start, Tensor[] inj_cell, "
          "Tensor[] inj_col, Tensor[] inj_amp, Tensor[] smp_cell, Tensor(b!)? q_write, Tensor? q_read, Tensor(c!)? image, "
          "int[] shot0, Tensor meta, int n_total, int nsteps, bool adjoint, int order, int pml, float dz, float dy, "
          "int chunk_rows, int save_every=1) -> Tensor[]");
    m.def("scalar_fold(Tensor image, Tensor? gs, Tensor? amp, Tensor src_cell, Tensor v, Tensor meta, int n_shots, int order, "
          "int pml, float dz, float dy) -> Tensor");
    m.def("flag_begin(Tensor bad) -> Tensor");
    m.def("flag_end(Tensor(a!) host) -> ()");
    m.def("abi_version() -> int");
}

// one catch-all registration: the operators do their own device checks (a CPU tensor gets the "no CPU path" message)
TORCH_LIBRARY_IMPL(b200fwi, CompositeExplicitAutograd, m) {
    m.impl("scalar_plan", scalar_plan);
    m.impl("scalar_prepare", scalar_prepare);
    m.impl("scalar_plan_points", scalar_plan_points);
    m.impl("scalar_run", scalar_run);
    m.impl("scalar_run_multi", scalar_run_multi);
    m.impl("scalar_fold", scalar_fold);
    m.impl("flag_begin", flag_begin);
    m.impl("flag_end", flag_end);
    m.impl("abi_version", abi_version);
}
