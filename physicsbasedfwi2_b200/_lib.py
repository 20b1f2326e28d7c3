"""ctypes binding of libfwi_b200.so (the C ABI declared in include/fwi_b200.h).

The product path has no CPU fallback: if the library is missing or a call fails, a
RuntimeError is raised (deepwave raises RuntimeError for bad arguments too).
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(_HERE, "csrc")
# (FWI_B200_LIB: a tuning variant for the ctypes entry points - elastic, misfit, slab, glue; _b200fwi_torch.so, i.e. the
# scalar propagator, is linked against the in-tree file whatever this says: scripts/dev_variants.sh)
SO_PATH = os.environ.get("FWI_B200_LIB") or os.path.join(_HERE, "libfwi_b200.so")
_SOURCES = ["common.cu", "acoustic.cu", "acoustic_small.cu", "elastic.cu", "misfit.cu", "host_kernels.cu", "glue.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-fmad=false", "-Xcompiler", "-fPIC", "-shared"]


class AcousticDesc(C.Structure):
    _fields_ = [("nzp", C.c_int32), ("nyp", C.c_int32), ("order", C.c_int32),
                ("pml_width", C.c_int32), ("n_shots", C.c_int32), ("chunk_rows", C.c_int32),
                ("nsteps", C.c_int32), ("adjoint", C.c_int32), ("n_inj", C.c_int32),
                ("inj_scale_a", C.c_int32), ("n_smp", C.c_int32), ("smp_div_a", C.c_int32),
                ("dz", C.c_double), ("dy", C.c_double), ("dt", C.c_double), ("flags", C.c_int32), ("save_every", C.c_int32)]


class AcousticLayout(C.Structure):
    _fields_ = [("pitch", C.c_int32), ("col0", C.c_int32), ("shot_stride", C.c_int64),
                ("n_images", C.c_int32), ("wave_fields", C.c_int32), ("chunk_rows", C.c_int32),
                ("units_per_shot", C.c_int32), ("unit_table_bytes", C.c_int64)]


class AcousticBuffers(C.Structure):
    _fields_ = [("w", C.c_void_p), ("sigz", C.c_void_p), ("sigy", C.c_void_p),
                ("wave", C.c_void_p), ("inj_row_start", C.c_void_p), ("inj_cell", C.c_void_p),
                ("inj_col", C.c_void_p), ("inj_amp", C.c_void_p), ("smp_cell", C.c_void_p),
                ("smp_out", C.c_void_p), ("q_write", C.c_void_p), ("q_read", C.c_void_p),
                ("image", C.c_void_p), ("unit_table", C.c_void_p), ("q_step_stride", C.c_int64)]


class ModelDesc(C.Structure):
    _fields_ = [("nz", C.c_int32), ("ny", C.c_int32), ("order", C.c_int32), ("pml_width", C.c_int32),
                ("c0", C.c_int32), ("c1", C.c_int32), ("dz", C.c_double), ("dy", C.c_double), ("dt", C.c_double),
                ("vmax", C.c_double)]


class SlabLink(C.Structure):
    _fields_ = [("halo_steps", C.c_int32), ("ghost_cols", C.c_int32), ("has_left", C.c_int32),
                ("has_right", C.c_int32), ("send_left_col", C.c_int32), ("send_right_col", C.c_int32),
                ("ghost_left_col", C.c_int32), ("ghost_right_col", C.c_int32), ("comm", C.c_void_p),
                ("peer_left", C.c_void_p), ("peer_right", C.c_void_p), ("epoch", C.c_uint32),
                ("timeout_ms", C.c_uint32)]


SLAB_HEADER_BYTES = 256


class MisfitDesc(C.Structure):
    _fields_ = [("nt", C.c_int32), ("n_traces", C.c_int32), ("kind", C.c_int32), ("normalize", C.c_int32),
                ("eps", C.c_float)]


class ButterDesc(C.Structure):
    _fields_ = [("nt", C.c_int32), ("n_traces", C.c_int32), ("order", C.c_int32), ("zero_phase", C.c_int32),
                ("dt", C.c_double), ("fc_low", C.c_double), ("fc_high", C.c_double)]


class GradPostDesc(C.Structure):
    _fields_ = [("nz", C.c_int32), ("ny", C.c_int32), ("water_rows", C.c_int32), ("depth_pow", C.c_double),
                ("rho_factor", C.c_double), ("rho_offset", C.c_double)]


class ElasticDesc(C.Structure):
    _fields_ = [("nzp", C.c_int32), ("nyp", C.c_int32), ("order", C.c_int32), ("pml_width", C.c_int32),
                ("n_shots", C.c_int32), ("nsteps", C.c_int32), ("n_src", C.c_int32), ("n_rec", C.c_int32),
                ("dz", C.c_double), ("dy", C.c_double), ("dt", C.c_double), ("source_kind", C.c_int32),
                ("free_surface", C.c_int32), ("adj_parity", C.c_int32)]


class ElasticBuffers(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("l2m", "lam", "mu", "by", "bz", "prof_z", "prof_y", "wave", "src_cell", "src_link",
                                          "src_amp", "rec_cell", "rec_link", "rec_out", "snap", "scratch", "grad", "grad_rec",
                                          "grad_src")]


def sources():
    return [os.path.join(_CSRC, s) for s in _SOURCES if os.path.exists(os.path.join(_CSRC, s))]


def is_stale() -> bool:
    if not os.path.exists(SO_PATH):
        return True
    t = os.path.getmtime(SO_PATH)
    deps = sources() + [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".h", ".cuh"))]
    deps.append(os.path.join(_HERE, "..", "include", "fwi_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a into the in-tree libfwi_b200.so (nvcc cross-compiles
    without a GPU)."""
    if not force and not is_stale():
        build_ext()
        return SO_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build libfwi_b200.so")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", SO_PATH] + sources()
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    build_ext(force=True)
    return SO_PATH


EXT_PATH = os.path.join(_HERE, "_b200fwi_torch.so")
_EXT_SRC = os.path.join(_CSRC, "torch_ext.cpp")


def ext_is_stale() -> bool:
    if not os.path.exists(EXT_PATH):
        return True
    t = os.path.getmtime(EXT_PATH)
    return any(os.path.getmtime(d) > t for d in (_EXT_SRC, os.path.join(_HERE, "..", "include", "fwi_b200.h")))


def build_ext(force: bool = False) -> str:
    """Compile csrc/torch_ext.cpp (TORCH_LIBRARY(b200fwi)) into the in-tree _b200fwi_torch.so, linked against
    libfwi_b200.so next to it ($ORIGIN rpath) and the torch libraries of this interpreter."""
    if not force and not ext_is_stale():
        return EXT_PATH
    import sysconfig
    import torch
    from torch.utils import cpp_extension as CE
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else (shutil.which("g++") or "g++")
    cuda_home = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    inc = [f"-I{p}" for p in CE.include_paths()] + [f"-I{cuda_home}/include", f"-I{sysconfig.get_paths()['include']}"]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           "-DTORCH_EXTENSION_NAME=_b200fwi_torch", *inc, _EXT_SRC, "-o", EXT_PATH,
           f"-L{_HERE}", "-l:libfwi_b200.so", f"-L{tlib}", "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_cuda", "-lc10_cuda",
           f"-L{cuda_home}/lib64", "-lcudart", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tlib}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building the TORCH_LIBRARY extension failed:\n" + r.stdout[-3000:] + r.stderr[-3000:])
    return EXT_PATH


_ops = None


def ops():
    """torch.ops.b200fwi (loads _b200fwi_torch.so).  Never falls back to anything else."""
    global _ops
    if _ops is None:
        import torch
        if not os.path.exists(EXT_PATH):
            raise RuntimeError(f"{EXT_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU or PyTorch fallback for the propagator)")
        lib()                                  # libfwi_b200.so first: the extension links against it
        torch.ops.load_library(EXT_PATH)
        _ops = torch.ops.b200fwi
        if int(_ops.abi_version()) != int(lib().fwi_version()):
            raise RuntimeError("_b200fwi_torch.so and libfwi_b200.so disagree on the ABI version: rebuild")
    return _ops


_lib = None


def lib() -> C.CDLL:
    """Loads libfwi_b200.so.  Never falls back to anything else."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise RuntimeError(
            f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU or PyTorch fallback for the propagator)")
    L = C.CDLL(SO_PATH)
    L.fwi_last_error.restype = C.c_char_p
    L.fwi_version.restype = C.c_int
    L.fwi_kernel_launches.restype = C.c_int64
    L.fwi_device_ok.argtypes = [C.c_int]
    L.fwi_acoustic_get_layout.argtypes = [C.POINTER(AcousticDesc), C.POINTER(AcousticLayout)]
    for name in ("fwi_acoustic_run", "fwi_acoustic_forward", "fwi_acoustic_backward"):
        getattr(L, name).argtypes = [C.POINTER(AcousticDesc), C.POINTER(AcousticBuffers), C.c_void_p]
        getattr(L, name).restype = C.c_int
    L.fwi_acoustic_run_slab.argtypes = [C.POINTER(AcousticDesc), C.POINTER(AcousticBuffers), C.POINTER(SlabLink),
                                        C.c_void_p]
    L.fwi_acoustic_run_slab.restype = C.c_int
    L.fwi_peer_alloc.argtypes = [C.c_int64, C.POINTER(C.c_void_p), C.c_char_p]
    L.fwi_peer_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    L.fwi_peer_close.argtypes = [C.c_void_p]
    L.fwi_peer_free.argtypes = [C.c_void_p]
    L.fwi_slab_status.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32)]
    L.fwi_slab_exchanges.argtypes = [C.c_int32, C.c_int32]
    L.fwi_slab_exchanges.restype = C.c_int32
    L.fwi_acoustic_prepare_model.argtypes = [C.POINTER(ModelDesc)] + [C.c_void_p] * 5
    L.fwi_acoustic_prepare_model.restype = C.c_int
    L.fwi_acoustic_plan_points.argtypes = [C.POINTER(ModelDesc), C.c_void_p, C.c_int32, C.c_int32, C.c_int32] + [C.c_void_p] * 6
    L.fwi_acoustic_plan_points.restype = C.c_int
    L.fwi_acoustic_fold_workspace_floats.argtypes = [C.POINTER(ModelDesc), C.c_int32, C.c_int32]
    L.fwi_acoustic_fold_workspace_floats.restype = C.c_int64
    L.fwi_acoustic_fold_gradient.argtypes = ([C.POINTER(ModelDesc), C.c_int32] + [C.c_void_p] * 4 + [C.c_int32, C.c_int32]
                                             + [C.c_void_p] * 4)
    L.fwi_acoustic_fold_gradient.restype = C.c_int
    L.fwi_trace_misfit_workspace_bytes.argtypes = [C.POINTER(MisfitDesc)]
    L.fwi_trace_misfit_workspace_bytes.restype = C.c_int64
    L.fwi_trace_misfit.argtypes = [C.POINTER(MisfitDesc)] + [C.c_void_p] * 6
    L.fwi_trace_misfit.restype = C.c_int
    L.fwi_trace_misfit_backward.argtypes = [C.POINTER(MisfitDesc)] + [C.c_void_p] * 7
    L.fwi_trace_misfit_backward.restype = C.c_int
    L.fwi_trace_misfit_noalloc.argtypes = [C.POINTER(MisfitDesc)] + [C.c_void_p] * 5
    L.fwi_trace_misfit_noalloc.restype = C.c_int
    L.fwi_butterworth_lowpass.argtypes = [C.POINTER(ButterDesc), C.c_void_p, C.c_void_p, C.c_void_p]
    L.fwi_butterworth_lowpass.restype = C.c_int
    L.fwi_elastic_gradient_post.argtypes = [C.POINTER(GradPostDesc)] + [C.c_void_p] * 8
    L.fwi_elastic_gradient_post.restype = C.c_int
    L.fwi_elastic_wave_floats.argtypes = [C.POINTER(ElasticDesc)]
    L.fwi_elastic_wave_floats.restype = C.c_int64
    for name in ("fwi_elastic_forward", "fwi_elastic_backward"):
        getattr(L, name).argtypes = [C.POINTER(ElasticDesc), C.POINTER(ElasticBuffers), C.c_void_p]
        getattr(L, name).restype = C.c_int
    _lib = L
    return L


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().fwi_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (status {rc}): {msg}")


def kernel_launches() -> int:
    return int(lib().fwi_kernel_launches())


# ---- optional time-loop timing (bench.py's roofline leg) ---------------------------------------
# When LOOP_TIMES is a list, every native time loop is bracketed by CUDA events recorded on the
# stream the kernels are launched on; entries are (tag, start_event, end_event, n_launches, n_shots).
LOOP_TIMES = None


def time_loop_begin(stream):
    if LOOP_TIMES is None:
        return None
    import torch
    ev = torch.cuda.Event(enable_timing=True)
    ev.record(stream)
    return ev


def time_loop_end(stream, ev, tag, n_launches, n_shots=1):
    if ev is None:
        return
    import torch
    end = torch.cuda.Event(enable_timing=True)
    end.record(stream)
    LOOP_TIMES.append((tag, ev, end, int(n_launches), int(n_shots)))
