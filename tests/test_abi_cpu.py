"""The C-ABI library builds for sm_100a, loads without a GPU, and exports every symbol that
include/fwi_b200.h declares.  No compute call is made here (there is no CPU path to call)."""
import ctypes
import os
import re

import pytest

import physicsbasedfwi2_b200 as pkg
from physicsbasedfwi2_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "fwi_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fwi_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def so():
    return ctypes.CDLL(pkg.build())


def test_header_declares_entry_points():
    syms = _declared_symbols()
    for must in ("fwi_acoustic_get_layout", "fwi_acoustic_forward", "fwi_acoustic_backward",
                 "fwi_elastic_forward", "fwi_elastic_backward", "fwi_elastic_wave_floats", "fwi_acoustic_plan_stats",
                 "fwi_last_error", "fwi_version", "fwi_kernel_launches", "fwi_device_ok"):
        assert must in syms


def test_library_exports_every_declared_symbol(so):
    for name in _declared_symbols():
        assert hasattr(so, name), f"{name} declared in include/fwi_b200.h but not exported"


def test_version_and_error_string(so):
    so.fwi_last_error.restype = ctypes.c_char_p
    assert so.fwi_version() >= 5
    assert so.fwi_last_error() is not None


def test_layout_is_pure_host_code_and_checks_arguments():
    L = pkg.lib()
    d = _lib.AcousticDesc(124, 324, 4, 10, 10, 0, 0, 0, 0, 0, 0, 0, 10.0, 10.0, 1e-3)
    lay = _lib.AcousticLayout()
    assert L.fwi_acoustic_get_layout(ctypes.byref(d), ctypes.byref(lay)) == 0
    assert lay.pitch % 32 == 0 and lay.pitch >= lay.col0 + 324
    assert lay.shot_stride == 124 * lay.pitch
    assert (lay.col0 + 2 * 2 + 10) % 32 == 0          # first interior column is 128-byte aligned
    assert lay.n_images == 10 and lay.wave_fields == 6
    assert lay.chunk_rows == 0 and lay.units_per_shot >= 1 and lay.unit_table_bytes > 0
    nopml = _lib.AcousticDesc(104, 304, 8, 0, 3, 0, 0, 0, 0, 0, 0, 0, 10.0, 10.0, 1e-3)
    assert L.fwi_acoustic_get_layout(ctypes.byref(nopml), ctypes.byref(lay)) == 0
    assert lay.wave_fields == 2 and (lay.col0 + 4) % 32 == 0
    bad = _lib.AcousticDesc(124, 324, 5, 10, 10, 0, 0, 0, 0, 0, 0, 0, 10.0, 10.0, 1e-3)
    assert L.fwi_acoustic_get_layout(ctypes.byref(bad), ctypes.byref(lay)) == -1
    assert b"order" in L.fwi_last_error()


def test_ctypes_structs_match_header_field_order():
    src = open(os.path.join(ROOT, "include", "fwi_b200.h")).read()
    for struct, cls in (("fwi_acoustic_desc", _lib.AcousticDesc), ("fwi_acoustic_layout", _lib.AcousticLayout),
                        ("fwi_acoustic_buffers", _lib.AcousticBuffers), ("fwi_elastic_desc", _lib.ElasticDesc),
                        ("fwi_elastic_buffers", _lib.ElasticBuffers)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (struct, struct), src, flags=re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            for part in decl.split(","):
                names.append(re.findall(r"[A-Za-z_][A-Za-z0-9_]*", part)[-1])
        assert names == [f[0] for f in cls._fields_], struct


def test_product_refuses_cpu_tensors():
    import torch
    from physicsbasedfwi2_b200.scalar import Propagator
    v = torch.full((20, 20), 1500.0)
    prop = Propagator({"vp": v}, 10.0)
    with pytest.raises(RuntimeError, match="no CPU path"):
        prop(torch.zeros(10, 1, 1), torch.zeros(1, 1, 2), torch.zeros(1, 3, 2), 1e-3)


def test_product_does_not_import_oracle():
    import subprocess
    import sys
    code = ("import sys; import physicsbasedfwi2_b200, physicsbasedfwi2_b200.scalar, "
            "physicsbasedfwi2_b200.distributed; "
            "assert not any(m == 'oracle' or m.startswith('oracle.') for m in sys.modules), 'oracle imported'")
    subprocess.check_call([sys.executable, "-c", code], cwd=ROOT)
    for dirpath, _, files in os.walk(os.path.join(ROOT, "physicsbasedfwi2_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, f


@pytest.mark.parametrize("shape", [
    (799, 2349, 8, 20, 6, 0), (799, 2349, 8, 20, 1, 0), (124, 324, 4, 10, 10, 0), (124, 324, 4, 10, 10, 2),
    (58, 139, 4, 0, 2, 0), (40, 64, 8, 6, 2, 0), (30, 30, 8, 10, 1, 0), (1549, 3549, 8, 20, 2, 7),
])
def test_work_table_tiles_the_grid_and_is_balanced(shape):
    """Host-only planner: the segments cover every computational cell exactly once (checked inside
    the library by area) and the heaviest warp unit is within 10 % of the mean."""
    L = pkg.lib()
    L.fwi_acoustic_plan_stats.argtypes = [ctypes.POINTER(_lib.AcousticDesc), ctypes.c_int,
                                          ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int32),
                                          ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
    nzp, nyp, order, pml, S, chunk = shape
    d = _lib.AcousticDesc(nzp, nyp, order, pml, S, chunk, 0, 0, 0, 0, 0, 0, 10.0, 10.0, 1e-3)
    ns, nu, mx, mean = ctypes.c_int64(), ctypes.c_int32(), ctypes.c_double(), ctypes.c_double()
    rc = L.fwi_acoustic_plan_stats(ctypes.byref(d), 148 * 16, ctypes.byref(ns), ctypes.byref(nu),
                                   ctypes.byref(mx), ctypes.byref(mean))
    assert rc == 0, L.fwi_last_error()
    assert 1 <= nu.value <= 148 * 16 and ns.value >= nu.value
    if chunk == 0:
        assert mx.value <= 1.1 * mean.value + 6.0


def test_misfit_and_slab_entry_points_check_arguments_on_the_host():
    L = pkg.lib()
    d = _lib.MisfitDesc(3000, 6 * 576, 0, 1, 1e-10)
    ws = L.fwi_trace_misfit_workspace_bytes(ctypes.byref(d))
    assert ws > 0 and ws % (16 * 6 * 576) == 0          # 4 partial arrays of [C][M] words
    assert L.fwi_trace_misfit_workspace_bytes(ctypes.byref(_lib.MisfitDesc(0, 5, 0, 1, 1e-10))) == -1
    bad = _lib.MisfitDesc(10, 5, 2, 1, 1e-10)
    assert L.fwi_trace_misfit(ctypes.byref(bad), 1, None, 1, 1, 1, None) == -1
    assert b"kind" in L.fwi_last_error()
    big = _lib.MisfitDesc(4000, 400000, 0, 1, 1e-10)          # 6.4 MB of partials: beyond the static array of the no-alloc variant
    assert L.fwi_trace_misfit_noalloc(ctypes.byref(big), 256, None, 256, 256, None) == -1
    assert b"4194304" in L.fwi_last_error()
    assert L.fwi_slab_exchanges(8000, 8) == 999 and L.fwi_slab_exchanges(8, 8) == 0 and L.fwi_slab_exchanges(9, 8) == 1
    assert L.fwi_slab_exchanges(0, 8) == 0
    ad = _lib.AcousticDesc(124, 324, 4, 10, 1, 0, 10, 0, 0, 0, 0, 0, 10.0, 10.0, 1e-3)
    link = _lib.SlabLink(8, 15, 1, 0, 0, 0, 0, 0, None, None, None, 0, 0)      # ghost_cols != 8 * 2
    assert L.fwi_acoustic_run_slab(ctypes.byref(ad), ctypes.byref(_lib.AcousticBuffers()), ctypes.byref(link), None) == -1
    assert b"ghost_cols" in L.fwi_last_error()


def test_torch_library_extension_loads_and_registers_every_operator():
    """csrc/torch_ext.cpp: TORCH_LIBRARY(b200fwi) - the PyTorch C++ operators over the C ABI (no compute without a GPU)."""
    import torch
    from physicsbasedfwi2_b200 import _lib
    ops = _lib.ops()
    for name in ("scalar_plan", "scalar_prepare", "scalar_plan_points", "scalar_run", "scalar_fold", "flag_begin", "flag_end",
                 "abi_version"):
        assert hasattr(ops, name), name
    assert int(ops.abi_version()) == int(_lib.lib().fwi_version())
    with pytest.raises(RuntimeError, match="CUDA tensor"):           # TORCH_CHECK -> RuntimeError, like deepwave
        ops.scalar_plan(torch.zeros(4, 4), 1e-3, 10.0, 10.0, 4, 10, -1.0, 1, 0, 0, 0)
    with pytest.raises(RuntimeError, match="2-D"):
        ops.scalar_plan(torch.zeros(4), 1e-3, 10.0, 10.0, 4, 10, -1.0, 1, 0, 0, 0)


def test_no_wavefield_load_is_scheduled_ahead_of_the_dependent_launch_wait():
    """Kernels of the time loops start under programmatic dependent launch and wait (griddepcontrol.wait, SASS ACQBULK)
    before touching a wavefield.  ld.global.nc loads may be moved above that wait by the compiler, so wavefields must be
    read with plain loads (csrc/elastic.cu, csrc/acoustic.cu:ld_field); what sits ahead of the wait in the SASS may only
    be LDG.E.CONSTANT (__ldg of tables and material grids that no kernel of a time loop writes), never a plain LDG, a
    bulk-tensor load or a store."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    src = open(os.path.join(ROOT, "physicsbasedfwi2_b200", "csrc", "elastic.cu")).read()
    assert "__ldg(p.f[" not in src and "__ldg(f +" not in src, "elastic.cu reads a wavefield with __ldg"
    sass = subprocess.run([cuobjdump, "-sass", pkg.build()], capture_output=True, text=True, check=True).stdout
    name, before_wait, waits, bad = None, [], 0, []
    for line in sass.splitlines():
        if "Function :" in line:
            name, before_wait = line.split(":", 1)[1].strip(), []
        elif name and "ACQBULK" in line:
            waits += 1
            bad += [(name, ins) for ins in before_wait]
            name = None                                   # only what precedes the first wait matters
        elif name:
            m = re.search(r"\b(LDG\.[A-Z0-9.]*|STG\.[A-Z0-9.]*|UTMALDG[A-Z0-9.]*|RED\.[A-Z0-9.]*|ATOMG?\.[A-Z0-9.]*)", line)
            if m and ".CONSTANT" not in m.group(1):
                before_wait.append(m.group(1))
    assert waits >= 20, f"only {waits} kernels with a dependent-launch wait found: did the SASS format change?"
    assert not bad, f"memory operations ahead of griddepcontrol.wait: {bad[:5]}"
