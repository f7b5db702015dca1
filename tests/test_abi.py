"""CPU tests of the drop-in boundary: libpcb200.so loads without a GPU and exports every symbol
that include/pcb200.h declares (plain + the reference's trailing-underscore FTNIZE names).
No compute calls are made here."""
import ctypes as C
import os
import re

from pencil_code_b200 import gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "pcb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"^\s*(?:void|int|long long|void \*)\s*\*?\s*([a-z_0-9]+)\s*\(", src, flags=re.M)
    return sorted(set(names))


def test_library_loads_and_exports_header_symbols():
    L = gpu.load_library()
    decl = _declared_functions()
    assert len(decl) >= 35, decl
    missing = [n for n in decl if not hasattr(L, n)]
    assert not missing, missing
    # the python mirror's list is the same set
    assert sorted(set(gpu.ABI_SYMBOLS)) == decl


def test_fortran_underscore_aliases():
    L = gpu.load_library()
    missing = [n for n in gpu.FTN_SYMBOLS if not hasattr(L, n)]
    assert not missing, missing


def test_config_struct_layout():
    L = gpu.load_library()
    assert L.pcb200_sizeof_config() == C.sizeof(gpu.Config)


def test_module_gpu_public_names():
    """every public procedure of src/gpu.h:1-12 that the hot path needs exists on the host mirror"""
    for name in ("register_GPU", "initialize_GPU", "finalize_GPU", "get_farray_ptr_gpu", "rhs_GPU",
                 "copy_farray_from_GPU", "load_farray_to_GPU", "reload_GPU_config", "update_on_gpu",
                 "before_boundary_gpu", "update_after_substep_gpu", "gpu_prepare_for_first_substep",
                 "get_gpu_reduced_vars", "test_gpu_bcs", "split_update_gpu"):
        assert hasattr(gpu.GPU, name), name
    assert gpu.GPU.n_gpu_reduced_vars == 20


def test_fortran_module_lists_gpu_h_names():
    """pencil_code_b200/fortran/gpu_b200.f90 must define every public name of gpu.h"""
    path = os.path.join(ROOT, "pencil_code_b200", "fortran", "gpu_b200.f90")
    src = open(path).read().lower()
    assert "lgpu = .true." in src and "include 'gpu.h'" in src
    for name in ("register_gpu", "initialize_gpu", "finalize_gpu", "get_farray_ptr_gpu", "rhs_gpu",
                 "copy_farray_from_gpu", "read_gpu_run_pars", "write_gpu_run_pars", "load_farray_to_gpu",
                 "reload_gpu_config", "update_on_gpu", "get_ptr_gpu", "get_ptr_gpu_training", "before_boundary_gpu",
                 "update_on_gpu_vec", "update_after_substep_gpu", "gpu_prepare_for_first_substep", "radtransfer_gpu",
                 "get_gpu_reduced_vars", "test_gpu_bcs", "split_update_gpu", "pushpars2c", "tf_train_gpu",
                 "tf_infer_gpu", "tf_create_model", "tau_snapshots", "tf_load_model", "tf_load_model_checkpoint",
                 "tf_save_model", "tf_save_checkpoint"):
        assert re.search(r"(subroutine|function)\s+%s\b" % name, src), name


def test_start_stub_resolves_every_bound_name(tmp_path):
    """nogpu_b200_ansi.c (linked into start.x, Makefile.src:3139) compiles against include/pcb200.h and
    defines every C name gpu_b200.f90 binds; each of those names is also exported by libpcb200.so or by the
    in-tree shim gpu_b200_ansi.c."""
    import subprocess
    fdir = os.path.join(ROOT, "pencil_code_b200", "fortran")
    obj = str(tmp_path / "stub.o")
    subprocess.check_call(["gcc", "-c", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(fdir, "nogpu_b200_ansi.c"), "-o", obj])
    nm = subprocess.run(["nm", obj], capture_output=True, text=True, check=True).stdout
    stub = {l.split()[2] for l in nm.splitlines() if len(l.split()) == 3 and l.split()[1] == "T"}
    bound = set(re.findall(r'bind\(c,\s*name="([a-z0-9_]+)"\)', open(os.path.join(fdir, "gpu_b200.f90")).read().lower()))
    assert bound and bound <= stub, sorted(bound - stub)
    shim = set(re.findall(r"^(?:void|int) (\w+)\(", open(os.path.join(fdir, "gpu_b200_ansi.c")).read(), re.M))
    L = gpu.load_library()
    for name in bound:
        assert name in shim or hasattr(L, name), name
    patch = open(os.path.join(fdir, "Makefile.src.patch")).read()
    assert "gpu_b200_ansi.o" in patch and "nogpu_b200_ansi.o" in patch and "-lpcb200" in patch
    # the diagnostics of a GPU run are computed by the OpenMP helper thread (Makefile.src:509-515 for gpu_astaroth)
    assert "override MULTITHREADING=openmp" in patch
