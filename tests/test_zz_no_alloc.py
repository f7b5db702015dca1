"""GPU test of pcb200_dt1_max_noalloc (include/pcb200.h): the Courant reduction whose scratch is a file-scope static
__device__ array instead of cudaMalloc'ed memory (no counterpart in the reference).
Collected last on purpose (file name), so that a failure here cannot hide other tests under `pytest -x`.

The new entry point is called through the C ABI on the device f-array and compared with the existing path
(Courant substep 1 through rhs_gpu_c, then pcb200_get_dt1_max) on the same state, twice in a row, the second time with a
smaller reduction grid so that the scratch array still holds stale values of the first call."""
import ctypes as C

import numpy as np
import pytest

from pencil_code_b200 import config as pcfg
from pencil_code_b200 import synthetic
from pencil_code_b200.gpu import GPU

pytestmark = pytest.mark.gpu


def test_dt1_max_without_allocation_matches_the_existing_entry_point_twice_in_a_row():
    cfg = pcfg.make_config(32, entropy=True, magnetic=True, nu=2e-3, eta=2e-3, chi=1e-3, cdt=0.4, cdtv=0.3)
    f = synthetic.initial_condition(cfg, seed=7)
    rt = C.CDLL("libcudart.so.12")
    rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
    rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
    rt.cudaFree.argtypes = [C.c_void_p]
    g = GPU(cfg)
    g.register_GPU()
    g.initialize_GPU(f)
    out_dev = C.c_void_p()
    assert rt.cudaMalloc(C.byref(out_dev), 8) == 0  # the CALLER owns the output; the library allocates nothing
    try:
        max_blocks = 4096  # PCB200_NOALLOC_MAX_BLOCKS
        for nblocks in (1184, 37):  # second call: smaller grid, scratch slots 37.. keep the first call's values
            p_in, _ = g.get_farray_ptr_gpu()
            rc = g.L.pcb200_dt1_max_noalloc(C.c_void_p(p_in), out_dev, C.byref(C.c_int(nblocks)))
            assert rc == 0
            g.synchronize()
            got = np.zeros(1)
            assert rt.cudaMemcpy(got.ctypes.data, out_dev, 8, 2) == 0
            # existing entry point on the same state: Courant substep 1, then the reduced dt1
            g.set_dt(0.0, True)
            g.rhs_GPU(f, 1)
            want = g.get_dt1_max()
            assert want > 0 and np.isfinite(got[0])
            assert abs(got[0] - want) <= 1e-12 * want, (nblocks, got[0], want)
            g.rhs_GPU(f, 2)
            g.rhs_GPU(f, 3)  # a different state for the second round
        # sizes beyond the static array are rejected with an error code (and a message on stderr), nothing is launched
        p_in, _ = g.get_farray_ptr_gpu()
        assert g.L.pcb200_dt1_max_noalloc(C.c_void_p(p_in), out_dev, C.byref(C.c_int(max_blocks + 1))) == 1
        assert g.L.pcb200_dt1_max_noalloc(C.c_void_p(p_in), out_dev, C.byref(C.c_int(0))) == 1
        assert g.L.pcb200_dt1_max_noalloc(C.c_void_p(p_in), out_dev, C.byref(C.c_int(max_blocks))) == 0
        g.synchronize()
    finally:
        rt.cudaFree(out_dev)
        g.finalize_GPU()
