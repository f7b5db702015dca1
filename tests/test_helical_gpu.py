"""GPU test: the CUDA path reproduces the reference's own CPU run of samples/helical-MHDturb (config[1]'s physics at
32^3) through the C ABI -- every digit reference.out prints of umax, urms, brms, jrms, jmax, abm, jbm, oum at
it = 0, 2, 4, 6, 8 via get_gpu_reduced_vars, and aa(5,5,0:4,1) of the final snapshot (tests/read_data.ref) via
copy_farray_from_GPU.  Forcing coefficients come from the library's own host routine pcb200_forcing_coefs_hel, fed with
the run's nr_f90 stream (tests/helical_mhdturb.py)."""
import ctypes as C

import numpy as np
import pytest

import helical_mhdturb as hm

pytestmark = pytest.mark.gpu


class _Coefs:
    pass


def _coefs(L, fs, dt):
    phase, kvec, phi = fs.draw()
    a3 = C.c_double * 3
    dp = C.POINTER(C.c_double)
    L.pcb200_forcing_coefs_hel.argtypes = [C.c_double] * 5 + [dp, C.c_double, C.c_double, dp, dp, dp, dp, dp]
    c1, c2, fda, kk, fact = a3(), a3(), a3(), a3(), C.c_double()
    L.pcb200_forcing_coefs_hel(hm.FORCE, hm.RELHEL, fs.kav, 1.0, dt, a3(*kvec), phase, phi, c1, c2, fda, kk, C.byref(fact))
    o = _Coefs()
    o.coef1, o.coef2, o.fda, o.kk = list(c1), list(c2), list(fda), list(kk)
    o.phase, o.fact, o.k1_ff = phase, fact.value, 1.0
    return o


@pytest.mark.parametrize("variant", [0, 1, 5])
def test_helical_mhdturb_reference_series_on_the_gpu(variant):
    from pencil_code_b200.gpu import GPU
    g = hm.golden()
    cfg, f = hm.setup()
    cfg["kernel_variant"] = variant
    rows = {int(r["it"]): r for r in g["rows"]}
    fs = hm.ForcingStream(g["kdat"])
    gpu = GPU(cfg)
    gpu.initialize_GPU(f)
    t, dt = 0.0, hm.DT
    try:
        for it in range(10):
            frc = _coefs(gpu.L, fs, dt)
            dt, t = gpu.time_step(f, dt, False, t, forcing=frc, diagnostics=it in rows)
            if it in rows:
                got = hm.format_reduced(gpu.get_gpu_reduced_vars())
                assert got == {k: rows[it][k] for k in got}, (it, got, rows[it])
        gpu.copy_farray_from_GPU(f)
    finally:
        gpu.finalize_GPU()
    assert ["%g" % v for v in f[cfg["iax"] + 1, 3:8, 3 + 5, 3 + 5]] == g["aa_5_5_0to4_y"]
