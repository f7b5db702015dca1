"""GPU test: BASELINE configs[1] at its own size against reference output -- the CUDA path reproduces the time series the
reference holds for samples/turbulence/helical-MHDturb-nopart and helical-MHDturb32-4procs (128^3 forced isothermal MHD, Courant
step from the device reduction, forcing coefficients from the library's pcb200_forcing_coefs_hel with the dt of the step)."""
import ctypes as C

import pytest

import helical128 as h
from test_oracle_helical128 import kdat

pytestmark = pytest.mark.gpu


class _Coefs:
    pass


def _coefs(L, fs, dt):
    phase, kvec, phi = fs.draw()
    a3 = C.c_double * 3
    dp = C.POINTER(C.c_double)
    L.pcb200_forcing_coefs_hel.argtypes = [C.c_double] * 5 + [dp, C.c_double, C.c_double, dp, dp, dp, dp, dp]
    c1, c2, fda, kk, fact = a3(), a3(), a3(), a3(), C.c_double()
    L.pcb200_forcing_coefs_hel(h.FORCE, h.RELHEL, fs.kav, 1.0, dt, a3(*kvec), phase, phi, c1, c2, fda, kk, C.byref(fact))
    o = _Coefs()
    o.coef1, o.coef2, o.fda, o.kk = list(c1), list(c2), list(fda), list(kk)
    o.phase, o.fact, o.k1_ff = phase, fact.value, 1.0
    return o


@pytest.mark.parametrize("name", sorted(h.LAYOUT))
def test_helical128_reference_series_on_the_gpu(name):
    from pencil_code_b200.gpu import GPU
    rows = h.golden()[name]
    cfg, f = h.setup(name)
    fs = h.ForcingStream(kdat())
    g = GPU(cfg)
    g.initialize_GPU(f)
    alpha = (0.0, -5.0 / 9.0, -153.0 / 128.0)
    beta = (1.0 / 3.0, 15.0 / 16.0, 8.0 / 15.0)
    t, dt = 0.0, 0.0
    try:
        for it in range(6):
            t0 = t
            if it in rows:
                g.request_diagnostics()
            # gpu.py:time_step with the forcing coefficients formed once the Courant dt of this step is known (after substep 1)
            g.set_dt(dt, True)
            ds = 0.0
            for itsub in (1, 2, 3):
                ds = 0.0 if itsub == 1 else alpha[itsub - 1] * ds
                if itsub == 3:
                    g.set_forcing_hel(_coefs(g.L, fs, dt))
                g.before_boundary_gpu(f, False, itsub, t)
                g.rhs_GPU(f, itsub, t)
                if itsub == 1:
                    dt = g.get_dt()
                ds += 1.0
                t += ds * (dt * beta[itsub - 1])
                g.update_after_substep_gpu()
            if it in rows:
                got = h.format_reduced(g.get_gpu_reduced_vars(), cfg["eta"], rows[it])
                assert got == {k: rows[it][k] for k in got}, (it, got, rows[it])
                assert "%.3f" % t0 == rows[it]["t"] and "%.2E" % dt == rows[it]["dt"]
    finally:
        g.finalize_GPU()
