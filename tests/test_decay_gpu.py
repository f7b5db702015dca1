"""GPU test: the CUDA path reproduces the reference's own CPU run of samples/decay-reconnection through the C ABI -- every
digit reference.out prints of t, dt (Courant step from the device reduction), urms, brms, jrms, epsK, epsM, oum, abm, jbm at
it = 0, 10, ..., 50 (tests/golden/decay_reconnection_rows.txt): 51 RK3 steps of isothermal MHD with 6th-order upwinding of u,
lnrho and A."""
import pytest

import decay_reconnection as dr

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("variant", [0, 1, 5])
def test_decay_reconnection_reference_series_on_the_gpu(variant):
    from pencil_code_b200.gpu import GPU
    rows = {int(r["it"]): r for r in dr.golden()}
    cfg, f = dr.setup()
    cfg["kernel_variant"] = variant
    gpu = GPU(cfg)
    gpu.initialize_GPU(f)
    t, dt = 0.0, 0.0
    try:
        for it in range(dr.NSTEPS):
            t0 = t
            dt, t = gpu.time_step(f, dt, True, t, diagnostics=it in rows)
            if it in rows:
                got = dr.format_reduced(gpu.get_gpu_reduced_vars(), cfg["eta"])
                assert got == {k: rows[it][k] for k in got}, (it, got, rows[it])
                assert "%.3E" % t0 == rows[it]["t"] and "%.3E" % dt == rows[it]["dt"], (it, t0, dt)
    finally:
        gpu.finalize_GPU()
