"""GPU parity of the vertically stratified path through the C ABI:
* samples/conv-slab for 41 Courant steps: the time series read back with get_gpu_reduced_vars must agree with the
  reference's own double-precision series (tests/golden/convslab_rows.txt) and the final f-array with the oracle;
* one substep and the ghost zones (z boundary conditions s / a / a2 / c1 / cT) against the oracle on a perturbed state."""
import numpy as np
import pytest

import convslab
from parity_util import GPU, rel_linf
from test_oracle_convslab import check_rows, run_series

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cp", [1.0, 2.0])
def test_convslab_41_steps_on_the_gpu_match_the_reference_series_and_the_oracle(oracle_mod, cp):
    c = convslab.config(cp=cp)  # cp = 2: samples/conv-slab_cp_2, same series with ssm/cp
    f0 = convslab.initial_condition(c)
    fg = f0.copy()
    g = GPU(c)
    g.register_GPU()
    g.initialize_GPU(fg)
    try:
        def step(dt, t, want):
            dt, t = g.time_step(fg, dt, True, t, diagnostics=want)
            return dt, t, (g.get_gpu_reduced_vars() if want else None)
        rows = run_series(step)
        g.copy_farray_from_GPU(fg)
    finally:
        g.finalize_GPU()
    # same absolute tolerance as for the oracle: one unit of the last-but-one digit the reference prints (measured on B200:
    # 0.05 units in every column, i.e. the CUDA path reproduces the reference's 13-14 digits)
    for r in rows:
        r["ssm"] /= cp
    worst = check_rows(rows)
    print("worst deviation in units of the last-but-one printed digit:", worst)
    o = oracle_mod.Oracle(c)
    fc, df = f0.copy(), np.zeros_like(f0)
    dt = t = 0.0
    for _ in range(41):
        dt, t, _ = o.time_step(fc, df, True, dt, t)
    o.update_ghosts(fc)
    err = rel_linf(fg, fc)
    assert err.max() <= 1e-9, err
    # ghost zones (boundary conditions) as well
    for k in range(5):
        den = np.abs(fc[k]).max()
        assert np.abs(fg[k] - fc[k]).max() <= 1e-9 * den, k


@pytest.mark.parametrize("n", [16, 24])
def test_one_substep_and_closed_boundaries_against_the_oracle(oracle_mod, n):
    c = convslab.config(n)
    f0 = convslab.initial_condition(c)
    rng = np.random.default_rng(11)
    f0[:, 3:-3, 3:-3, 3:-3] += 0.02 * rng.standard_normal((5, n, n, n))
    o = oracle_mod.Oracle(c)
    for ldt in (False, True):
        fg, fc, df = f0.copy(), f0.copy(), np.zeros_like(f0)
        g = GPU(c)
        g.register_GPU()
        g.initialize_GPU(fg)
        try:
            g.set_dt(1e-3, ldt)
            g.rhs_GPU(fg, 1)
            dtg = g.get_dt()
            g.copy_farray_from_GPU(fg)
        finally:
            g.finalize_GPU()
        dtc, _, _ = o.substep(fc, df, 1, ldt, 1e-3, 0.0)
        o.update_ghosts(fc)
        assert abs(dtg - dtc) <= 1e-13 * dtc
        assert rel_linf(fg, fc).max() <= 1e-12
        for k in range(5):
            assert np.abs(fg[k] - fc[k]).max() <= 1e-12 * np.abs(fc[k]).max(), ("ghost zones", k)


def test_rotating_periodic_box_with_gravity_uses_the_stratified_kernel(oracle_mod):
    """gravity + Coriolis force in a triply periodic box (no closed boundaries): kernel variant 1 with the STRAT terms"""
    from pencil_code_b200 import config as pcfg
    from parity_util import run_both
    c = pcfg.make_config(24, entropy=True, magnetic=True, nu=2e-3, eta=2e-3, chi=1e-3)
    c.update(gravz=-0.3, Omega=0.7)
    fg, fc, _, _, _ = run_both(c, oracle_mod, 1, dt=1e-3, substeps_only=3)
    assert rel_linf(fg, fc).max() <= 1e-12


def test_unsupported_closed_boundary_set_ups_are_refused():
    """the module aborts on what it does not implement (gpu_astaroth.f90:211-228 convention): checked in a subprocess"""
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, 'tests'); import convslab; from pencil_code_b200.gpu import GPU; import numpy as np;"
            "c = convslab.config(16); c['kernel_variant'] = 10; f = convslab.initial_condition(c); g = GPU(c); g.register_GPU(); g.initialize_GPU(f)")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode != 0 and "variant 1" in (r.stderr + r.stdout)
