"""CPU tests pinning the ORACLE to the reference's 14-digit golden of a vertically stratified run:
samples/conv-slab/reference.out.double (tests/golden/convslab_rows.txt, made by tests/golden/make_convslab_golden.py).

The run exercises, over 41 RK3 steps with the Courant time step: advection, Coriolis force, nu-const viscosity and
viscous heating, continuity and entropy equations with 6th-order upwinding, the ideal-gas pressure gradient, K-profile heat
conduction, the cooling layer, constant gravity and the z boundary conditions s / a / a2 / c1 / cT.  Every printed column
must agree with the reference to the last digits it prints."""
import os

import numpy as np
import pytest

import convslab

HERE = os.path.dirname(os.path.abspath(__file__))
# absolute tolerance per column = one unit in the last-but-one printed digit (print.in.double formats);
# the oracle's measured deviations are 10-100 times smaller
TOL = dict(t=2e-13, dt=2e-15, urms=1e-13, umax=1e-13, rhom=1e-12, ssm=1e-13, dtc=1e-12, dtu=1e-12, dtnu=1e-12, dtchi=1e-12)


def run_series(step_fn, nsteps=41, it1=10):
    rows = []
    dt, t = 0.0, 0.0
    for it in range(nsteps):
        want = it % it1 == 0
        t0 = t
        dt, t, d = step_fn(dt, t, want)
        if want:
            rows.append(convslab.series_row(it, t0, dt, d))
    return rows


def check_rows(rows, tol_scale=1.0):
    gold = convslab.golden_rows(os.path.join(HERE, "golden", "convslab_rows.txt"))
    assert [r["it"] for r in rows] == [g["it"] for g in gold]
    worst = {}
    for r, g in zip(rows, gold):
        for k, tol in TOL.items():
            err = abs(r[k] - g[k])
            worst[k] = max(worst.get(k, 0.0), err / tol)
            assert err <= tol * tol_scale, (r["it"], k, r[k], g[k])
    return worst


@pytest.mark.parametrize("cp", [1.0, 2.0])
def test_convslab_41_steps_match_the_reference_double_precision_series(oracle_mod, cp):
    """cp = 2: samples/conv-slab_cp_2/reference.out.double -- the same 41 rows digit for digit (checked by
    tests/golden/make_convslab_golden.py), its entropy column being ssbycpm = ssm/cp."""
    c = convslab.config(cp=cp)
    f = convslab.initial_condition(c)
    o = oracle_mod.Oracle(c)
    df = np.zeros_like(f)
    rows = run_series(lambda dt, t, want: o.time_step(f, df, True, dt, t, diag=want))
    for r in rows:
        r["ssm"] /= cp
    check_rows(rows)


def test_initial_stratification_is_hydrostatic_and_in_thermal_equilibrium(oracle_mod):
    """piecewise polytropic start (density.f90:2111-2177, entropy.f90:2186-2260): away from the blended interfaces the
    pressure gradient balances gravity and the conductive flux is constant, so du_z/dt and ds/dt are small there."""
    c = convslab.config()
    f = convslab.initial_condition(c)
    f[c["iux"]:c["iux"] + 3] = 0.0
    o = oracle_mod.Oracle(dict(c, cool=0.0))
    o.update_ghosts(f)
    df = np.zeros_like(f)
    o.rhs(f, df)
    z = c["z0"] + c["dz"] * np.arange(c["nz"])
    inner = (np.abs(z - c["z1"]) > 0.3) & (np.abs(z - c["z2"]) > 0.3) & (z > c["z0"] + 0.2) & (z < c["ztop"] - 0.1)
    assert inner.sum() >= 6
    duz = df[c["iux"] + 2, 3:-3, 3:-3, 3:-3][inner]
    dss = df[c["iss"], 3:-3, 3:-3, 3:-3][inner]
    assert np.abs(duz).max() < 2e-3 * abs(c["gravz"])
    assert np.abs(dss).max() < 2e-3


@pytest.mark.parametrize("bc", ["s", "a", "a2"])
def test_symmetry_conditions_in_z(oracle_mod, bc):
    """bc_sym_z boundcond.f90:3842-3900: ghost planes as (anti)symmetric images; 'a' zeroes the boundary plane"""
    c = convslab.config(16)
    c["bcz"] = [bc] * 4 + ["s"]
    c["lheatc_Kprof"] = 0
    rng = np.random.default_rng(5)
    f = rng.standard_normal((5, 22, 22, 22))
    g = f.copy()
    oracle_mod.Oracle(c).update_ghosts(g)
    n1, n2 = 3, 18
    for i in (1, 2, 3):
        if bc == "s":
            assert np.array_equal(g[0, n1 - i], g[0, n1 + i]) and np.array_equal(g[0, n2 + i], g[0, n2 - i])
        elif bc == "a":
            assert np.array_equal(g[0, n1 - i], -g[0, n1 + i]) and np.array_equal(g[0, n2 + i], -g[0, n2 - i])
        else:
            assert np.allclose(g[0, n1 - i] + g[0, n1 + i], 2 * g[0, n1], rtol=0, atol=1e-14)
            assert np.allclose(g[0, n2 + i] + g[0, n2 - i], 2 * g[0, n2], rtol=0, atol=1e-14)
    if bc == "a":
        assert not g[0, n1].any() and not g[0, n2].any()
    # x and y ghosts stay periodic images of the (possibly modified) interior
    assert np.array_equal(g[0, 3:-3, 3:-3, :3], g[0, 3:-3, 3:-3, 16:19])
    assert np.array_equal(g[0, 3:-3, :3, 3:-3], g[0, 3:-3, 16:19, 3:-3])


def test_flux_and_temperature_conditions_for_entropy(oracle_mod):
    """c1 (entropy_bcs.f90:93-235): the ghost values make the centred gradient of ln T equal -Fbot/(K T) at the bottom;
    cT (eos_idealgas.f90:3112-3260): cs2 on the top plane equals cs2top and ln cs2 is antisymmetric about it."""
    c = convslab.config(16)
    f = convslab.initial_condition(c)
    rng = np.random.default_rng(6)
    f[3:5, 3:-3, 3:-3, 3:-3] += 0.01 * rng.standard_normal((2, 16, 16, 16))
    o = oracle_mod.Oracle(c)
    o.update_ghosts(f)
    g1, cv1, cp = c["gamma"] - 1.0, c["gamma"] / c["cp"], c["cp"]
    lncs2 = np.log(c["cs20"]) + cv1 * f[4] + g1 * (f[3] - c["lnrho0"])
    n1, n2 = 3, 18
    cs2b = np.exp(lncs2[n1])
    for i in (1, 2, 3):
        # d lnTT/dz by the two-point difference the condition enforces: -(gamma-1)... = -FbotKbot*cp*(gamma-1)/cs2 * ... in lncs2 units
        lhs = (lncs2[n1 + i] - lncs2[n1 - i]) / (2 * i * c["dz"])
        rhs = -cv1 * (cp - cp / c["gamma"]) * cp * c["FbotKbot"] / cs2b
        assert np.allclose(lhs, rhs, rtol=1e-11, atol=0)
        assert np.allclose(lncs2[n2 + i] + lncs2[n2 - i], 2 * np.log(c["cs2top"]), rtol=0, atol=1e-13)
    assert np.allclose(np.exp(lncs2[n2]), c["cs2top"], rtol=1e-13)
