"""CPU test pinning the ORACLE's upwinded MHD path and its Courant time step to reference output: all six rows of the time
series the reference holds for its own CPU run of samples/decay-reconnection (tests/golden/decay_reconnection_rows.txt; set-up,
random field and generator in tests/decay_reconnection.py, tests/nr_f90.py).  51 RK3 steps with dt from the advective
(|u| + sqrt(cs2 + vA2)), viscous and resistive limits; 6th-order upwinding of u, lnrho and A (the upwinded induction equation,
magnetic.f90:6257-6313); Lorentz force; nu-const viscosity; eta-const resistivity."""
import math

import numpy as np

import decay_reconnection as dr


def _curl(o, f, k):
    d = lambda v, j: o.deriv(f, k + v, j, 1)  # noqa: E731
    return np.array([d(2, 2) - d(1, 3), d(0, 3) - d(2, 1), d(1, 1) - d(0, 2)])


def test_decay_reconnection_series(oracle_mod):
    rows = {int(r["it"]): r for r in dr.golden()}
    cfg, f = dr.setup()
    o = oracle_mod.Oracle(cfg)
    o.update_ghosts(f)
    df = np.zeros_like(f)
    t, dt = 0.0, 0.0
    iux, iax = cfg["iux"], cfg["iax"]
    I = (slice(3, -3),) * 3
    for it in range(dr.NSTEPS):
        if it in rows:  # columns the reduced slots do not carry, from the oracle's own stencils
            row = rows[it]
            om, b = _curl(o, f, iux), _curl(o, f, iax)
            u = f[iux:iux + 3][(slice(None),) + I]
            divu = sum(o.deriv(f, iux + j, j + 1, 1) for j in range(3))
            assert "%.3E" % math.sqrt((om ** 2).sum(0).mean()) == row["orms"]
            assert "%.3E" % (divu ** 2).mean() == row["divu2m"]
            assert "%.3E" % (u * b).sum(0).mean() == row["ubm"]
        t0 = t
        dt, t, d = o.time_step(f, df, True, dt, t, diag=it in rows)
        o.update_ghosts(f)
        if it in rows:
            got = dr.format_reduced(d, cfg["eta"])
            assert got == {k: rows[it][k] for k in got}, (it, got, rows[it])
            assert "%.3E" % t0 == rows[it]["t"] and "%.3E" % dt == rows[it]["dt"], (it, t0, dt)


def test_ran0_properties():
    """general.f90:867-889 'minimal standard' generator: numbers in (0,1), reproducible, negative seeds accepted."""
    import nr_f90
    a = nr_f90.Ran0(-10)
    u = np.array([a.next() for _ in range(5000)])
    assert 0.0 < u.min() and u.max() < 1.0 and abs(u.mean() - 0.5) < 0.02
    b = nr_f90.Ran0(-10)
    assert [b.next() for _ in range(50)] == list(u[:50])
    # Park-Miller known answer (Numerical Recipes ran0): the state after one call from seed 1 (without the mask) is 16807
    c = nr_f90.Ran0(1 ^ nr_f90.MASK)
    assert abs(c.next() - 16807.0 / nr_f90.IM) < 1e-18
