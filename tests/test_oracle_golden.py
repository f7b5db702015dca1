"""CPU tests pinning the ORACLE (oracle/pc_oracle.cpp) against the reference's own artefacts:

* tests/golden/deriv_golden.npz -- outputs of the reference's numpy stencils
  (python/pencil/math/derivatives/der_6th_order_w_ghosts.py), made by make_deriv_golden.py;
* analytic derivatives of trigonometric fields (mirrors python/tests/derivatives.py:17-50);
* the closed-form it=0 row of samples/gputest/cpu-test.out:273 (urms, umax, epsK, rhom);
* internal consistency: Courant dt formula, RK3 coefficients, periodic ghosts.
"""
import math
import os

import numpy as np
import pytest

from pencil_code_b200 import config as pcfg
from pencil_code_b200 import synthetic

HERE = os.path.dirname(os.path.abspath(__file__))
NG = 3


def _periodic(f):
    nz, ny, nx = f.shape[-3] - 6, f.shape[-2] - 6, f.shape[-1] - 6
    f[..., :3] = f[..., nx:nx + 3]; f[..., nx + 3:] = f[..., 3:6]
    f[..., :3, :] = f[..., ny:ny + 3, :]; f[..., ny + 3:, :] = f[..., 3:6, :]
    f[..., :3, :, :] = f[..., nz:nz + 3, :, :]; f[..., nz + 3:, :, :] = f[..., 3:6, :, :]
    return f


def test_deriv_against_reference_python_golden(oracle_mod):
    g = np.load(os.path.join(HERE, "golden", "deriv_golden.npz"))
    f = g["f"]
    dx, dy, dz = g["dxyz"]
    nz, ny, nx = (s - 6 for s in f.shape)
    cfg = pcfg.make_config(nx, False, False, ny=ny, nz=nz)
    cfg.update(dx=float(dx), dy=float(dy), dz=float(dz), nvar=1)
    o = oracle_mod.Oracle(cfg)
    f4 = np.ascontiguousarray(f[None])
    # der5 is excluded: the python post-processing xder5_6th sums f(+k)+f(-k) (symmetric, cannot be a
    # 5th derivative) whereas the Fortran hot path (deriv.f90:855-935) uses differences; the Fortran
    # form is authoritative and is checked analytically in test_der5_der6_analytic below.
    for kind, name in ((1, "der"), (2, "der2"), (6, "der6")):
        for j, ax in ((1, "x"), (2, "y"), (3, "z")):
            got = o.deriv(f4, 0, j, kind)
            ref = g["%s_%s" % (name, ax)]
            scale = np.abs(ref).max()
            # the python reference uses 1/(60 dx) where Fortran/oracle use (1/60)*dx_1: last-ulp differences
            assert np.abs(got - ref).max() <= 4e-15 * scale, (name, ax)


@pytest.mark.parametrize("n", [16, 32])
def test_deriv_analytic_convergence(oracle_mod, n):
    cfg = pcfg.make_config(n, False, False)
    cfg["nvar"] = 1
    o = oracle_mod.Oracle(cfg)
    i = np.arange(n)
    x = (cfg["x0"] + i * cfg["dx"])[None, None, :]
    y = (cfg["y0"] + i * cfg["dy"])[None, :, None]
    z = (cfg["z0"] + i * cfg["dz"])[:, None, None]
    f = np.zeros((1, n + 6, n + 6, n + 6))
    f[0, 3:-3, 3:-3, 3:-3] = np.sin(x) * np.cos(2 * y) * np.sin(z + 0.3)
    _periodic(f)
    ex = {
        (1, 1): np.cos(x) * np.cos(2 * y) * np.sin(z + 0.3),
        (1, 2): -2 * np.sin(x) * np.sin(2 * y) * np.sin(z + 0.3),
        (1, 3): np.sin(x) * np.cos(2 * y) * np.cos(z + 0.3),
        (2, 1): -np.sin(x) * np.cos(2 * y) * np.sin(z + 0.3),
        (2, 2): -4 * np.sin(x) * np.cos(2 * y) * np.sin(z + 0.3),
        (12, 0): -2 * np.cos(x) * np.sin(2 * y) * np.sin(z + 0.3),
        (23, 0): -2 * np.sin(x) * np.sin(2 * y) * np.cos(z + 0.3),
        (31, 0): np.cos(x) * np.cos(2 * y) * np.cos(z + 0.3),
    }
    tol = 40.0 * (2 * cfg["dx"]) ** 6 / 140.0  # 6th order: error ~ k^7 h^6/140
    for (kind, j), e in ex.items():
        got = o.deriv(f, 0, max(j, 1), kind)
        assert np.abs(got - e).max() < tol, (kind, j, np.abs(got - e).max(), tol)


def test_der5_der6_analytic(oracle_mod):
    """deriv.f90:855-935 / :936-1072 on sin(x): d5 = cos x, d6 = -sin x (2nd-order accurate stencils)."""
    n = 64
    cfg = pcfg.make_config(n, False, False, ny=4, nz=4)
    cfg["nvar"] = 1
    o = oracle_mod.Oracle(cfg)
    x = (cfg["x0"] + np.arange(n) * cfg["dx"])[None, None, :]
    f = np.zeros((1, 10, 10, n + 6))
    f[0, 3:-3, 3:-3, 3:-3] = np.sin(x)
    _periodic(f)
    h = cfg["dx"]
    assert np.abs(o.deriv(f, 0, 1, 5) - np.cos(x)).max() < 0.5 * h * h
    assert np.abs(o.deriv(f, 0, 1, 6) + np.sin(x)).max() < 0.5 * h * h
    # upwind variant: D6*dx^5/60 (deriv.f90:978-981)
    assert np.allclose(o.deriv(f, 0, 1, 7), o.deriv(f, 0, 1, 6) * h**5 / 60.0, rtol=1e-12, atol=1e-18)


def test_gputest_known_answer_row0(oracle_mod):
    """samples/gputest/cpu-test.out:273: it=0  urms=1.73205 umax=2.43390 epsK=4.105E-02 rhom=3.4212
    (32^3, inituu='ABC' ampluu=1 k=2, lnrho=1.23, nu=1e-3; hydro.f90:2214-2221)."""
    n = 32
    cfg = pcfg.make_config(n, entropy=True, magnetic=True, nu=1e-3, eta=1e-3, chi=1e-4, cdt=0.4, cdtv=0.3)
    o = oracle_mod.Oracle(cfg)
    f = np.zeros(synthetic.farray_shape(cfg))
    i = np.arange(n)
    x = (cfg["x0"] + i * cfg["dx"])[None, None, :]
    y = (cfg["y0"] + i * cfg["dy"])[None, :, None]
    z = (cfg["z0"] + i * cfg["dz"])[:, None, None]
    I = (slice(3, -3),) * 3
    f[0][I] = np.sin(2 * z) + np.cos(2 * y) + 0 * x
    f[1][I] = np.sin(2 * x) + np.cos(2 * z) + 0 * y
    f[2][I] = np.sin(2 * y) + np.cos(2 * x) + 0 * z
    f[3][I] = 1.23
    o.update_ghosts(f)
    df = np.zeros_like(f)
    _, d = o.rhs(f, df, lcourant=True, diag=True)
    assert "%.5f" % d["urms"] == "1.73205"
    assert "%.5f" % d["umax"] == "2.43390"
    assert "%.3E" % d["epsK"] == "4.105E-02"
    assert "%.4f" % d["rhom"] == "3.4212"


def _gputest_setup():
    """samples/gputest as its CPU run was configured (cpu-test.out:98-272): 32^3, [-pi,pi]^3, ABC flow k=2 amplitude 1,
    lnrho=1.23, ss=0, A=0, nu=1e-3, eta=1e-3, heat conduction 'nothing', viscous + Ohmic heating on, dt=1e-3 fixed."""
    n = 32
    cfg = pcfg.make_config(n, entropy=True, magnetic=True, nu=1e-3, eta=1e-3, chi=0.0, cdt=0.4, cdtv=0.3)
    f = np.zeros(synthetic.farray_shape(cfg))
    i = np.arange(n)
    x = (cfg["x0"] + i * cfg["dx"])[None, None, :]
    y = (cfg["y0"] + i * cfg["dy"])[None, :, None]
    z = (cfg["z0"] + i * cfg["dz"])[:, None, None]
    I = (slice(3, -3),) * 3
    f[0][I] = np.sin(2 * z) + np.cos(2 * y) + 0 * x  # hydro.f90:2214-2221 (inituu='ABC')
    f[1][I] = np.sin(2 * x) + np.cos(2 * z) + 0 * y
    f[2][I] = np.sin(2 * y) + np.cos(2 * x) + 0 * z
    f[3][I] = 1.23
    return cfg, f


def gputest_rows():
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gputest_cpu_rows.txt")
    rows = [l.split() for l in open(path) if not l.startswith("#")]
    assert len(rows) == 20
    return rows


def format_gputest_row(d):
    """print.in formats: urms(f10.5) umax(f10.5) epsK(e10.3) rhom(f9.4); Fortran e10.3 prints 0.xxxE+yy only without
    the 1p scale factor -- legend.f90 adds 1p, so the mantissa is d.ddd as in C's %.3E."""
    return ["%.5f" % d["urms"], "%.5f" % d["umax"], "%.3E" % d["epsK"], "%.4f" % d["rhom"]]


def test_gputest_20_rows_of_the_reference_cpu_binary(oracle_mod):
    """All 20 time-series rows the reference's own CPU build printed for samples/gputest (cpu-test.out:275-294,
    committed as tests/golden/gputest_cpu_rows.txt): 20 RK3 steps of hydro + continuity + entropy (viscous heating)
    + induction, every printed digit of urms, umax, epsK, rhom."""
    cfg, f = _gputest_setup()
    o = oracle_mod.Oracle(cfg)
    o.update_ghosts(f)
    df = np.zeros_like(f)
    t, dt = 0.0, 1e-3
    for row in gputest_rows():
        dt, t, d = o.time_step(f, df, False, dt, t, diag=True)
        assert format_gputest_row(d) == row[2:6], (row, d)
        assert "%.2E" % dt == row[1]


def test_periodic_ghosts(oracle_mod):
    cfg = pcfg.make_config(8, True, True, ny=6, nz=5)
    o = oracle_mod.Oracle(cfg)
    rng = np.random.default_rng(1)
    f = rng.standard_normal(synthetic.farray_shape(cfg))
    o.update_ghosts(f)
    g = f.copy()
    _periodic(g)
    assert np.array_equal(f, g)
    # corners must equal the diagonally opposite interior cell
    assert f[2, 0, 0, 0] == f[2, 5, 6, 8]


def test_courant_dt_formula(oracle_mod):
    """uniform state: dt1 = sqrt((|u|.dline_1 + sqrt(cs2 dxyz_2))^2/cdt^2 + (nu dxyz_2/cdtv)^2) (equ.f90:1872-1972)."""
    n = 8
    cfg = pcfg.make_config(n, entropy=False, magnetic=False, nu=0.01, cdt=0.4, cdtv=0.3)
    o = oracle_mod.Oracle(cfg)
    f = np.zeros(synthetic.farray_shape(cfg))
    f[0] = 0.3; f[1] = -0.2; f[2] = 0.1
    df = np.zeros_like(f)
    dt1, _ = o.rhs(f, df, lcourant=True)
    d1 = 1.0 / cfg["dx"]
    adv = (0.3 + 0.2 + 0.1) * d1 + math.sqrt(1.0 * 3 * d1 * d1)
    ref = math.sqrt((adv / 0.4) ** 2 + (0.01 * 3 * d1 * d1 / 0.3) ** 2)
    assert abs(dt1 - ref) <= 1e-14 * ref
    # uniform state -> RHS vanishes
    assert np.abs(df[:, 3:-3, 3:-3, 3:-3]).max() < 1e-13


def test_rk3_linear_advection_order(oracle_mod):
    """2N-RK3 (timestep.f90:39-44) on the pure-diffusion limit: A_t = eta del2 A is integrated with
    3rd-order accuracy; halving dt reduces the error ~8x."""
    n = 16
    cfg = pcfg.make_config(n, entropy=False, magnetic=True, nu=0.0, eta=0.05)
    cfg.update(llorentzforce=0)
    o = oracle_mod.Oracle(cfg)
    i = np.arange(n)
    x = (cfg["x0"] + i * cfg["dx"])[None, None, :]

    def run(nsteps, T=0.4):
        f = np.zeros(synthetic.farray_shape(cfg))
        f[5, 3:-3, 3:-3, 3:-3] = np.sin(x)  # A_y(x): uu=0 -> dA/dt = eta del2 A
        df = np.zeros_like(f)
        dt, t = T / nsteps, 0.0
        for _ in range(nsteps):
            dt, t, _ = o.time_step(f, df, False, dt, t)
        assert abs(t - T) < 1e-12
        return f[5, 3:-3, 3:-3, 3:-3]

    # semi-discrete exact: der2 eigenvalue for k=1
    h = cfg["dx"]
    lam = (-490 + 2 * (270 * math.cos(h) - 27 * math.cos(2 * h) + 2 * math.cos(3 * h))) / 180.0 / h**2
    exact = np.sin(x) * math.exp(0.05 * lam * 0.4)
    e1 = np.abs(run(4) - exact).max()
    e2 = np.abs(run(8) - exact).max()
    assert e2 < e1 / 6.0, (e1, e2)


def test_forcing_is_solenoidal_and_normalised(oracle_mod):
    """forcing.f90:1903-1964: coef1 = k (k x e), coef2 = relhel k x (k x e) are both perpendicular to k."""
    cfg = pcfg.make_config(8, False, False)
    o = oracle_mod.Oracle(cfg)
    kv = (3.0, -2.0, 4.0)
    fr = o.forcing_coefs(0.07, 1.0, 5.0, 1.0, 0.01, kv, 0.3, 1.1)
    k = np.array(kv)
    assert abs(np.dot(k, np.array(fr.coef1[:]))) < 1e-12
    assert abs(np.dot(k, np.array(fr.coef2[:]))) < 1e-12
    # |coef1| = k*|k x e| , |coef2| = k*|k x e| for relhel=1 -> equal magnitudes (maximally helical)
    assert abs(np.linalg.norm(fr.coef1[:]) - np.linalg.norm(fr.coef2[:])) < 1e-12
