"""CPU test pinning the ORACLE's magnetic + forcing path to reference output: the time series, the snapshot values and
the xy-averages the reference holds for its own CPU run of samples/helical-MHDturb (tests/golden/helical_mhdturb.json,
made by tests/golden/make_helical_golden.py; set-up and random streams in tests/helical_mhdturb.py, tests/nr_f90.py).
This is config[1]'s physics (isothermal MHD, nu-const, eta-const, helical forcing) over 10 RK3 steps: induction, Lorentz
force, resistive term through j = (grad div A - del2 A)/mu0, the forcing coefficients, their application after the
third substep and the generator behind both the initial condition and the forcing."""
import math

import numpy as np

import helical_mhdturb as hm


def _curl(o, f, k):
    d = lambda v, j: o.deriv(f, k + v, j, 1)  # noqa: E731
    return np.array([d(2, 2) - d(1, 3), d(0, 3) - d(2, 1), d(1, 1) - d(0, 2)])


def test_helical_mhdturb_series_snapshot_and_averages(oracle_mod):
    g = hm.golden()
    cfg, f = hm.setup()
    o = oracle_mod.Oracle(cfg)
    o.update_ghosts(f)
    fs = hm.ForcingStream(g["kdat"])
    assert fs.nk == 98 and abs(fs.kav - 3.13416) < 1e-12
    df = np.zeros_like(f)
    t, dt = 0.0, hm.DT
    rows = {int(r["it"]): r for r in g["rows"]}
    iux, ilnrho, iax = cfg["iux"], cfg["ilnrho"], cfg["iax"]
    I = (slice(3, -3),) * 3
    for it in range(10):
        if it in rows:
            # columns that get_gpu_reduced_vars does not carry, formed here from the oracle's own stencils
            row = rows[it]
            b, om = _curl(o, f, iax), _curl(o, f, iux)
            rho = np.exp(f[ilnrho][I])
            assert "%.3E" % math.sqrt((om ** 2).sum(0).mean()) == row["orms"]
            assert "%.3E" % np.sqrt((om ** 2).sum(0)).max() == row["omax"]
            assert "%.3E" % np.sqrt((b ** 2).sum(0) / rho).max() == row["vAmax"]   # mu0 = 1
            assert "%.4f" % (rho.sum() * cfg["dx"] * cfg["dy"] * cfg["dz"]) == row["mass"]
            for name, axes in (("bmx", (1, 2)), ("bmy", (1, 3)), ("bmz", (2, 3))):     # b is (3, z, y, x)
                assert "%.3E" % math.sqrt((b.mean(axis=axes) ** 2).sum(0).mean()) == row[name], name
            if it == 6:  # 4th record of xyaverages.dat, t = 0.24
                assert ["%g" % v for v in b[1].mean(axis=(1, 2))[:4]] == g["bymz_rec3_z0to3"]
        if it in (0, 3):  # records 0 and 1 of yaverages.dat / zaverages.dat (d2davg = 0.1: t = 0 and t = 0.12)
            bymxy = (_curl(o, f, iax)[1]).mean(axis=0)           # [y][x]
            assert ["%g" % v for v in bymxy[4:9, 10]] == g["bymxy_rec%d_x10_y4to8" % (it // 3)]
            if it == 3:
                bz = _curl(o, f, iax)[2]
                assert ["%g" % v for v in bz[0, 4:9, 13]] == g["bb3_xy_rec1_y4to8_x13"]  # slice_bb3.xy (z = z(n1))
                ux2mxz = (f[iux][I] ** 2).mean(axis=1)           # [z][x]
                assert ["%g" % v for v in ux2mxz[4:9, 10]] == g["ux2mxz_rec1_x10_z4to8"]
        phase, kvec, phi = fs.draw()
        frc = o.forcing_coefs(hm.FORCE, hm.RELHEL, fs.kav, 1.0, dt, kvec, phase, phi)
        dt, t, d = o.time_step(f, df, False, dt, t, forcing=frc, diag=it in rows)
        o.update_ghosts(f)
        if it in rows:
            got = hm.format_reduced(d)
            assert got == {k: rows[it][k] for k in got}, (it, got, rows[it])
            assert "%.3f" % (t - dt) == rows[it]["t"] and "%.2E" % dt == rows[it]["dt"]
    # var.dat after nt = 10 steps: aa(5,5,0:4,1) of tests/read_data.ref
    assert ["%g" % v for v in f[iax + 1, 3:8, 3 + 5, 3 + 5]] == g["aa_5_5_0to4_y"]


def test_mars_ran_properties():
    """general.f90:891-942: numbers lie in (0,1), re-seeding reproduces the stream, PUT of a saved state resumes it."""
    import nr_f90
    a = nr_f90.MarsRan()
    a.seed_put(-10, 0)
    u = a.vector(5000)
    assert 0.0 < u.min() and u.max() < 1.0 and abs(u.mean() - 0.5) < 0.02
    b = nr_f90.MarsRan()
    b.seed_put(-10, 0)
    assert list(b.vector(100)) == list(u[:100])
    c = nr_f90.MarsRan()
    c.seed_put(b.s1, b.s2)
    assert list(c.vector(50)) == list(u[100:150])
    assert -2**31 <= b.s1 < 2**31 and 0 < b.s2 < nr_f90.IM
