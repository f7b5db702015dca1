"""CPU test: the oracle reproduces the two time series the reference holds for BASELINE configs[1] at its own size (128^3
forced isothermal MHD with the Courant step; tests/golden/helical128_rows.txt, set-up in tests/helical128.py)."""
import json
import os

import numpy as np
import pytest

import helical128 as h


def kdat():
    return json.load(open(os.path.join(h.HERE, "golden", "helical_mhdturb.json")))["kdat"]


@pytest.mark.parametrize("name", sorted(h.LAYOUT))
def test_helical128_series(oracle_mod, name):
    rows = h.golden()[name]
    cfg, f = h.setup(name)
    o = oracle_mod.Oracle(cfg)
    o.update_ghosts(f)
    fs = h.ForcingStream(kdat())
    df = np.zeros_like(f)
    t, dt = 0.0, 0.0
    for it in range(6):
        t0 = t
        dt, t, d = o.time_step(f, df, True, dt, t, diag=it in rows)
        phase, kvec, phi = fs.draw()
        o.forcing_hel(f, o.forcing_coefs(h.FORCE, h.RELHEL, fs.kav, 1.0, dt, kvec, phase, phi))  # addforce: run.f90:575
        o.update_ghosts(f)
        if it in rows:
            got = h.format_reduced(d, cfg["eta"], rows[it])
            assert got == {k: rows[it][k] for k in got}, (it, got, rows[it])
            assert "%.3f" % t0 == rows[it]["t"] and "%.2E" % dt == rows[it]["dt"]


def test_ran0_stream_equals_scalar_generator():
    import nr_f90
    for seed in (-10, -41, 1812):
        r = nr_f90.Ran0(seed)
        assert np.array_equal(np.array([r.next() for _ in range(3000)]), nr_f90.ran0_stream(seed, 3000))
