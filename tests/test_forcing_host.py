"""CPU tests of the library's host-side helical forcing (no GPU needed: pcb200_forcing_coefs_hel is plain C++):
the product's coefficient routine and splitmix64 draw order against the oracle's restatement of
fconst_coefs_hel (forcing.f90:1773-2001, draws :1804-1811)."""
import math
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from pencil_code_b200 import config as pcfg  # noqa: E402
from pencil_code_b200.gpu import HelicalForcing  # noqa: E402


@pytest.fixture(scope="module")
def oracle():
    import pc_oracle
    return pc_oracle.Oracle(pcfg.make_config(16, True, True, nu=1e-2, eta=1e-2, chi=1e-2))


def test_kvector_shell_and_kav():
    kv = pcfg.kvectors()
    assert all(4.5 <= math.sqrt(a * a + b * b + c * c) < 5.5 for a, b, c in kv)
    assert len(kv) == len(set(kv)) and len(kv) > 100
    frc = HelicalForcing(kv)
    assert abs(frc.kav - np.mean([math.sqrt(a * a + b * b + c * c) for a, b, c in kv])) < 1e-14


def test_coefficients_match_oracle(oracle):
    kv = pcfg.kvectors()
    frc = HelicalForcing(kv, force=0.07, relhel=1.0, seed=2024)
    twin = HelicalForcing(kv, force=0.07, relhel=1.0, seed=2024)  # replays the same uniform stream
    dt = 3.1e-3
    for _ in range(25):
        c = frc.draw(dt)
        f1, f2, f3 = twin._uniform(), twin._uniform(), twin._uniform()
        phase = math.pi * (2 * f1 - 1.0)
        ik = int(len(kv) * (0.9999 * f2) + 1) - 1
        phi = f3 * 2 * math.pi
        assert c.phase == phase and list(c.kk) == list(kv[ik])
        ref = oracle.forcing_coefs(0.07, 1.0, frc.kav, 1.0, dt, kv[ik], phase, phi)
        for name in ("coef1", "coef2", "fda"):
            a, b = np.array(getattr(c, name)), np.array(list(getattr(ref, name)))
            assert np.abs(a - b).max() <= 1e-14 * max(1.0, np.abs(b).max()), name
        assert abs(c.fact - ref.fact) <= 1e-15 * abs(ref.fact)
        # solenoidal, normalised: the forcing direction is perpendicular to k (forcing.f90:1880-1960)
        k = np.array(c.kk)
        assert abs(np.dot(k, np.array(c.coef1))) <= 1e-10 * np.linalg.norm(k) * np.linalg.norm(c.coef1)


def test_stream_is_reproducible_and_uniform():
    frc = HelicalForcing(pcfg.kvectors(), seed=7)
    u = np.array([frc._uniform() for _ in range(20000)])
    assert 0.0 <= u.min() and u.max() < 1.0
    assert abs(u.mean() - 0.5) < 0.01 and abs(u.var() - 1 / 12) < 0.005
    frc2 = HelicalForcing(pcfg.kvectors(), seed=7)
    assert [frc2._uniform() for _ in range(10)] == list(u[:10])
