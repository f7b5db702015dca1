"""GPU tests (-m gpu) of the boundary towards the Fortran driver: entry points of include/pcb200.h that the time loop
of run.f90 / equ.f90 reaches besides rhs_gpu -- the C++ driver mirror, RELOAD, run-time parameter pushes, device
pointers, the ltest_bcs / ltest_rhs test modes -- each against the oracle or an independent path."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from pencil_code_b200 import config as pcfg
from pencil_code_b200 import synthetic
from pencil_code_b200.gpu import GPU, HelicalForcing, LIB_PATH

import parity_util as pu

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cfg(n=32, **kw):
    return pcfg.make_config(n, True, True, nu=5e-3, eta=5e-3, chi=5e-3, cdt=0.4, cdtv=0.3, **kw)


def _oracle_forcing(oracle_mod, c):
    f = oracle_mod.Forcing()
    for i in range(3):
        f.coef1[i], f.coef2[i], f.fda[i], f.kk[i] = c.coef1[i], c.coef2[i], c.fda[i], c.kk[i]
    f.phase, f.fact, f.k1_ff = c.phase, c.fact, c.k1_ff
    return f


def test_driver_run_equals_module_calls_and_oracle(oracle_mod):
    """pcb200_driver_run (C++ mirror of run.f90's loop: forcing draw, set_dt, 3 x {before_boundary, rhs}, diagnostics
    every it1 steps) == the same calls made one by one through the module-GPU mirror (bitwise), and both follow the
    oracle fed with the same forcing coefficients."""
    cfg = _cfg()
    dt = pcfg.fixed_dt(cfg)
    f0 = synthetic.initial_condition(cfg, seed=7)
    nt = 6
    # (a) the C++ driver
    fa = f0.copy()
    g = GPU(cfg); g.initialize_GPU(fa)
    ta, dta, ita, rows = g.driver_run(nt, dt, forcing=HelicalForcing(pcfg.kvectors()), it1=2)
    g.copy_farray_from_GPU(fa); g.finalize_GPU()
    assert ita == nt and [r[0] for r in rows] == [1, 3, 5]
    # (b) the same sequence through gpu.py, (c) the oracle with the same coefficients
    fb = f0.copy()
    g = GPU(cfg); g.initialize_GPU(fb)
    fs = HelicalForcing(pcfg.kvectors())
    o = oracle_mod.Oracle(cfg)
    fc, dfc = f0.copy(), np.zeros_like(f0)
    tb = tc = 0.0
    series = []
    for it in range(1, nt + 1):
        c = fs.draw(dt)
        want = (it - 1) % 2 == 0
        _, tb = g.time_step(fb, dt, False, tb, forcing=c, diagnostics=want)
        _, tc, d = o.time_step(fc, dfc, False, dt, tc, forcing=_oracle_forcing(oracle_mod, c), diag=want)
        if want:
            series.append(d)
    g.copy_farray_from_GPU(fb); g.finalize_GPU()
    o.update_ghosts(fc)
    assert np.array_equal(fa, fb) and ta == tb
    assert pu.rel_linf(fa, fc).max() <= 1e-11
    for (it, t, r), d in zip(rows, series):
        for k in ("urms", "umax", "rhom", "brms", "jrms", "abm", "ssm"):
            assert abs(r[k] - d[k]) <= 1e-11 * max(abs(d[k]), 1e-300), (it, k)


def test_update_on_gpu_and_reload_change_the_physics_mid_run(oracle_mod):
    """update_on_gpu('nu') (updateInConfigScalName) and RELOAD (pcb200_reconfigure + reload_gpu_config_c) while the run
    is initialized: the following steps use the new values, as an oracle run that changes them at the same step does."""
    cfg = _cfg()
    dt = pcfg.fixed_dt(cfg)
    f0 = synthetic.initial_condition(cfg, seed=11)
    fg = f0.copy()
    g = GPU(cfg); g.initialize_GPU(fg)
    o = oracle_mod.Oracle(cfg)
    fc, dfc = f0.copy(), np.zeros_like(f0)
    tg = tc = 0.0
    for phase in range(3):
        for _ in range(2):
            _, tg = g.time_step(fg, dt, False, tg)
            _, tc, _ = o.time_step(fc, dfc, False, dt, tc)
        if phase == 0:
            assert g.update_on_gpu(0, "nu", 2e-2) >= 0
            o.p.nu = 2e-2
            with pytest.raises(KeyError):
                g.update_on_gpu(0, "no_such_parameter", 1.0)
        elif phase == 1:
            g.reconfigure(eta=1.5e-2, chi=1e-3, lviscosity_heat=0)
            o.p.eta, o.p.chi, o.p.lviscosity_heat = 1.5e-2, 1e-3, 0
    g.copy_farray_from_GPU(fg); g.finalize_GPU()
    o.update_ghosts(fc)
    assert pu.rel_linf(fg, fc).max() <= 1e-11
    # the changes were not no-ops: an unchanged run differs
    fu = f0.copy()
    g = GPU(cfg); g.initialize_GPU(fu)
    t = 0.0
    for _ in range(6):
        _, t = g.time_step(fu, dt, False, t)
    g.copy_farray_from_GPU(fu); g.finalize_GPU()
    assert pu.rel_linf(fu, fc).max() > 1e-6


def test_device_pointers_address_the_variable_slots():
    """get_farray_ptr_gpu / pos_real_ptr_c (gpu_astaroth_ansi.c): the returned device addresses are the f-array slots:
    reading them back with cudaMemcpy gives the host values (row pitch of the device layout taken into account)."""
    cfg = _cfg(24)
    f = synthetic.initial_condition(cfg, seed=3)
    g = GPU(cfg); g.initialize_GPU(f)
    g.copy_farray_from_GPU(f)  # host copy with the ghost zones the device filled
    p_in, p_out = g.get_farray_ptr_gpu()
    assert p_in and p_out and p_in != p_out
    rt = C.CDLL("libcudart.so.12")
    rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
    mx = cfg["nx"] + 6
    for ind in (0, 3, 7):
        p = g.pos_real_ptr(p_in, ind)
        row = np.zeros(mx)
        assert rt.cudaMemcpy(row.ctypes.data, C.c_void_p(p), mx * 8, 2) == 0
        assert np.array_equal(row, f[ind, 0, 0, :])
    g.finalize_GPU()


@pytest.mark.parametrize("mode", ["plain", "loopback", "loopback_p2p"])
def test_bcs_check_of_the_reference_test_mode(monkeypatch, mode):
    """ltest_bcs -> test_bcs_c (testBCs, astaroth/gpu_astaroth.cc:2283-2454): randomised mesh, device ghost/halo path
    against the CPU boundary conditions to 1e-12; also through the halo-exchange code path on one GPU.  The state of
    the run is not disturbed."""
    if mode != "plain":
        monkeypatch.setenv("PCB200_LOOPBACK", "1")
    if mode == "loopback_p2p":
        monkeypatch.setenv("PCB200_LOOPBACK_P2P", "1")
    cfg = _cfg(24, ny=20, nz=18)
    f = synthetic.initial_condition(cfg, seed=5)
    g = GPU(cfg); g.initialize_GPU(f)
    dt = pcfg.fixed_dt(cfg)
    g.time_step(f, dt, False, 0.0)
    before = np.empty_like(f); g.copy_farray_from_GPU(before)
    assert g.test_bcs_count() == 0
    g.test_gpu_bcs()  # would exit(EXIT_FAILURE) on failure
    after = np.empty_like(f); g.copy_farray_from_GPU(after)
    assert np.array_equal(before, after)
    g.finalize_GPU()


def test_rhs_replay_in_c_through_the_fortran_symbols(tmp_path, oracle_mod):
    """ltest_rhs (equ.f90:1974-2143 test_rhs_gpu) replayed by a C program that dlopen()s libpcb200.so and resolves ONLY
    the underscore-mangled names the reference's Fortran binds (FTNIZE): all three substeps on a CPU copy (oracle as
    cpu_version) and on the device, per-variable maximum relative difference printed in the reference's format."""
    import pc_oracle
    cfg = _cfg(32)
    src = os.path.join(ROOT, "tests", "c", "test_rhs_replay.c")
    exe = str(tmp_path / "test_rhs_replay")
    subprocess.check_call(["gcc", "-O1", "-std=c11", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "oracle"),
                           "-o", exe, src, "-ldl", "-lm"])
    from pencil_code_b200.gpu import make_struct
    c = make_struct(cfg)
    (tmp_path / "config.bin").write_bytes(bytes(c))
    (tmp_path / "oracle_pars.bin").write_bytes(bytes(pc_oracle.pars_from_dict(cfg)))
    f0 = synthetic.initial_condition(cfg, seed=21)
    f0.tofile(tmp_path / "f0.bin")
    r = subprocess.run([exe, LIB_PATH, pc_oracle.build(), str(tmp_path / "config.bin"), str(tmp_path / "oracle_pars.bin"),
                        str(tmp_path / "f0.bin"), repr(pcfg.fixed_dt(cfg))], capture_output=True, text=True, timeout=300)
    print(r.stdout[-3000:], r.stderr[-2000:])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "Max comp diff:" in r.stdout and "Passed BC test" in r.stderr
    worst = max(float(l.split(":")[-1]) for l in r.stdout.splitlines() if l.startswith(" rel Linf"))
    assert worst <= 1e-12
