"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI with host buffers, against the
CPU oracle on identical seeded inputs.  Tolerances are those of BASELINE.json's north_star:
f-array after 1 substep <= 1e-12 relative L-inf, after 100 steps <= 1e-9, time series likewise."""
import numpy as np
import pytest

from pencil_code_b200 import config as pcfg
from pencil_code_b200 import synthetic

import parity_util as pu

pytestmark = pytest.mark.gpu

TOL_SUBSTEP = 1e-12
TOL_100 = 1e-9

CASES = {
    "iso_mhd": dict(n=32, entropy=False, magnetic=True, nu=5e-3, eta=5e-3),
    "mhd_entropy": dict(n=32, entropy=True, magnetic=True, nu=5e-3, eta=5e-3, chi=5e-3),
    "hydro_entropy": dict(n=24, entropy=True, magnetic=False, nu=5e-3, chi=2e-3),
    "hydro_hyper3": dict(n=32, entropy=False, magnetic=False, hyper3=True),
    "mhd_entropy_upwind": dict(n=24, entropy=True, magnetic=True, nu=5e-3, eta=5e-3, chi=5e-3, upwind=True),
    "ragged": dict(n=37, ny=19, nz=11, entropy=True, magnetic=True, nu=4e-3, eta=3e-3, chi=2e-3),
}


def _cfg(name, **over):
    kw = dict(CASES[name])
    n = kw.pop("n")
    ent, mag = kw.pop("entropy"), kw.pop("magnetic")
    c = pcfg.make_config(n, ent, mag, cdt=0.4, cdtv=0.3, **kw)
    c.update(over)
    return c


VARIANTS = [1, 4, 5, 10]


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("name", sorted(CASES))
def test_one_substep(oracle_mod, name, variant):
    """variant 1 = global-load kernel (fallback for odd mx, cross-check), 4 = TMA/shared-memory z-marching kernel with
    lead/trail z-mixed derivatives (anisotropic grids), 5 = its isotropic-spacing specialisation (round-1 default),
    9 = all-trail z-mixed derivatives with delayed finalisation, z histories and accumulators in tensor memory (default)"""
    cfg = _cfg(name)
    if variant >= 2 and cfg["nx"] % 2:
        pytest.skip("TMA variant needs 16-byte row strides (even mx); odd grids use variant 1")
    fg, fc, _, _, _ = pu.run_both(cfg, oracle_mod, 0, dt=pcfg.fixed_dt(cfg), substeps_only=1, kernel_variant=variant)
    err = pu.rel_linf(fg, fc)
    assert err.max() <= TOL_SUBSTEP, (name, variant, err)


@pytest.mark.parametrize("variant", [4, 5, 10])
@pytest.mark.parametrize("shape", [(32, 8, 4), (64, 16, 40), (70, 9, 5), (6, 6, 6), (34, 10, 17), (40, 24, 3), (32, 8, 100)])
def test_tma_variants_tile_edges(oracle_mod, shape, variant):
    """grids that are not multiples of the 32x8 tile, shorter than one tile, the minimum nz = 3, and z-chunk remainders"""
    nx, ny, nz = shape
    cfg = pcfg.make_config(nx, True, True, ny=ny, nz=nz, nu=5e-3, eta=4e-3, chi=3e-3, cdt=0.4, cdtv=0.3)
    cfg.update(dx=0.2, dy=0.21, dz=0.19)
    if variant >= 5:  # variants 5 and 10 are isotropic-spacing specialisations (anisotropic grids fall back to 4)
        cfg.update(dx=0.2, dy=0.2, dz=0.2)
    fg, fc, _, _, _ = pu.run_both(cfg, oracle_mod, 0, dt=0.01, substeps_only=3, kernel_variant=variant)
    assert pu.rel_linf(fg, fc).max() <= 5e-12, (shape, variant)


@pytest.mark.parametrize("name", ["iso_mhd", "mhd_entropy", "ragged"])
def test_three_substeps_and_ghosts(oracle_mod, name):
    cfg = _cfg(name)
    fg, fc, _, _, _ = pu.run_both(cfg, oracle_mod, 0, dt=pcfg.fixed_dt(cfg), substeps_only=3)
    err = pu.rel_linf(fg, fc)
    assert err.max() <= 5e-12, (name, err)
    # copy_farray_from_GPU returns valid periodic ghost zones (incl. edges and corners)
    g = fg.copy()
    nz, ny, nx = (s - 6 for s in g.shape[1:])
    g[..., :3] = g[..., nx:nx + 3]; g[..., nx + 3:] = g[..., 3:6]
    g[..., :3, :] = g[..., ny:ny + 3, :]; g[..., ny + 3:, :] = g[..., 3:6, :]
    g[..., :3, :, :] = g[..., nz:nz + 3, :, :]; g[..., nz + 3:, :, :] = g[..., 3:6, :, :]
    assert np.array_equal(g, fg)


def test_nonisotropic_grid_spacing(oracle_mod):
    cfg = _cfg("mhd_entropy")
    cfg.update(dx=0.21, dy=0.17, dz=0.26)
    fg, fc, _, _, _ = pu.run_both(cfg, oracle_mod, 0, dt=0.01, substeps_only=2)
    assert pu.rel_linf(fg, fc).max() <= 2e-12


@pytest.mark.parametrize("name", ["iso_mhd", "mhd_entropy"])
def test_100_steps_fixed_dt_with_time_series(oracle_mod, name):
    cfg = _cfg(name)
    fg, fc, sg, sc, (dtg, dtc, tg, tc) = pu.run_both(cfg, oracle_mod, 100, dt=pcfg.fixed_dt(cfg), diag_every=10)
    err = pu.rel_linf(fg, fc)
    assert err.max() <= TOL_100, (name, err)
    assert abs(tg - tc) <= 1e-12 * abs(tc)
    keys = ["urms", "umax", "rhom", "dt"] + (["brms", "bmax", "jrms", "abm", "jbm"] if cfg["iax"] >= 0 else [])
    for a, b in zip(sg, sc):
        for k in keys:
            assert abs(a[k] - b[k]) <= TOL_100 * max(abs(b[k]), 1e-300), (k, a[k], b[k])


def test_100_steps_courant_dt(oracle_mod):
    """ldt=T: dt comes from the CFL reduction inside substep 1 of every step (equ.f90:408)."""
    cfg = _cfg("mhd_entropy")
    fg, fc, sg, sc, (dtg, dtc, tg, tc) = pu.run_both(cfg, oracle_mod, 100, dt=0.0, ldt=True, diag_every=20)
    assert abs(dtg - dtc) <= 1e-12 * dtc
    assert abs(tg - tc) <= 1e-10 * tc
    assert pu.rel_linf(fg, fc).max() <= TOL_100
    for a, b in zip(sg, sc):
        for k in ("urms", "brms", "dt", "dtu", "dtc", "dtb", "dtnu", "epsK", "oum"):
            assert abs(a[k] - b[k]) <= TOL_100 * max(abs(b[k]), 1e-300), (k, a[k], b[k])


def test_helical_forcing_30_steps(oracle_mod):
    cfg = _cfg("iso_mhd")
    kv = pcfg.kvectors()
    fg, fc, _, _, _ = pu.run_both(cfg, oracle_mod, 30, dt=pcfg.fixed_dt(cfg), forcing=kv)
    assert pu.rel_linf(fg, fc).max() <= TOL_100
    # the forcing did something: compare with an unforced run
    fg0, _, _, _, _ = pu.run_both(cfg, oracle_mod, 30, dt=pcfg.fixed_dt(cfg))
    assert np.abs(pu.interior(fg)[:3] - pu.interior(fg0)[:3]).max() > 1e-4


def test_helical_forcing_300_steps_64_and_no_host_sync(oracle_mod):
    """300 forced steps at 64^3 against the oracle, and the fixed-dt time loop queues its steps without a single
    blocking host synchronisation (dt is uploaded only when it changes, the forcing tables are built on the device)."""
    from pencil_code_b200.gpu import GPU, HelicalForcing
    cfg = pcfg.make_config(64, False, True, nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3)
    kv = pcfg.kvectors()
    fg, fc, _, _, _ = pu.run_both(cfg, oracle_mod, 300, dt=pcfg.fixed_dt(cfg), forcing=kv)
    assert pu.rel_linf(fg, fc).max() <= TOL_100
    f = synthetic.initial_condition(cfg)
    g = GPU(cfg)
    g.initialize_GPU(f)
    try:
        hf = HelicalForcing(kv)
        dt, t = pcfg.fixed_dt(cfg), 0.0
        _, t = g.time_step(f, dt, False, t, forcing=hf.draw(dt))  # first step uploads dt
        n0 = g.host_sync_count()
        for _ in range(20):
            _, t = g.time_step(f, dt, False, t, forcing=hf.draw(dt))
        assert g.host_sync_count() == n0
    finally:
        g.finalize_GPU()


def test_hyper3_20_steps(oracle_mod):
    cfg = _cfg("hydro_hyper3")
    fg, fc, _, _, _ = pu.run_both(cfg, oracle_mod, 20, dt=pcfg.fixed_dt(cfg))
    assert pu.rel_linf(fg, fc).max() <= TOL_100


def test_config1_helical_mhd_128_three_steps(oracle_mod):
    """BASELINE config[1] at full size (128^3 isothermal MHD): 3 full steps against the oracle."""
    cfg = pcfg.config_helical_mhd_128()
    fg, fc, sg, sc, _ = pu.run_both(cfg, oracle_mod, 3, dt=pcfg.fixed_dt(cfg), diag_every=1)
    assert pu.rel_linf(fg, fc).max() <= 1e-11
    for a, b in zip(sg, sc):
        for k in ("urms", "brms", "jrms"):
            assert abs(a[k] - b[k]) <= 1e-11 * abs(b[k])


def test_device_synthetic_ic_matches_numpy(oracle_mod):
    """pcb200_init_synthetic (bench workload generator) == pencil_code_b200.synthetic up to libm ulps"""
    from pencil_code_b200.gpu import GPU
    cfg = _cfg("mhd_entropy")
    f = np.zeros(synthetic.farray_shape(cfg))
    g = GPU(cfg)
    g.initialize_GPU(f, nt=0)
    g.init_synthetic(12345)
    g.copy_farray_from_GPU(f)
    g.finalize_GPU()
    ref = synthetic.initial_condition(cfg, seed=12345)
    assert np.abs(pu.interior(f) - pu.interior(ref)).max() < 1e-15


def test_full_size_512_independent_kernels_agree_and_ghosts_wrap():
    """BASELINE.json's full size (512^3, 8 variables, 26.7 GB on the device), size-independent properties:
    (1) the default TMA/shared-memory kernel and the global-load kernel (variant 1, an independent access path that
    is oracle-checked at small sizes) agree to the one-substep tolerance after a full RK3 step with forcing on the
    device-generated workload; (2) the ghost zones handed back are the periodic images of the interior."""
    import torch
    from pencil_code_b200.gpu import GPU, HelicalForcing
    if torch.cuda.mem_get_info()[1] < 60e9:
        pytest.skip("needs a B200-class device (26.7 GB per instance)")
    cfg = pcfg.config_mhd_entropy(512)
    dt = pcfg.fixed_dt(cfg)
    out = []
    for variant in (0, 1):
        c = dict(cfg, kernel_variant=variant)
        f = np.zeros(synthetic.farray_shape(c))
        g = GPU(c)
        g.initialize_GPU(f, nt=0)
        g.init_synthetic(12345)
        g.time_step(f, dt, False, 0.0, forcing=HelicalForcing(pcfg.kvectors()).draw(dt))
        g.copy_farray_from_GPU(f)
        g.finalize_GPU()
        out.append(f)
    a, b = out
    assert np.isfinite(a).all()
    for k in range(a.shape[0]):
        ia, ib = a[k, 3:-3, 3:-3, 3:-3], b[k, 3:-3, 3:-3, 3:-3]
        err = np.abs(ia - ib).max() / np.abs(ib).max()
        assert err <= TOL_SUBSTEP, (k, err)
    n = 512
    assert np.array_equal(a[..., :3], a[..., n:n + 3]) and np.array_equal(a[..., n + 3:], a[..., 3:6])
    assert np.array_equal(a[:, :, :3, :], a[:, :, n:n + 3, :]) and np.array_equal(a[:, :, n + 3:, :], a[:, :, 3:6, :])
    assert np.array_equal(a[:, :3], a[:, n:n + 3]) and np.array_equal(a[:, n + 3:], a[:, 3:6])



def test_full_size_512_one_substep_against_the_oracle_on_z_windows(oracle_mod):
    """BASELINE config[2] at its full size against the ORACLE (not against another CUDA kernel): one substep of the
    512^3 MHD+entropy run on the device-generated state; the oracle evaluates the RHS on three 16-plane z windows
    (bottom, middle, top of the box) whose ghost planes are the neighbouring planes of the same host array --
    pc_oracle_rhs only needs valid ghosts, so a window is an exact sub-problem for one substep."""
    import torch
    from pencil_code_b200.gpu import GPU
    if torch.cuda.mem_get_info()[1] < 60e9:
        pytest.skip("needs a B200-class device (26.7 GB per instance)")
    n, w = 512, 16
    cfg = pcfg.config_mhd_entropy(n)
    dt = pcfg.fixed_dt(cfg)
    f0 = np.zeros(synthetic.farray_shape(cfg))
    g = GPU(cfg)
    g.initialize_GPU(f0, nt=0)
    g.init_synthetic(12345)
    g.copy_farray_from_GPU(f0)  # state before the substep, ghost zones filled
    g.set_dt(dt, False)
    g.rhs_GPU(f0, 1)
    f1 = np.empty_like(f0)
    g.copy_farray_from_GPU(f1)
    g.finalize_GPU()
    wcfg = dict(cfg, nz=w)
    o = oracle_mod.Oracle(wcfg)
    beta1 = 1.0 / 3.0
    for z0 in (3, 3 + n // 2 - w // 2, 3 + n - w):  # first interior plane of the window (array index)
        fw = np.ascontiguousarray(f0[:, z0 - 3:z0 + w + 3])
        dfw = np.zeros_like(fw)
        o.rhs(fw, dfw)
        ref = fw + dt * beta1 * dfw
        got = f1[:, z0 - 3:z0 + w + 3]
        err = pu.rel_linf(np.ascontiguousarray(got), ref)
        assert err.max() <= TOL_SUBSTEP, (z0, err)


@pytest.mark.parametrize("which", ["hydro_hyper3_256", "mhd_entropy_256"])
def test_configs_at_256_three_steps_against_the_oracle(oracle_mod, which):
    """BASELINE config[4] (hydro + hyperdiffusion, 256^3) at its full size and config[2]'s physics at 256^3: three
    forced RK3 steps, whole domain, against the oracle."""
    cfg = pcfg.config_hydro_hyper3(256) if which == "hydro_hyper3_256" else pcfg.config_mhd_entropy(256)
    fg, fc, _, _, _ = pu.run_both(cfg, oracle_mod, 3, dt=pcfg.fixed_dt(cfg), forcing=pcfg.kvectors())
    assert pu.rel_linf(fg, fc).max() <= 1e-11


def test_hyper3_simplified_viscosity(oracle_mod):
    """ivisc='hyper3-simplified' (viscosity.f90:1982-1989: fvisc += nu_hyper3 del6 u) alone and together with nu-const"""
    for extra in (dict(lvisc_nu_const=0, nu=0.0), dict(lvisc_nu_const=1, nu=3e-3)):
        cfg = _cfg("hydro_hyper3", lvisc_hyper3_simplified=1, lvisc_hyper3_nu_const=0, nu_hyper3=1e-6, **extra)
        fg, fc, _, _, _ = pu.run_both(cfg, oracle_mod, 0, dt=pcfg.fixed_dt(cfg), substeps_only=1)
        assert pu.rel_linf(fg, fc).max() <= TOL_SUBSTEP, extra
        fg, fc, _, _, _ = pu.run_both(cfg, oracle_mod, 20, dt=pcfg.fixed_dt(cfg))
        assert pu.rel_linf(fg, fc).max() <= TOL_100, extra


@pytest.mark.parametrize("name", ["mhd_entropy_upwind", "ragged"])
def test_upwind_and_ragged_20_steps(oracle_mod, name):
    """upwinded advection of u, lnrho, s and the upwinded induction term; odd mx (global-load kernel): 20 steps"""
    cfg = _cfg(name)
    fg, fc, sg, sc, _ = pu.run_both(cfg, oracle_mod, 20, dt=pcfg.fixed_dt(cfg), diag_every=5)
    assert pu.rel_linf(fg, fc).max() <= TOL_100
    for a, b in zip(sg, sc):
        for k in ("urms", "brms", "jrms"):
            assert abs(a[k] - b[k]) <= TOL_100 * abs(b[k])


def test_config1_helical_mhd_128_hundred_steps(oracle_mod):
    """BASELINE config[1] at full size, the north-star's long-run criterion: 100 forced steps, f-array and the
    urms/brms/dt time series within 1e-9 of the oracle (about 2 minutes of oracle time on 16 cores)."""
    cfg = pcfg.config_helical_mhd_128()
    fg, fc, sg, sc, _ = pu.run_both(cfg, oracle_mod, 100, dt=pcfg.fixed_dt(cfg), forcing=pcfg.kvectors(), diag_every=10)
    err = pu.rel_linf(fg, fc)
    print("config[1] 128^3, 100 steps: rel Linf per variable", err)
    assert err.max() <= TOL_100
    for a, b in zip(sg, sc):
        for k in ("urms", "brms", "jrms"):
            assert abs(a[k] - b[k]) <= TOL_100 * abs(b[k])


def test_async_copy_is_a_snapshot(tmp_path):
    """pcb200_copy_farray_async: the host array receives the state at the time of the call even though the time loop
    keeps running while the transfer is in flight; written as var.dat it restarts a run bit-identically"""
    from pencil_code_b200 import snapshot
    from pencil_code_b200.gpu import GPU
    cfg = pcfg.make_config(64, True, True, nu=5e-3, eta=5e-3, chi=5e-3)
    f = synthetic.initial_condition(cfg)
    g = GPU(cfg)
    g.initialize_GPU(f)
    dt, t = pcfg.fixed_dt(cfg), 0.0
    _, t = g.time_step(f, dt, False, t)
    ref = np.empty_like(f)
    g.copy_farray_from_GPU(ref)            # synchronous copy of the state after step 1
    snap = np.zeros_like(f)
    g.copy_farray_from_GPU_async(snap)     # asynchronous snapshot of the same state ...
    for _ in range(3):                     # ... while three more steps run
        _, t = g.time_step(f, dt, False, t)
    g.copy_farray_wait()
    later = np.empty_like(f)
    g.copy_farray_from_GPU(later)
    g.finalize_GPU()
    assert np.array_equal(snap, ref)
    assert not np.array_equal(later, ref)
    # restart from the snapshot file (rsnap -> load_farray_to_GPU): three more steps reproduce `later`
    path = snapshot.snapshot_path(str(tmp_path), 0)
    snapshot.output_snap(path, snap, dt, cfg)
    f2, t2, _, _ = snapshot.input_snap(path, cfg)
    assert t2 == dt
    g = GPU(cfg)
    g.initialize_GPU(f2)
    for _ in range(3):
        _, t2 = g.time_step(f2, dt, False, t2)
    again = np.empty_like(f)
    g.copy_farray_from_GPU(again)
    g.finalize_GPU()
    assert np.array_equal(again, later)


def test_zero_velocity_uniform_state_is_steady(oracle_mod):
    """edge case: uniform state has zero RHS (up to FMA-contraction rounding of the der2 weights)"""
    from pencil_code_b200.gpu import GPU
    cfg = _cfg("mhd_entropy")
    f = np.zeros(synthetic.farray_shape(cfg))
    f[3] = 0.5
    f0 = f.copy()
    g = GPU(cfg)
    g.initialize_GPU(f)
    g.time_step(f, 0.01, False, 0.0)
    g.copy_farray_from_GPU(f)
    g.finalize_GPU()
    assert np.abs(f - f0).max() <= 1e-15


@pytest.mark.parametrize("overlap", [True, False])
def test_p2p_put_path_loopback(oracle_mod, monkeypatch, overlap):
    """The peer-memory z halo on ONE GPU (PCB200_LOOPBACK_P2P=1: this rank is both neighbours): copy-engine puts
    into the ghost planes, the arrival counters and the cuStreamWaitValue64 ordering of the boundary kernels must
    reproduce the periodic single-domain result, Courant dt included."""
    monkeypatch.setenv("PCB200_LOOPBACK", "1")
    monkeypatch.setenv("PCB200_LOOPBACK_P2P", "1")
    if not overlap:
        monkeypatch.setenv("PCB200_NO_OVERLAP", "1")
    cfg = pcfg.make_config(34, True, True, ny=10, nz=17, nu=5e-3, eta=4e-3, chi=3e-3, cdt=0.4, cdtv=0.3)
    fg, fc, sg, sc, (dtg, dtc, tg, tc) = pu.run_both(cfg, oracle_mod, 4, dt=0.0, ldt=True, diag_every=2)
    assert pu.rel_linf(fg, fc).max() <= 1e-11
    assert abs(dtg - dtc) <= 1e-12 * dtc
    for a, b in zip(sg, sc):
        for k in ("urms", "brms", "jrms", "dtu"):
            assert abs(a[k] - b[k]) <= 1e-11 * abs(b[k])


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("overlap", [True, False])
def test_exchange_code_path_loopback(oracle_mod, monkeypatch, variant, overlap):
    """The multi-GPU code path on ONE GPU: PCB200_LOOPBACK=1 makes every neighbour this rank, so the y-face
    pack/unpack kernels, the plane transfers on the communication stream and the interior/boundary split of the
    fused kernel (overlap) all run and must reproduce the periodic single-domain result."""
    monkeypatch.setenv("PCB200_LOOPBACK", "1")
    if not overlap:
        monkeypatch.setenv("PCB200_NO_OVERLAP", "1")
    cfg = pcfg.make_config(34, True, True, ny=10, nz=17, nu=5e-3, eta=4e-3, chi=3e-3, cdt=0.4, cdtv=0.3)
    fg, fc, sg, sc, (dtg, dtc, tg, tc) = pu.run_both(cfg, oracle_mod, 4, dt=0.0, ldt=True, diag_every=2,
                                                     kernel_variant=variant)
    assert pu.rel_linf(fg, fc).max() <= 1e-11
    assert abs(dtg - dtc) <= 1e-12 * dtc
    for a, b in zip(sg, sc):
        for k in ("urms", "brms", "jrms", "dtu"):
            assert abs(a[k] - b[k]) <= 1e-11 * abs(b[k])
    # ghost zones returned by copy_farray_from_GPU are the periodic images (x wrap, y faces, z planes + corners)
    g = fg.copy()
    nz, ny, nx = (s - 6 for s in g.shape[1:])
    g[..., :3] = g[..., nx:nx + 3]; g[..., nx + 3:] = g[..., 3:6]
    g[..., :3, :] = g[..., ny:ny + 3, :]; g[..., ny + 3:, :] = g[..., 3:6, :]
    g[..., :3, :, :] = g[..., nz:nz + 3, :, :]; g[..., nz + 3:, :, :] = g[..., 3:6, :, :]
    assert np.array_equal(g, fg)


def test_multi_gpu_decomposition_invariance():
    """N>=2 GPUs: z-slabs and yz-pencils over NCCL against the single-domain oracle (tests/multi_gpu_check.py
    under torchrun); skipped on a single-GPU box (the gloo tests cover the host logic there)."""
    import subprocess
    import sys
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 4 if n >= 4 else 2
    script = __file__.replace("test_parity_gpu.py", "multi_gpu_check.py")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                        "--master-addr", "127.0.0.1", "--master-port", "29533", script], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0 and "MULTI_GPU_CHECK PASS" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_gputest_20_rows_of_the_reference_cpu_binary_on_the_gpu():
    """The reference's own known-answer run (samples/gputest/cpu-test.out:275-294, tests/golden/gputest_cpu_rows.txt)
    through the C ABI: 20 RK3 steps, get_gpu_reduced_vars formatted with print.in's formats must reproduce every digit
    the reference's CPU binary printed."""
    from pencil_code_b200.gpu import GPU
    from test_oracle_golden import _gputest_setup, format_gputest_row, gputest_rows
    cfg, f = _gputest_setup()
    g = GPU(cfg)
    g.initialize_GPU(f)
    t, dt = 0.0, 1e-3
    try:
        for row in gputest_rows():
            dt, t = g.time_step(f, dt, False, t, diagnostics=True)
            assert format_gputest_row(g.get_gpu_reduced_vars()) == row[2:6], row
    finally:
        g.finalize_GPU()
