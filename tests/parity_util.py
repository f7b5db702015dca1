"""Shared helpers of the GPU parity tests: run the same steps through the C ABI (GPU) and through
the oracle (CPU) from identical initial conditions and compare."""
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from pencil_code_b200 import synthetic  # noqa: E402
from pencil_code_b200.gpu import GPU  # noqa: E402

NG = 3


def interior(f):
    return f[:, NG:-NG, NG:-NG, NG:-NG]


def rel_linf(a, b):
    """relative L-infinity per variable: max|a-b| / max|b| over the interior"""
    out = []
    for k in range(a.shape[0]):
        den = np.abs(interior(b)[k]).max()
        num = np.abs(interior(a)[k] - interior(b)[k]).max()
        out.append(num / den if den > 0 else num)
    return np.array(out)


class ForcingStream:
    """host RNG + coefficient generation shared by both arms (forcing.f90:1804-1811 draw order)."""

    def __init__(self, oracle, kvecs, force=0.07, relhel=1.0, seed=2024):
        self.o, self.kvecs, self.force, self.relhel = oracle, kvecs, force, relhel
        self.rng = np.random.default_rng(seed)
        self.kav = float(np.mean([math.sqrt(sum(c * c for c in k)) for k in kvecs]))

    def draw(self, dt):
        f1, f2, f3 = self.rng.random(3)
        phase = math.pi * (2 * f1 - 1.0)
        ik = int(len(self.kvecs) * (0.9999 * f2) + 1) - 1
        phi = f3 * 2 * math.pi
        return self.o.forcing_coefs(self.force, self.relhel, self.kav, 1.0, dt, self.kvecs[ik], phase, phi)


def run_both(cfg, oracle_mod, nsteps, dt=None, ldt=False, forcing=None, substeps_only=None, seed=12345,
             diag_every=0, kernel_variant=None, f0=None):
    """Returns (f_gpu, f_cpu, series_gpu, series_cpu).  substeps_only=k stops after k substeps of step 1."""
    cfg = dict(cfg)
    if kernel_variant is not None:
        cfg["kernel_variant"] = kernel_variant
    o = oracle_mod.Oracle(cfg)
    f0 = synthetic.initial_condition(cfg, seed=seed) if f0 is None else f0
    # --- GPU arm through the module-GPU mirror / C ABI (host buffers)
    fg = f0.copy()
    g = GPU(cfg)
    g.register_GPU()
    g.initialize_GPU(fg)
    # --- CPU arm
    fc = f0.copy()
    dfc = np.zeros_like(fc)
    sg, sc = [], []
    dtg = dtc = (dt if dt is not None else 0.0)
    tg = tc = 0.0
    fs = ForcingStream(o, forcing) if forcing else None
    try:
        if substeps_only:
            g.set_dt(dtg, ldt)
            for it in range(1, substeps_only + 1):
                g.rhs_GPU(fg, it)
                dtc, tc, _ = o.substep(fc, dfc, it, ldt, dtc, tc)
            if ldt:
                dtg = g.get_dt()
        else:
            for step in range(nsteps):
                want = bool(diag_every) and step % diag_every == 0
                frc = None
                # forcing coefficients need the dt of the step: Courant dt is known after substep 1 only,
                # so (as the reference does for fixed dt) use the dt entering the step
                if fs:
                    frc = fs.draw(dtg if not ldt else max(dtg, 1e-30))
                dtg, tg = g.time_step(fg, dtg, ldt, tg, forcing=frc, diagnostics=want)
                dtc, tc, dc = o.time_step(fc, dfc, ldt, dtc, tc, forcing=frc, diag=want)
                if want:
                    r = g.get_gpu_reduced_vars()
                    r["dt"] = dtg
                    dc["dt"] = dtc
                    sg.append(r)
                    sc.append(dc)
        g.copy_farray_from_GPU(fg)
    finally:
        g.finalize_GPU()
    o.update_ghosts(fc)
    return fg, fc, sg, sc, (dtg, dtc, tg, tc)
