"""samples/decay-reconnection as the reference's CPU run was configured (start.in, run.in, src/cparam.local, src/Makefile.local):
32^3, cubic periodic box of size 2 pi / wav1 with wav1 = 0.3 (start.f90:211-216), isothermal (gamma = 1, no energy equation),
u = 0, lnrho = 0, A = 'power_randomphase_hel' (amplaa = 0.2, relhel_aa = 0, initpower_aa = 2, kpeak_aa = 20; the other
parameters at the defaults of magnetic.f90:144-276), nu = 2e-5 ('nu-const'), eta = 4e-6 ('eta-const'), lupw_uu = lupw_lnrho =
lupw_aa = T, Courant time step with the default cdt = 0.9, cdtv = 0.25 (cdata.f90:235-239), 51 steps, it1 = 10.
The run used 4 processors (nprocy = nprocz = 2) and the default generator 'min_std': every processor draws the random phases of
its own block of k-space from seed -(10 + iproc) (start.f90:389)."""
import math
import os

import numpy as np

import nr_f90
from pencil_code_b200 import config as pcfg
from pencil_code_b200 import synthetic

HERE = os.path.dirname(os.path.abspath(__file__))
NSTEPS, IT1 = 51, 10


def golden():
    lines = open(os.path.join(HERE, "golden", "decay_reconnection_rows.txt")).read().splitlines()
    names = lines[0].lstrip("# ").split()[:16]
    return [dict(zip(names, l.split())) for l in lines[1:]]


def setup():
    n = 32
    L = 2 * math.pi / 0.3
    cfg = pcfg.make_config(n, entropy=False, magnetic=True, nu=2e-5, eta=4e-6, upwind=True, Lxyz=L, xyz0=-L / 2,
                           gamma=1.0, cdt=0.9, cdtv=0.25)
    f = np.zeros(synthetic.farray_shape(cfg))
    A = nr_f90.power_randomphase_hel(lambda iproc: nr_f90.Ran0(-(10 + iproc)), n, 2, 2, ampl=2e-1, initpower=2.0,
                                     initpower2=-11. / 3., kpeak=20.0, relhel=0.0)
    f[cfg["iax"]:cfg["iax"] + 3, 3:-3, 3:-3, 3:-3] = A
    return cfg, f


# print.in: t(e10.3) dt(e10.3) urms brms jrms epsK epsM (e10.3), oum abm jbm (e11.3): all d.dddE+xx
def format_reduced(d, eta):
    out = {k: "%.3E" % d[k] for k in ("urms", "brms", "jrms", "epsK", "oum", "abm", "jbm")}
    out["epsM"] = "%.3E" % (eta * d["jrms"] ** 2)  # <eta mu0 j^2>, mu0 = 1 (magnetic.f90: idiag_epsM)
    return out
