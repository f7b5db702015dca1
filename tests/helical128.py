"""samples/turbulence/helical-MHDturb-nopart and helical-MHDturb32-4procs as the reference's CPU runs were configured: this
IS BASELINE.json's configs[1] -- 128^3, [-pi,pi]^3 periodic, isothermal (gamma = 1), u = 0, lnrho = 0, A = 1e-4 x Gaussian
noise, nu = eta = 2e-3, helical forcing (force = 0.05, relhel = 1, the k.dat of samples/helical-MHDturb), Courant time step
with the defaults cdt = 0.9, cdtv = 0.25, 10 steps, it1 = 5, default generator 'min_std'.
Random numbers: every processor draws its local noise from ran0 seeded -(10 + iproc) (start.f90:389); after the initial
condition the seed is set to 1812 on all processors (start.f90:498-501), which equals seed0 and is therefore not written with
the snapshot (persist.f90:260-265); run.x starts from seed0 as well (run.f90:1145-1147): the forcing stream is Ran0(1812).
The forcing is added after the time step with the dt of that step (run.f90:541-575, forcing.f90: fact ~ sqrt(dt))."""
import os

import numpy as np

import nr_f90
from pencil_code_b200 import config as pcfg
from pencil_code_b200 import synthetic

HERE = os.path.dirname(os.path.abspath(__file__))
FORCE, RELHEL = 0.05, 1.0
LAYOUT = {"helical-MHDturb-nopart": (4, 8), "helical-MHDturb32-4procs": (2, 2)}  # (nprocy, nprocz) that drew the noise
FORMATS = {"urms": "%.3E", "brms": "%.3E", "abm": "%.2E", "jbm": "%.2E", "oum": "%.2E", "epsK": "%.2E", "epsM": "%.2E",
           "rhom": "%.6f"}


def golden():
    out, cols = {}, {}
    for l in open(os.path.join(HERE, "golden", "helical128_rows.txt")):
        if l.startswith("#"):
            name, c = l[1:].split(":")
            cols[name.strip()] = c.split()
        else:
            t = l.split()
            out.setdefault(t[0], {})[int(t[1])] = dict(zip(cols[t[0]], t[1:]))
    return out


def setup(name):
    cfg = pcfg.make_config(128, entropy=False, magnetic=True, nu=2e-3, eta=2e-3, gamma=1.0, cdt=0.9, cdtv=0.25)
    f = np.zeros(synthetic.farray_shape(cfg))
    nr_f90.gaunoise_vect_decomposed(f, cfg["iax"], 1e-4, *LAYOUT[name])
    return cfg, f


class ForcingStream:
    def __init__(self, kdat_text):
        self.nk, self.kav, self.kvecs = nr_f90.read_kdat(kdat_text)
        self.rng = nr_f90.Ran0(1812)

    def draw(self):
        phase, ik, phi = nr_f90.helical_forcing_draw(self.rng, self.nk)
        return phase, self.kvecs[ik], phi


def format_reduced(d, eta, row):
    d = dict(d)
    d["epsM"] = eta * d["jrms"] ** 2  # <eta mu0 j^2>, mu0 = 1
    return {k: FORMATS[k] % d[k] for k in FORMATS if k in row}
