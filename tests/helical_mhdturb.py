"""samples/helical-MHDturb as the reference's CPU run was configured, restated for the tests (start.in, run.in,
src/cparam.local, src/Makefile.local of that sample).  What moves (u, lnrho, A) there is exactly config[1]'s physics:
32^3, [-pi,pi]^3 periodic, isothermal (ENTROPY=noentropy, gamma=1), nu = eta = 5e-3 ('nu-const', 'eta-const'), helical
forcing (force=0.07, relhel=1, k.dat of the sample), fixed dt = 4e-2.  The sample's other modules do not act back on
these variables: the passive scalar and the tracer/stalker particles are passive, `lcoulomb=T` only fills the auxiliary
slots Lam and diva (magnetic.f90:4095-4103, :4331-4334).

Random numbers (tests/nr_f90.py): start.x seeds channel 1 with PUT=(/-10,0/) (start.f90:389-390, iproc=0), draws the
Gaussian-noise A (magnetic.f90:2336 -> initcond.f90:5005-5047) and, after the particles, re-seeds with (/1812,0/)
(start.f90:498-501, lseed_global=T); that state is saved with var.dat (persist.f90:252-265) and read back by run.x
(run.f90:1168), so the forcing stream of the run is mars_ran seeded with 1812, one number discarded.
"""
import json
import math
import os

import numpy as np

import nr_f90
from pencil_code_b200 import config as pcfg
from pencil_code_b200 import synthetic

HERE = os.path.dirname(os.path.abspath(__file__))
DT = 4e-2
FORCE, RELHEL = 0.07, 1.0


def golden():
    g = json.load(open(os.path.join(HERE, "golden", "helical_mhdturb.json")))
    g["rows"] = [dict(zip(g["columns"], r.split())) for r in g["rows"]]
    return g


def setup():
    """-> (cfg, f) with the reference's initial state: u = 0, lnrho = 0, A = 1e-4 x Gaussian noise."""
    cfg = pcfg.make_config(32, entropy=False, magnetic=True, nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3, gamma=1.0)
    f = np.zeros(synthetic.farray_shape(cfg))
    rng = nr_f90.MarsRan()
    rng.seed_put(-10, 0)
    nr_f90.gaunoise_vect(rng, 1e-4, f, cfg["iax"], cfg["iax"] + 2)
    return cfg, f


class ForcingStream:
    """(phase, wavevector, phi) per step from the run's own generator state."""

    def __init__(self, kdat_text):
        self.nk, self.kav, self.kvecs = nr_f90.read_kdat(kdat_text)
        self.rng = nr_f90.MarsRan()
        self.rng.seed_put(1812, 0)

    def draw(self):
        phase, ik, phi = nr_f90.helical_forcing_draw(self.rng, self.nk)
        return phase, self.kvecs[ik], phi


# print.in of the sample: umax(e10.3) urms(e10.3) brms(e10.3) jrms(e10.3) jmax(e10.2) abm(e10.2) jbm(e10.2) oum(e10.2);
# legend.f90 adds the 1p scale factor, so eN.d prints d.ddd as C's %.dE does
REDUCED_FORMATS = (("umax", "%.3E"), ("urms", "%.3E"), ("brms", "%.3E"), ("jrms", "%.3E"), ("jmax", "%.2E"),
                   ("abm", "%.2E"), ("jbm", "%.2E"), ("oum", "%.2E"))


def format_reduced(d):
    return {k: fmt % d[k] for k, fmt in REDUCED_FORMATS}
