"""Test helper (NOT product code): the reference's machine-independent random numbers and the two consumers that
samples/helical-MHDturb needs, restated so that the ORACLE can be started from the reference's own initial state and
driven with the reference's own forcing stream.

* MarsRan            -- general.f90:891-942 `mars_ran` (random_gen='nr_f90'): Marsaglia shift sequence x Park-Miller
                        sequence on default (32-bit) integers, `am = nearest(1.0,-1.0)/im` in the build's real kind
                        (REAL_PRECISION=double), seeding by `random_seed_wrapper(PUT=(/s,0/))` (general.f90:799-865:
                        put(2)==0 -> `dummy = mars_ran(put(1))`, i.e. initialise and discard one number);
* gaunoise_vect      -- initcond.f90:5005-5047: per (n, m) row and per variable pair one vector r(mx), one vector p(mx),
                        sin branch for the 1st and 3rd component, cos branch (same r, p) for the 2nd; ghosts included;
* helical_forcing_draw -- forcing.f90:1804-1815 (fran(2): phase, wavevector index) and :1909 (phi), in that order.
"""
import math

import numpy as np

IA, IM, IQ, IR = 16807, 2147483647, 127773, 2836
_M32 = 0xFFFFFFFF


def _i32(v):
    """two's-complement wrap of a Python int to a Fortran default integer"""
    v &= _M32
    return v - (1 << 32) if v & 0x80000000 else v


class MarsRan:
    def __init__(self):
        self.s1 = 0  # rstate(1)
        self.s2 = 0  # rstate(2)
        self.init1 = 1812
        # am = nearest(1.0,-1.0)/real(im) in double precision
        self.am = np.nextafter(1.0, 0.0) / float(IM)

    def seed_put(self, put1, put2=0):
        """random_seed_wrapper(PUT=(/put1,put2/)) for 'nr_f90' (general.f90:838-860)"""
        if put2 == 0:
            self.next(init=put1)
        else:
            self.s1, self.s2 = put1, put2

    def next(self, init=None):
        if init is not None or self.s1 == 0 or self.s2 <= 0:
            if init is not None:
                self.init1 = init
            self.s1 = _i32(777755555 ^ abs(self.init1))
            self.s2 = _i32((888889999 ^ abs(self.init1)) | 1)
        # Marsaglia shift sequence (ishft is a logical shift on the 32-bit pattern)
        u = self.s1 & _M32
        u ^= (u << 13) & _M32
        u ^= u >> 17
        u ^= (u << 5) & _M32
        self.s1 = _i32(u)
        # Park-Miller by Schrage's method
        k = self.s2 // IQ
        self.s2 = IA * (self.s2 - k * IQ) - IR * k
        if self.s2 < 0:
            self.s2 += IM
        return self.am * float(((self.s1 ^ self.s2) & IM) | 1)

    def vector(self, n):
        return np.array([self.next() for _ in range(n)])


def gaunoise_vect(rng, ampl, f, i1, i2):
    """f: (nvar, mz, my, mx) array, variables i1..i2 (0-based, inclusive) get ampl * Gaussian noise."""
    _, mz, my, mx = f.shape
    for n in range(mz):
        for m in range(my):
            r = p = None
            for i in range(i1, i2 + 1):
                if (i - i1) % 2 == 0:
                    r = rng.vector(mx)
                    p = rng.vector(mx)
                    tmp = np.sqrt(-2 * np.log(r)) * np.sin(2 * math.pi * p)
                else:
                    tmp = np.sqrt(-2 * np.log(r)) * np.cos(2 * math.pi * p)
                f[i, n, m, :] += ampl * tmp


def read_kdat(text):
    """k.dat as forcing.f90:892-900 reads it: `nk kav`, then nk values each of kkx, kky, kkz (list-directed)."""
    tok = text.split()
    nk, kav = int(tok[0]), float(tok[1])
    v = [float(t) for t in tok[2:2 + 3 * nk]]
    assert len(v) == 3 * nk
    return nk, kav, list(zip(v[:nk], v[nk:2 * nk], v[2 * nk:]))


def helical_forcing_draw(rng, nk):
    """-> (phase, ik0, phi): ik0 is the 0-based wavevector index."""
    fran1, fran2 = rng.next(), rng.next()
    phase = math.pi * (2 * fran1 - 1.0)
    ik = int(nk * (0.9999 * fran2) + 1)
    phi = rng.next() * 2 * math.pi
    return phase, ik - 1, phi
