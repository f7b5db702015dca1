"""Test helper (NOT product code): the reference's machine-independent random numbers and the two consumers that
samples/helical-MHDturb needs, restated so that the ORACLE can be started from the reference's own initial state and
driven with the reference's own forcing stream.

* MarsRan            -- general.f90:891-942 `mars_ran` (random_gen='nr_f90'): Marsaglia shift sequence x Park-Miller
                        sequence on default (32-bit) integers, `am = nearest(1.0,-1.0)/im` in the build's real kind
                        (REAL_PRECISION=double), seeding by `random_seed_wrapper(PUT=(/s,0/))` (general.f90:799-865:
                        put(2)==0 -> `dummy = mars_ran(put(1))`, i.e. initialise and discard one number);
* gaunoise_vect      -- initcond.f90:5005-5047: per (n, m) row and per variable pair one vector r(mx), one vector p(mx),
                        sin branch for the 1st and 3rd component, cos branch (same r, p) for the 2nd; ghosts included;
* helical_forcing_draw -- forcing.f90:1804-1815 (fran(2): phase, wavevector index) and :1909 (phi), in that order;
* Ran0               -- general.f90:867-889 `ran0` (random_gen='min_std', the default: general.f90:86), state = the single
                        seed integer set by `random_seed_wrapper(PUT=...)` (general.f90:817-821), negative seeds included
                        (Fortran integer division truncates toward zero);
* power_randomphase_hel -- initcond.f90:5772-6653 for the arguments magnetic.f90:2326-2331 passes by default (3-D vector
                        potential, no cutoff, no Gaussian, lscale_tobox=T, projection on, nfact=4, lfactors=F, no random
                        amplitudes): per-processor random phases in the processor's block of k-space (draw order of
                        random_number_wrapper_3, general.f90:722-768: first index outermost), broken power law, solenoidal
                        projection, unnormalised backward FFT (fourier_fftpack.f90:3561-3710: cfftb, natural
                        index order) -- here one numpy FFT of the global array;
* ran0_stream, gaunoise_vect_decomposed -- the same generator vectorised, and gaunoise_vect as a run on several processors
                        draws it (every processor its local array from its own seed).
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


MASK = 123459876


def _tdiv(a, b):
    q = abs(a) // abs(b)
    return q if (a >= 0) == (b >= 0) else -q


class Ran0:
    def __init__(self, seed):
        self.s = seed  # rstate(1)
        self.am = 1.0 / float(IM)

    def next(self):
        d = _i32(self.s ^ MASK)
        k = _tdiv(d, IQ)
        d = _i32(IA * (d - k * IQ) - IR * k)
        if d < 0:
            d = _i32(d + IM)
        r = self.am * d
        self.s = _i32(d ^ MASK)
        return r


def power_randomphase_hel(make_rng, n, nprocy, nprocz, ampl, initpower, initpower2, kpeak, relhel, nfact=4.0):
    """-> A[3][z][y][x] on the global n^3 grid.  make_rng(iproc) returns the generator of processor iproc as start.x seeds it
    (start.f90:389: seed = -(10 + iproc)); iproc = ipy + nprocy*ipz (nprocx = 1)."""
    nx, ny, nz = n, n // nprocy, n // nprocz
    U = np.zeros((3, n, n, n), dtype=complex)  # [component][kz][ky][kx]
    for iproc in range(nprocy * nprocz):
        ipy, ipz = iproc % nprocy, iproc // nprocy
        rng = make_rng(iproc)
        for c in range(3):
            r = np.array([rng.next() for _ in range(nx * ny * nz)]).reshape(nx, ny, nz)  # do i; do j; do k: a(i,j,k)
            ph = math.pi * (2 * r - 1)
            blk = ampl * (np.cos(ph) + 1j * np.sin(ph))
            U[c, ipz * nz:(ipz + 1) * nz, ipy * ny:(ipy + 1) * ny, :] = blk.transpose(2, 1, 0)
    kk = np.roll(np.arange(n) - n // 2, n // 2).astype(float)  # cshift((/(i-n/2,i=0,n-1)/), n/2)
    KZ, KY, KX = np.meshgrid(kk, kk, kk, indexing="ij")
    k2 = KX ** 2 + KY ** 2 + KZ ** 2
    k2[0, 0, 0] = 1.0  # "avoid division by zero"
    mhalf = .25 * (initpower - 3 + 1)
    nexp1, nexp2 = .25 * nfact * (initpower - initpower2), 1. / nfact
    kpeak1, kpeak21 = 1. / kpeak, 1. / kpeak ** 2
    fact = (kpeak1 * 1.0) ** 1.5  # scale_factor = 1 (lscale_tobox=T), D1 = 0, D2 = 1
    fact *= kpeak1                # lvectorpotential
    U *= fact * ((k2 * kpeak21) ** mhalf) / (1. + (k2 * kpeak21) ** nexp1) ** nexp2
    kdotu = (KX * U[0] + KY * U[1] + KZ * U[2]) / k2  # projection: v_i = u_i - k_i k_j u_j / k^2
    U[0] -= KX * kdotu
    U[1] -= KY * kdotu
    U[2] -= KZ * kdotu
    assert relhel == 0.0, "helical part (initcond.f90:6459-6530) not restated: samples/decay-reconnection has relhel_aa=0"
    return np.fft.ifftn(U, axes=(1, 2, 3)).real * float(n) ** 3


def ran0_stream(seed, count):
    """The first `count` numbers of Ran0(seed) as an array.  With the mask applied on the way in and out, ran0 is the plain
    Lehmer sequence d_{n+1} = 16807 d_n mod (2^31 - 1) (Schrage's method is exact for positive d), so after the first number
    (scalar routine: the seed may be negative) the sequence is built by doubling: d[k:2k] = 16807^k d[0:k]."""
    r0 = Ran0(seed)
    r0.next()
    d = np.empty(count, dtype=np.uint64)
    d[0] = r0.s ^ MASK
    k, g = 1, IA
    while k < count:
        m = min(k, count - k)
        d[k:k + m] = (d[:m] * np.uint64(g)) % np.uint64(IM)
        g = (g * g) % IM
        k *= 2
    return d.astype(np.float64) * (1.0 / float(IM))


def gaunoise_vect_decomposed(f, i1, ampl, nprocy, nprocz):
    """gaunoise_vect (initcond.f90:5005-5047) for the three variables i1..i1+2 of a run on nprocy x nprocz processors with the
    default generator: processor iproc = ipy + nprocy*ipz draws its LOCAL array (ghost rows and planes included, they are
    overwritten by the boundary exchange) from Ran0(-(10 + iproc)) (start.f90:389).  f is the global array (nvar, mz, my, mx)."""
    _, mzg, myg, mx = f.shape
    ny, nz = (myg - 6) // nprocy, (mzg - 6) // nprocz
    my, mz = ny + 6, nz + 6
    for iproc in range(nprocy * nprocz):
        ipy, ipz = iproc % nprocy, iproc // nprocy
        u = ran0_stream(-(10 + iproc), mz * my * 4 * mx).reshape(mz, my, 4, mx)  # per (n, m): r, p for (1st, 2nd), r, p for 3rd
        a1, a2 = np.sqrt(-2 * np.log(u[:, :, 0])), np.sqrt(-2 * np.log(u[:, :, 2]))
        loc = np.stack([a1 * np.sin(2 * math.pi * u[:, :, 1]), a1 * np.cos(2 * math.pi * u[:, :, 1]),
                        a2 * np.sin(2 * math.pi * u[:, :, 3])])
        f[i1:i1 + 3, 3 + ipz * nz:3 + (ipz + 1) * nz, 3 + ipy * ny:3 + (ipy + 1) * ny, :] += ampl * loc[:, 3:-3, 3:-3, :]
