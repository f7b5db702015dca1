"""ctypes binding of the CPU ORACLE (oracle/pc_oracle.cpp).

TEST INFRASTRUCTURE ONLY: may be imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs -- never by the product package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


class Pars(C.Structure):
    _fields_ = [
        ("nx", C.c_int), ("ny", C.c_int), ("nz", C.c_int), ("nvar", C.c_int),
        ("iux", C.c_int), ("ilnrho", C.c_int), ("iss", C.c_int), ("iax", C.c_int),
        ("dx", C.c_double), ("dy", C.c_double), ("dz", C.c_double),
        ("x0", C.c_double), ("y0", C.c_double), ("z0", C.c_double),
        ("cs20", C.c_double), ("gamma", C.c_double), ("cp", C.c_double), ("lnrho0", C.c_double),
        ("mu0", C.c_double),
        ("nu", C.c_double), ("nu_hyper3", C.c_double),
        ("lvisc_nu_const", C.c_int), ("lvisc_hyper3_simplified", C.c_int), ("lvisc_hyper3_nu_const", C.c_int),
        ("diffrho_hyper3", C.c_double), ("ldiff_hyper3lnrho", C.c_int),
        ("chi", C.c_double), ("lheatc_chiconst", C.c_int), ("lviscosity_heat", C.c_int),
        ("eta", C.c_double), ("lresi_eta_const", C.c_int), ("lohmic_heat", C.c_int), ("llorentzforce", C.c_int),
        ("lupw_uu", C.c_int), ("lupw_lnrho", C.c_int), ("lupw_ss", C.c_int), ("lupw_aa", C.c_int),
        ("cdt", C.c_double), ("cdtv", C.c_double), ("cdtv2", C.c_double), ("cdtv3", C.c_double),
        ("dtmax", C.c_double), ("nthreads", C.c_int),
        # vertically stratified set-ups (conv-slab): z boundary conditions, gravity, rotation, K-profile, cooling
        ("bcz", (C.c_int * 8) * 2),
        ("gravz", C.c_double), ("Omega", C.c_double),
        ("lheatc_Kprof", C.c_int),
        ("hcond0", C.c_double), ("hcond1", C.c_double), ("hcond2", C.c_double),
        ("widthss", C.c_double), ("z1", C.c_double), ("z2", C.c_double),
        ("cool", C.c_double), ("wcool", C.c_double), ("cs2cool", C.c_double), ("ztop", C.c_double),
        ("FbotKbot", C.c_double), ("FtopKtop", C.c_double), ("cs2bot", C.c_double), ("cs2top", C.c_double),
    ]


BC_CODES = {"p": 0, "s": 1, "a": 2, "a2": 3, "c1": 4, "cT": 5}


class Forcing(C.Structure):
    _fields_ = [("coef1", C.c_double * 3), ("coef2", C.c_double * 3), ("fda", C.c_double * 3),
                ("kk", C.c_double * 3), ("phase", C.c_double), ("fact", C.c_double), ("k1_ff", C.c_double)]


class Diag(C.Structure):
    _fields_ = [(n, C.c_double) for n in
                ("urms", "umax", "rhom", "brms", "bmax", "jrms", "jmax", "abm", "jbm", "oum", "epsK", "ssm",
                 "dtu", "dtc", "dtb", "dtnu", "dt1_max", "dtchi")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def build(fast=False, outdir=None):
    """Compile the oracle with g++ (seconds). fast=True -> -O3 -march=native timing build."""
    outdir = outdir or _HERE
    name = "libpc_oracle_fast.so" if fast else "libpc_oracle.so"
    out = os.path.join(outdir, name)
    src = os.path.join(_HERE, "pc_oracle.cpp")
    if os.path.exists(out) and os.path.getmtime(out) >= max(os.path.getmtime(src), os.path.getmtime(os.path.join(_HERE, "pc_oracle.h"))) and not fast:
        return out
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    flags = (["-O3", "-march=native"] if fast else ["-O2", "-ffp-contract=off", "-fno-fast-math"])
    cmd = [cxx] + flags + ["-fopenmp", "-fPIC", "-std=c++17", "-shared", "-o", out, src]
    subprocess.check_call(cmd)
    return out


_libs = {}


def lib(fast=False, outdir=None):
    key = (fast, outdir)
    if key in _libs:
        return _libs[key]
    path = build(fast=fast, outdir=outdir)
    L = C.CDLL(path)
    dp = C.POINTER(C.c_double)
    L.pc_oracle_update_ghosts.argtypes = [C.POINTER(Pars), dp]
    L.pc_oracle_deriv.argtypes = [C.POINTER(Pars), dp, C.c_int, C.c_int, C.c_int, dp]
    L.pc_oracle_rhs.argtypes = [C.POINTER(Pars), dp, dp, C.c_int, C.POINTER(Diag)]
    L.pc_oracle_rhs.restype = C.c_double
    L.pc_oracle_substep.argtypes = [C.POINTER(Pars), dp, dp, C.c_int, C.c_int, dp, dp, C.POINTER(Diag)]
    L.pc_oracle_time_step.argtypes = [C.POINTER(Pars), dp, dp, C.c_int, dp, dp, C.POINTER(Forcing), C.POINTER(Diag)]
    L.pc_oracle_forcing_hel.argtypes = [C.POINTER(Pars), dp, C.POINTER(Forcing)]
    L.pc_oracle_forcing_coefs.argtypes = [C.c_double] * 5 + [dp, C.c_double, C.c_double, C.POINTER(Forcing)]
    L.pc_oracle_sizeof_pars.restype = C.c_int
    assert L.pc_oracle_sizeof_pars() == C.sizeof(Pars), "oracle struct layout mismatch"
    _libs[key] = L
    return L


def _dp(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_double))


def pars_from_dict(d, nthreads=0):
    p = Pars()
    for k, v in d.items():
        if k == "bcz":  # list of 'bot:top' strings per variable, as in run.in (boundcond.f90:175-260)
            for j, spec in enumerate(v):
                bot, _, top = spec.partition(":")
                p.bcz[0][j] = BC_CODES[bot]
                p.bcz[1][j] = BC_CODES[top or bot]
        elif hasattr(p, k):
            setattr(p, k, v)
    p.nthreads = nthreads
    return p


class Oracle:
    """Convenience wrapper. f arrays are numpy (nvar, mz, my, mx) float64 (x fastest)."""

    def __init__(self, pars_dict, nthreads=0, fast=False, outdir=None):
        self.L = lib(fast=fast, outdir=outdir)
        self.p = pars_from_dict(pars_dict, nthreads)

    def shape(self):
        p = self.p
        return (p.nvar, p.nz + 6, p.ny + 6, p.nx + 6)

    def update_ghosts(self, f):
        self.L.pc_oracle_update_ghosts(C.byref(self.p), _dp(f))

    def deriv(self, f, k, j, kind):
        p = self.p
        out = np.empty((p.nz, p.ny, p.nx))
        self.L.pc_oracle_deriv(C.byref(self.p), _dp(f), k, j, kind, _dp(out))
        return out

    def rhs(self, f, df, lcourant=False, diag=False):
        d = Diag() if diag else None
        dt1 = self.L.pc_oracle_rhs(C.byref(self.p), _dp(f), _dp(df), int(lcourant), C.byref(d) if diag else None)
        return dt1, (d.as_dict() if diag else None)

    def substep(self, f, df, itsub, ldt, dt, t, diag=False):
        d = Diag() if diag else None
        cdt, ct = C.c_double(dt), C.c_double(t)
        self.L.pc_oracle_substep(C.byref(self.p), _dp(f), _dp(df), itsub, int(ldt), C.byref(cdt), C.byref(ct),
                                 C.byref(d) if diag else None)
        return cdt.value, ct.value, (d.as_dict() if diag else None)

    def time_step(self, f, df, ldt, dt, t, forcing=None, diag=False):
        d = Diag() if diag else None
        cdt, ct = C.c_double(dt), C.c_double(t)
        self.L.pc_oracle_time_step(C.byref(self.p), _dp(f), _dp(df), int(ldt), C.byref(cdt), C.byref(ct),
                                   C.byref(forcing) if forcing is not None else None, C.byref(d) if diag else None)
        return cdt.value, ct.value, (d.as_dict() if diag else None)

    def forcing_hel(self, f, frc):
        self.L.pc_oracle_forcing_hel(C.byref(self.p), _dp(f), C.byref(frc))

    def forcing_coefs(self, force, relhel, kav, cs0eff, dt, kvec, phase, phi):
        out = Forcing()
        kv = (C.c_double * 3)(*kvec)
        self.L.pc_oracle_forcing_coefs(force, relhel, kav, cs0eff, dt, kv, phase, phi, C.byref(out))
        return out
