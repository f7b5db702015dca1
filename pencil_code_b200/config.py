"""Run configurations of the hot path (what src/cparam.local + src/Makefile.local + run.in
select in the reference), as plain dicts shared by the GPU module binding and the tests.

Keys follow the reference's names (cparam.f90, *_run_pars namelists).  Variable slots are
0-based (Fortran index - 1) in registration order ux,uy,uz,lnrho,[ss],[ax,ay,az]
(register.f90:135-147).
"""
import math

DEFAULTS = dict(
    nprocx=1, nprocy=1, nprocz=1, ipx=0, ipy=0, ipz=0,
    cs20=1.0, gamma=5.0 / 3.0, cp=1.0, lnrho0=0.0, mu0=1.0,           # eos_idealgas.f90:46-47, cdata.f90:64
    nu=0.0, nu_hyper3=0.0, lvisc_nu_const=0, lvisc_hyper3_simplified=0, lvisc_hyper3_nu_const=0,
    diffrho_hyper3=0.0, ldiff_hyper3lnrho=0,
    chi=0.0, lheatc_chiconst=0, lviscosity_heat=1,                    # entropy.f90:130
    eta=0.0, lresi_eta_const=0, lohmic_heat=1, llorentzforce=1,       # magnetic.f90:205,247
    lupw_uu=0, lupw_lnrho=0, lupw_ss=0, lupw_aa=0,
    cdt=0.9, cdtv=0.25, cdtv2=0.03, cdtv3=0.01, dtmax=1e37,           # cdata.f90:235-239
    device=-1, kernel_variant=0,
)


def make_config(n, entropy, magnetic, nu=0.0, eta=0.0, chi=0.0, hyper3=False, upwind=False,
                ny=None, nz=None, nprocz=1, ipz=0, nprocy=1, ipy=0, Lxyz=2 * math.pi, xyz0=-math.pi, **over):
    """Periodic cube of n^3 global cells (or n x ny x nz), decomposed into nprocy x nprocz yz-pencils
    (nprocy=1: z-slabs); rank = ipy + nprocy*ipz (mpicomm.f90, nprocx=1).
    Grid: dx=L/nxgrid, x_i=x0+(i+1/2)dx (grid.f90:123,149); defaults xyz0=-pi, Lxyz=2pi (cdata.f90:209-211)."""
    nxg = n
    nyg = ny if ny is not None else n
    nzg = nz if nz is not None else n
    assert nzg % nprocz == 0 and nyg % nprocy == 0
    c = dict(DEFAULTS)
    dx, dy, dz = Lxyz / nxg, Lxyz / nyg, Lxyz / nzg
    nzl = nzg // nprocz
    nyl = nyg // nprocy
    nvar = 4 + (1 if entropy else 0) + (3 if magnetic else 0)
    c.update(
        nx=nxg, ny=nyl, nz=nzl, nprocz=nprocz, ipz=ipz, nprocy=nprocy, ipy=ipy, nvar=nvar,
        iux=0, ilnrho=3, iss=4 if entropy else -1, iax=(5 if entropy else 4) if magnetic else -1,
        dx=dx, dy=dy, dz=dz,
        x0=xyz0 + 0.5 * dx, y0=xyz0 + 0.5 * dy + ipy * nyl * dy, z0=xyz0 + 0.5 * dz + ipz * nzl * dz,
        nu=nu, lvisc_nu_const=1 if nu > 0 else 0,
        eta=eta if magnetic else 0.0, lresi_eta_const=1 if (magnetic and eta > 0) else 0,
        chi=chi if entropy else 0.0, lheatc_chiconst=1 if (entropy and chi > 0) else 0,
    )
    if hyper3:
        c.update(lvisc_hyper3_nu_const=1, nu_hyper3=5e-12, ldiff_hyper3lnrho=1, diffrho_hyper3=5e-12)
    if upwind:
        c.update(lupw_uu=1, lupw_lnrho=1, lupw_ss=1 if entropy else 0, lupw_aa=1 if magnetic else 0)
    c.update(over)
    return c


# BASELINE.json configs 2-5
def config_helical_mhd_128(**kw):
    """config[1]: isothermal forced MHD turbulence 128^3 (lnrho, uu, aa), nu=eta=5e-3."""
    return make_config(128, entropy=False, magnetic=True, nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3, **kw)


def config_mhd_entropy(n=512, **kw):
    """config[2]: periodic MHD with entropy, 8 variables, nu=eta=chi=2e-4 (512^3)."""
    return make_config(n, entropy=True, magnetic=True, nu=2e-4, eta=2e-4, chi=2e-4, cdt=0.4, cdtv=0.3, **kw)


def config_hydro_hyper3(n=256, **kw):
    """config[4]: hydro-only forced turbulence with 6th-order hyperdiffusion (del6)."""
    return make_config(n, entropy=False, magnetic=False, nu=0.0, hyper3=True, cdt=0.4, cdtv=0.3, **kw)


def fixed_dt(cfg):
    """dt = 0.4*dx/1.5 for the fixed-dt timing mode."""
    return 0.4 * cfg["dx"] / 1.5


def kvectors(kmin=4.5, kmax=5.5):
    """All integer k with kmin <= |k| < kmax (stands in for k.dat of the helical forcing)."""
    out = []
    r = int(math.ceil(kmax))
    for kx in range(-r, r + 1):
        for ky in range(-r, r + 1):
            for kz in range(-r, r + 1):
                k = math.sqrt(kx * kx + ky * ky + kz * kz)
                if kmin <= k < kmax:
                    out.append((float(kx), float(ky), float(kz)))
    return out
