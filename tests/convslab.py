"""samples/conv-slab of the reference as a configuration + initial condition for the oracle and the CUDA path.

Restates, for tests only, what start.x does for this sample (start.in) and what run.in selects:
  * grid: 32^3, xyz0=(-0.5,-0.5,-0.68), Lxyz=(1,1,2), periodic in x and y (x_i = x0+(i+1/2)dx), closed in z
    (z_n = z0+n dz, dz = Lz/(nz-1)): grid.f90:119-149,208-219
  * hydro 'up-down' (hydro.f90:2633-2647), density and entropy 'piecew-poly' (density.f90:1578-1610,2111-2177;
    entropy.f90:1971-1992,2186-2260) with gravz=-1, z1=0, z2=1, zref=1.32, mpoly0/1/2 = 1/3/0, isothtop=1
  * run.in: bcz = s, s, a, a2, c1:cT; Omega=0.1; nu=4e-3; K-profile conduction hcond0=8e-3; cooling cool=15,
    wcool=0.2, cs2cool=1/3; upwinding of lnrho and ss; cdt=0.4, cdtv=0.3, Courant time step
The 14-digit time series this reproduces is samples/conv-slab/reference.out.double (tests/golden/convslab_rows.txt).
"""
import math

import numpy as np

NG = 3
COLUMNS = ("it", "t", "dt", "urms", "umax", "rhom", "ssm", "dtc", "dtu", "dtnu", "dtchi")


def config(n=32, cp=1.0, **over):
    """cp = 2 is samples/conv-slab_cp_2 (cp doubled, hcond0 doubled: the reference holds the SAME series for it, with ssm/cp
    in place of ssm -- its check that nothing assumes cp = 1)."""
    cs0, gamma, rho0 = .57735, 1.6666666, 1.0
    Lx = Ly = 1.0
    Lz = 2.0
    x0 = y0 = -0.5
    z0 = -0.68
    dx, dy, dz = Lx / n, Ly / n, Lz / (n - 1)
    gravz, z1, z2, zref = -1.0, 0.0, 1.0, 1.32
    mpoly0, mpoly1, mpoly2 = 1.0, 3.0, 0.0
    hcond0 = 8e-3 * cp
    gamma_m1 = gamma - 1.0
    hcond1 = (mpoly1 + 1.) / (mpoly0 + 1.)  # entropy.f90:849-850
    hcond2 = (mpoly2 + 1.) / (mpoly0 + 1.)
    Fbot = -gamma / (cp * gamma_m1) * hcond0 * gravz / (mpoly0 + 1)  # entropy.f90:854-859 (bcz(iss) bottom = 'c1')
    FbotKbot = Fbot / (hcond0 * hcond1)  # :864-866
    cool, cs2cool = 15.0, 0.333333
    c = dict(
        nx=n, ny=n, nz=n, nprocx=1, nprocy=1, nprocz=1, ipx=0, ipy=0, ipz=0, nvar=5,
        iux=0, ilnrho=3, iss=4, iax=-1,
        dx=dx, dy=dy, dz=dz, x0=x0 + 0.5 * dx, y0=y0 + 0.5 * dy, z0=z0,
        cs20=cs0 ** 2, gamma=gamma, cp=cp, lnrho0=math.log(rho0), mu0=1.0,
        nu=4e-3, nu_hyper3=0.0, lvisc_nu_const=1, lvisc_hyper3_simplified=0, lvisc_hyper3_nu_const=0,
        diffrho_hyper3=0.0, ldiff_hyper3lnrho=0, chi=0.0, lheatc_chiconst=0, lviscosity_heat=1,
        eta=0.0, lresi_eta_const=0, lohmic_heat=1, llorentzforce=1,
        lupw_uu=0, lupw_lnrho=1, lupw_ss=1, lupw_aa=0,
        cdt=0.4, cdtv=0.30, cdtv2=0.03, cdtv3=0.01, dtmax=1e37,
        device=-1, kernel_variant=0,
        bcz=["s", "s", "a", "a2", "c1:cT"],
        gravz=gravz, Omega=0.1,
        lheatc_Kprof=1, hcond0=hcond0, hcond1=hcond1, hcond2=hcond2, widthss=0.05, z1=z1, z2=z2,
        cool=cool, wcool=0.2, cs2cool=cs2cool, ztop=z0 + Lz,
        FbotKbot=FbotKbot, FtopKtop=0.0,
        # run: cs2top is replaced by (cool*cs2cool+cool2*cs2cool2)/(cool+cool2) = cs2cool (entropy.f90:818-838);
        # cs2bot (density.f90:1605) is not used by these boundary conditions
        cs2bot=1.0, cs2top=cs2cool,
    )
    c["_init"] = dict(cs0=cs0, gamma=gamma, cp=cp, gravz=gravz, z1=z1, z2=z2, zref=zref, mpoly=(mpoly0, mpoly1, mpoly2),
                      xyz0=(x0, y0, z0), Lxyz=(Lx, Ly, Lz), widthlnrho=0.05, widthss=0.05, ampluu=0.1, widthuu=0.1,
                      isothtop=1, fac_cs=1.0, lnrho0=math.log(rho0))
    c.update(over)
    return c


def _step(z, z0, width):  # sub.f90:4975-4987
    return 0.5 * (1 + np.tanh((z - z0) / width))


def initial_condition(c):
    """f[nvar, mz, my, mx] as start.x leaves it (ghost zones zero: run.x fills them before the first use)."""
    I = c["_init"]
    n = c["nx"]
    mx = n + 2 * NG
    f = np.zeros((c["nvar"], mx, mx, mx))
    x = c["x0"] + c["dx"] * np.arange(n)
    y = c["y0"] + c["dy"] * np.arange(n)
    z = c["z0"] + c["dz"] * np.arange(n)
    x0, y0, z0 = I["xyz0"]
    Lx, Ly, Lz = I["Lxyz"]
    gamma, gravz, cp1 = I["gamma"], I["gravz"], 1.0 / I["cp"]
    gamma_m1 = gamma - 1.0
    z1, z2, zref = I["z1"], I["z2"], I["zref"]
    mpoly0, mpoly1, mpoly2 = I["mpoly"]
    ztop = z0 + Lz

    # --- hydro.f90:2633-2647 'up-down'
    uz = np.zeros((n, n, n))
    for kz in range(n):
        prof = I["ampluu"] * math.exp(-0.5 * (z[kz] - z1) ** 2 / I["widthuu"] ** 2)
        for ky in range(n):
            tmp = np.sqrt((x - (x0 + 0.3 * Lx)) ** 2 + (y[ky] - (y0 + 0.3 * Ly)) ** 2)
            row = prof * np.exp(-0.5 * (tmp ** 2) / I["widthuu"] ** 2)
            tmp = np.sqrt((x - (x0 + 0.5 * Lx)) ** 2 + (y[ky] - (y0 + 0.8 * Ly)) ** 2)
            uz[kz, ky, :] = row - 0.7 * prof * np.exp(-0.5 * (tmp ** 2) / I["widthuu"] ** 2)
    f[c["iux"] + 2, NG:-NG, NG:-NG, NG:-NG] = uz

    # --- density.f90:2111-2177 polytropic_lnrho_z, called top -> bottom (density.f90:1596-1602)
    lnrho = np.full(n, I["lnrho0"])
    st = dict(cs2int=I["cs0"] ** 2, lnrhoint=I["lnrho0"])

    def poly_lnrho(mpoly, zint, zbot, zblend, isoth, fac_cs=None):
        nonlocal lnrho
        stp = _step(z, zblend, I["widthlnrho"])
        if isoth != 0:
            beta1 = 0.0
            tmp = st["lnrhoint"] + gamma * gravz / st["cs2int"] * (z - zint)
        else:
            beta1 = gamma * gravz / (mpoly + 1)
            tmp = 1.0 + beta1 * (z - zint) / st["cs2int"]
            tmp = np.maximum(tmp, 5 * np.finfo(float).eps)
            tmp = st["lnrhoint"] + mpoly * np.log(tmp)
        lnrho = stp * lnrho + (1 - stp) * tmp
        if isoth != 0:
            st["lnrhoint"] = st["lnrhoint"] + gamma * gravz / st["cs2int"] * (zbot - zint)
        else:
            st["lnrhoint"] = st["lnrhoint"] + mpoly * math.log(1 + beta1 * (zbot - zint) / st["cs2int"])
        if isoth != 0 and fac_cs is not None:
            st["cs2int"] = fac_cs ** 2 * st["cs2int"]
        else:
            st["cs2int"] = st["cs2int"] + beta1 * (zbot - zint)

    poly_lnrho(mpoly2, zref, z2, ztop + Lz, I["isothtop"], I["fac_cs"])
    poly_lnrho(mpoly0, z2, z1, z2, 0)
    poly_lnrho(mpoly1, z1, z0, z1, 0)
    f[c["ilnrho"], NG:-NG, NG:-NG, NG:-NG] = lnrho[:, None, None]

    # --- entropy.f90:2186-2260 polytropic_ss_z (entropy.f90:1971-1992)
    ss = np.zeros(n)
    se = dict(cs2int=I["cs0"] ** 2, ssint=0.0)

    def poly_ss(mpoly, zint, zbot, zblend, isoth, fac_cs=None):
        nonlocal ss
        stp = _step(z, zblend, I["widthss"])
        if isoth != 0:
            beta1 = 0.0
            tmp = se["ssint"] - gamma_m1 * gravz * (z - zint) / se["cs2int"] / cp1
        else:
            beta1 = gamma * gravz / (mpoly + 1)
            tmp = 1.0 + beta1 * (z - zint) / se["cs2int"]
            tmp = np.maximum(tmp, 5 * np.finfo(float).eps)
            tmp = se["ssint"] + (1 - mpoly * gamma_m1) / gamma * np.log(tmp) / cp1
        ss = stp * ss + (1 - stp) * tmp
        if isoth != 0:
            if fac_cs is not None:
                se["ssint"] = se["ssint"] - gamma_m1 * gravz * (zbot - zint) / se["cs2int"] / cp1 + 2 * math.log(fac_cs) / gamma / cp1
            else:
                se["ssint"] = se["ssint"] - gamma_m1 * gravz * (zbot - zint) / se["cs2int"] / cp1
        else:
            se["ssint"] = se["ssint"] + (1 - mpoly * gamma_m1) / gamma * math.log(1 + beta1 * (zbot - zint) / se["cs2int"]) / cp1
        if isoth != 0 and fac_cs is not None:
            se["cs2int"] = fac_cs ** 2 * se["cs2int"]
        else:
            se["cs2int"] = se["cs2int"] + beta1 * (zbot - zint)

    poly_ss(mpoly2, zref, z2, z0 + 2 * Lz, I["isothtop"], I["fac_cs"])
    poly_ss(mpoly0, z2, z1, z2, 0)
    poly_ss(mpoly1, z1, z0, z1, 0)
    f[c["iss"], NG:-NG, NG:-NG, NG:-NG] = ss[:, None, None]
    return f


def golden_rows(path):
    rows = []
    for line in open(path):
        if line.startswith("#") or not line.strip():
            continue
        v = line.split()
        rows.append(dict(zip(COLUMNS, [int(v[0])] + [float(s) for s in v[1:]])))
    return rows


def series_row(it, t, dt, d):
    """one row of the time series as print.in.double lists it; the dt* columns are l_dt=.true. diagnostics
    (diagnostics.f90: multiplied by dt when printed)"""
    return dict(it=it, t=t, dt=dt, urms=d["urms"], umax=d["umax"], rhom=d["rhom"], ssm=d["ssm"],
                dtc=d["dtc"] * dt, dtu=d["dtu"] * dt, dtnu=d["dtnu"] * dt, dtchi=d["dtchi"] * dt)
