"""Analytic pin of the oracle's right-hand side (CPU suite): fields that are a constant plus plane waves have CLOSED-FORM
discrete derivatives under the reference's stencils -- for f = cos(phi), phi = k.x + p, on an equidistant periodic grid

   der   (deriv.f90:89-171)    d_i f  = -K1_i sin(phi),  K1_i = (45 sin t - 9 sin 2t + sin 3t)/(30 d),  t = k_i d
   der2  (deriv.f90:472-561)   d_ii f = -K2_i cos(phi),  K2_i = (490 - 2(270 cos t - 27 cos 2t + 2 cos 3t))/(180 d^2)
   derij (deriv.f90:1400-1452, bidiagonal)  d_ij f = -Kx_ij cos(phi),
         Kx_ij = 4 (270 sin t_i sin t_j - 27 sin 2t_i sin 2t_j + 2 sin 3t_i sin 3t_j)/(720 d_i d_j)

-- so every term of the MHD equations the path evaluates (equ.f90:1386-1567 and the modules it calls) can be written
down pointwise without any stencil code and compared with pc_oracle_rhs to round-off.  This is independent of the
oracle's derivative and assembly code (it shares only the reading of the equations): induction u x B + eta del2 A,
Lorentz force with j = (grad div A - del2 A)/mu0 through the bidiagonal mixed derivatives, viscous force
nu (del2 u + grad div u / 3 + 2 S.grad lnrho), pressure gradient with and without entropy, viscous and Ohmic heating,
chi-const heat conduction, advection and continuity -- and, with

   der5  (deriv.f90:855-935)   d5_i f = -K5_i sin(phi),  K5_i = (5 sin t - 4 sin 2t + sin 3t)/d^5
   der6  (deriv.f90:936-1072)  d6_i f =  K6_i cos(phi),  K6_i = (-20 + 30 cos t - 12 cos 2t + 2 cos 3t)/d^6
         (upwind form: the same bracket times 1/(60 d))

the sixth-order hyperdiffusion terms (viscosity.f90: hyper3-simplified, hyper3_nu-const; density.f90: hyper3lnrho) and the
6th-order upwinding of u, lnrho, s and A (sub.f90:7541-7574 doupwind; magnetic.f90:6257-6313)."""
import numpy as np
import pytest

from pencil_code_b200 import config as pcfg


class Waves:
    """f(x) = c0 + sum_m a_m cos(k_m.x + p_m) on the interior grid of cfg; analytic discrete derivatives."""

    def __init__(self, cfg, c0, modes):
        self.cfg, self.c0, self.modes = cfg, c0, modes
        n = (cfg["nz"], cfg["ny"], cfg["nx"])
        z = cfg["z0"] + cfg["dz"] * np.arange(n[0])
        y = cfg["y0"] + cfg["dy"] * np.arange(n[1])
        x = cfg["x0"] + cfg["dx"] * np.arange(n[2])
        self.X = np.meshgrid(z, y, x, indexing="ij")[::-1]  # X[0]=x, X[1]=y, X[2]=z, arrays (nz,ny,nx)
        self.d = (cfg["dx"], cfg["dy"], cfg["dz"])

    def _phi(self, k, p):
        return k[0] * self.X[0] + k[1] * self.X[1] + k[2] * self.X[2] + p

    def val(self):
        return self.c0 + sum(a * np.cos(self._phi(k, p)) for a, k, p in self.modes)

    def d1(self, i):
        out = 0.0
        for a, k, p in self.modes:
            t = k[i] * self.d[i]
            K1 = (45 * np.sin(t) - 9 * np.sin(2 * t) + np.sin(3 * t)) / (30 * self.d[i])
            out = out - a * K1 * np.sin(self._phi(k, p))
        return out + 0.0 * self.X[0]

    def d2(self, i, j):
        out = 0.0
        for a, k, p in self.modes:
            ti, tj = k[i] * self.d[i], k[j] * self.d[j]
            if i == j:
                K = (490 - 2 * (270 * np.cos(ti) - 27 * np.cos(2 * ti) + 2 * np.cos(3 * ti))) / (180 * self.d[i] ** 2)
            else:
                K = 4 * (270 * np.sin(ti) * np.sin(tj) - 27 * np.sin(2 * ti) * np.sin(2 * tj)
                         + 2 * np.sin(3 * ti) * np.sin(3 * tj)) / (720 * self.d[i] * self.d[j])
            out = out - a * K * np.cos(self._phi(k, p))
        return out + 0.0 * self.X[0]

    def lap(self):
        return self.d2(0, 0) + self.d2(1, 1) + self.d2(2, 2)

    def d5(self, i):
        out = 0.0
        for a, k, p in self.modes:
            t = k[i] * self.d[i]
            K5 = (5 * np.sin(t) - 4 * np.sin(2 * t) + np.sin(3 * t)) / self.d[i] ** 5
            out = out - a * K5 * np.sin(self._phi(k, p))
        return out + 0.0 * self.X[0]

    def d6(self, i, upwind=False):
        out = 0.0
        for a, k, p in self.modes:
            t = k[i] * self.d[i]
            K6 = (-20 + 30 * np.cos(t) - 12 * np.cos(2 * t) + 2 * np.cos(3 * t))
            K6 = K6 / (60 * self.d[i]) if upwind else K6 / self.d[i] ** 6
            out = out + a * K6 * np.cos(self._phi(k, p))
        return out + 0.0 * self.X[0]

    def del6(self):
        return self.d6(0) + self.d6(1) + self.d6(2)

    def upw(self, u, mask=-1):
        """sum_j |u_j| D6up_j f over the directions j != mask (sub.f90 calc_del6_for_upwind)"""
        return sum(np.abs(u[j]) * self.d6(j, upwind=True) for j in range(3) if j != mask)


def analytic_rhs(cfg, W):
    """W: dict variable name -> Waves.  Returns dict name -> df/dt on the interior (the equations, term by term)."""
    ent, mag = cfg["iss"] >= 0, cfg["iax"] >= 0
    u = [W[n].val() for n in ("ux", "uy", "uz")]
    uij = [[W[n].d1(j) for j in range(3)] for n in ("ux", "uy", "uz")]
    lnrho, gl = W["lnrho"].val(), [W["lnrho"].d1(j) for j in range(3)]
    divu = uij[0][0] + uij[1][1] + uij[2][2]
    out = {}
    # continuity (density.f90:2887-3060): dlnrho/dt = -u.grad lnrho - div u
    out["lnrho"] = -sum(u[j] * gl[j] for j in range(3)) - divu
    if cfg.get("lupw_lnrho"):  # u.grad lnrho -> u.grad lnrho - sum |u_j| D6up_j lnrho
        out["lnrho"] = out["lnrho"] + W["lnrho"].upw(u)
    if cfg.get("ldiff_hyper3lnrho"):  # density.f90: diffrho_hyper3 del6 lnrho
        out["lnrho"] = out["lnrho"] + cfg["diffrho_hyper3"] * W["lnrho"].del6()
    # momentum: advection
    ru = [-sum(u[j] * uij[i][j] for j in range(3)) for i in range(3)]
    if cfg.get("lupw_uu"):
        ru = [ru[i] + W[("ux", "uy", "uz")[i]].upw(u) for i in range(3)]
    # viscosity nu-const (viscosity.f90:1633-1664): nu (del2 u + 1/3 grad div u + 2 S.glnrho)
    S = [[0.5 * (uij[i][j] + uij[j][i]) - (divu / 3.0 if i == j else 0.0) for j in range(3)] for i in range(3)]
    names = ("ux", "uy", "uz")
    nu = cfg["nu"] if cfg["lvisc_nu_const"] else 0.0
    for i in range(3):
        graddiv = sum(W[names[j]].d2(i, j) for j in range(3))
        ru[i] = ru[i] + nu * (W[names[i]].lap() + graddiv / 3.0 + 2.0 * sum(S[i][j] * gl[j] for j in range(3)))
    for i in range(3):  # viscosity.f90: hyper3-simplified nu3 del6 u; hyper3_nu-const nu3 (del6 u + uij5.glnrho)
        if cfg.get("lvisc_hyper3_simplified"):
            ru[i] = ru[i] + cfg["nu_hyper3"] * W[names[i]].del6()
        if cfg.get("lvisc_hyper3_nu_const"):
            ru[i] = ru[i] + cfg["nu_hyper3"] * (W[names[i]].del6() + sum(W[names[i]].d5(j) * gl[j] for j in range(3)))
    S2 = sum(S[i][j] ** 2 for i in range(3) for j in range(3))
    gam, cp = cfg["gamma"], cfg["cp"]
    gm1, cv1 = gam - 1.0, gam / cp
    if ent:
        ss, gss = W["ss"].val(), [W["ss"].d1(j) for j in range(3)]
        e = cv1 * ss + gm1 * (lnrho - cfg["lnrho0"])
        cs2 = cfg["cs20"] * np.exp(e)
        TT = cfg["cs20"] / (cp * gm1) * np.exp(e)  # eos_idealgas.f90: cs20 = gamma_m1 cp TT0
        glnTT = [gm1 * gl[j] + cv1 * gss[j] for j in range(3)]
        for i in range(3):  # entropy.f90:3596-3600: -cs2 (glnrho + glnTT)/gamma
            ru[i] = ru[i] - cs2 * (gl[i] + glnTT[i]) / gam
        rs = -sum(u[j] * gss[j] for j in range(3))
        if cfg.get("lupw_ss"):
            rs = rs + W["ss"].upw(u)
        if cfg["lviscosity_heat"]:
            rs = rs + 2.0 * nu * S2 / TT
        if cfg["lheatc_chiconst"]:  # entropy.f90 calc_heatcond_constchi: cp chi (del2 lnTT + (glnrho+glnTT).glnTT)
            del2lnTT = gm1 * W["lnrho"].lap() + cv1 * W["ss"].lap()
            rs = rs + cp * cfg["chi"] * (del2lnTT + sum((gl[j] + glnTT[j]) * glnTT[j] for j in range(3)))
    else:
        for i in range(3):  # noentropy.f90: -cs20 glnrho
            ru[i] = ru[i] - cfg["cs20"] * gl[i]
    if mag:
        an = ("ax", "ay", "az")
        aij = [[W[n].d1(j) for j in range(3)] for n in an]
        bb = [aij[2][1] - aij[1][2], aij[0][2] - aij[2][0], aij[1][0] - aij[0][1]]
        del2a = [W[n].lap() for n in an]
        graddiva = [sum(W[an[j]].d2(i, j) for j in range(3)) for i in range(3)]
        jj = [(graddiva[i] - del2a[i]) / cfg["mu0"] for i in range(3)]  # magnetic.f90: jj = mu01 (graddivA - del2A)
        eta = cfg["eta"] if cfg["lresi_eta_const"] else 0.0
        uxb = [u[1] * bb[2] - u[2] * bb[1], u[2] * bb[0] - u[0] * bb[2], u[0] * bb[1] - u[1] * bb[0]]
        for i in range(3):
            out[an[i]] = uxb[i] + eta * del2a[i]
            if cfg.get("lupw_aa"):  # magnetic.f90:6257-6313: + sum_{j != i} |u_j| D6up_j A_i (u x B itself is unchanged)
                out[an[i]] = out[an[i]] + W[an[i]].upw(u, mask=i)
        rho1 = np.exp(-lnrho)
        jxb = [jj[1] * bb[2] - jj[2] * bb[1], jj[2] * bb[0] - jj[0] * bb[2], jj[0] * bb[1] - jj[1] * bb[0]]
        if cfg["llorentzforce"]:
            for i in range(3):
                ru[i] = ru[i] + rho1 * jxb[i]
        if ent and cfg["lohmic_heat"]:
            rs = rs + eta * cfg["mu0"] * sum(c * c for c in jj) * rho1 / TT
    for i in range(3):
        out[names[i]] = ru[i]
    if ent:
        out["ss"] = rs
    return out


def _fill(cfg, W):
    nv = cfg["nvar"]
    f = np.zeros((nv, cfg["nz"] + 6, cfg["ny"] + 6, cfg["nx"] + 6))
    slot = {"ux": 0, "uy": 1, "uz": 2, "lnrho": 3}
    if cfg["iss"] >= 0:
        slot["ss"] = cfg["iss"]
    if cfg["iax"] >= 0:
        slot.update(ax=cfg["iax"], ay=cfg["iax"] + 1, az=cfg["iax"] + 2)
    for n, w in W.items():
        f[slot[n], 3:-3, 3:-3, 3:-3] = w.val()
    return f, slot


CASES = {
    # isothermal MHD, oblique waves in all three directions: curl, del2, all three bidiagonal pairs, u x B, j x B / rho
    "iso_mhd_oblique": (dict(entropy=False, magnetic=True, nu=7e-3, eta=4e-3), dict(
        ux=(0.11, [(0.05, (1, 2, -1), 0.3)]), uy=(-0.07, [(0.04, (2, -1, 1), 1.1)]), uz=(0.05, [(0.03, (-1, 1, 2), 2.0)]),
        lnrho=(0.2, [(0.08, (1, 1, 1), 0.7), (0.03, (2, 0, -1), 0.1)]),
        ax=(0.0, [(0.06, (1, -2, 1), 0.4), (0.02, (0, 1, 3), 1.3)]), ay=(0.0, [(0.05, (2, 1, -1), 2.2)]),
        az=(0.0, [(0.07, (-1, 1, 1), 0.9), (0.01, (3, 2, 0), 0.2)]))),
    # MHD with entropy: pressure gradient with cs2(lnrho, s), viscous + Ohmic heating, chi-const conduction
    "mhd_entropy_oblique": (dict(entropy=True, magnetic=True, nu=6e-3, eta=5e-3, chi=3e-3), dict(
        ux=(0.05, [(0.06, (1, 1, 0), 0.3), (0.02, (0, 2, 1), 0.8)]), uy=(0.02, [(0.05, (1, -1, 2), 1.7)]),
        uz=(-0.04, [(0.04, (2, 1, 1), 0.5)]), lnrho=(-0.1, [(0.07, (1, 0, 2), 0.2)]),
        ss=(0.05, [(0.04, (0, 1, 1), 1.0), (0.02, (2, -1, 0), 0.6)]),
        ax=(0.0, [(0.05, (1, 1, 1), 0.1)]), ay=(0.0, [(0.04, (-1, 2, 1), 0.7)]), az=(0.0, [(0.06, (1, -1, 2), 1.9)]))),
    # hydro with entropy on an anisotropic box (dx != dy != dz): per-direction weights
    "hydro_entropy_aniso": (dict(entropy=True, magnetic=False, nu=5e-3, chi=2e-3, ny=20, nz=24), dict(
        ux=(0.03, [(0.05, (1, 1, 1), 0.3)]), uy=(0.0, [(0.05, (2, -1, 1), 1.2)]), uz=(0.01, [(0.03, (1, 2, -1), 0.4)]),
        lnrho=(0.0, [(0.06, (1, -1, 1), 0.9)]), ss=(0.0, [(0.05, (-1, 1, 2), 0.2)]))),
    # sixth-order hyperdiffusion: hyper3_nu-const viscosity (del6 u + uij5.glnrho) and hyper3 density diffusion; the
    # coefficients are large enough for the terms to carry weight against the 2e-12 tolerance
    "hydro_hyper3_nu_const": (dict(entropy=False, magnetic=False, nu=0.0, lvisc_hyper3_nu_const=1, nu_hyper3=3e-3,
                                   ldiff_hyper3lnrho=1, diffrho_hyper3=2e-3), dict(
        ux=(0.03, [(0.05, (1, 2, -1), 0.3)]), uy=(-0.02, [(0.04, (2, -1, 1), 1.2)]), uz=(0.01, [(0.03, (1, 1, 2), 0.4)]),
        lnrho=(0.1, [(0.06, (2, -1, 1), 0.9), (0.02, (1, 0, 3), 0.5)]))),
    # hyper3-simplified on top of nu-const, with a magnetic field
    "iso_mhd_hyper3_simplified": (dict(entropy=False, magnetic=True, nu=4e-3, eta=3e-3, lvisc_hyper3_simplified=1,
                                       nu_hyper3=2e-3), dict(
        ux=(0.02, [(0.05, (2, 1, -1), 0.2)]), uy=(0.01, [(0.04, (1, -2, 1), 0.7)]), uz=(-0.03, [(0.03, (1, 1, 1), 1.4)]),
        lnrho=(0.0, [(0.05, (1, -1, 2), 0.3)]),
        ax=(0.0, [(0.05, (1, 2, 1), 0.6)]), ay=(0.0, [(0.04, (2, -1, 1), 1.1)]), az=(0.0, [(0.03, (-1, 1, 2), 0.8)]))),
    # 6th-order upwinding of u, lnrho, s and A (the |u_j| factors keep the closed form pointwise)
    "mhd_entropy_upwind": (dict(entropy=True, magnetic=True, nu=6e-3, eta=5e-3, chi=3e-3, upwind=True), dict(
        ux=(0.05, [(0.09, (1, 1, 0), 0.3), (0.04, (0, 2, 1), 0.8)]), uy=(-0.02, [(0.07, (1, -1, 2), 1.7)]),
        uz=(0.01, [(0.06, (2, 1, 1), 0.5)]), lnrho=(-0.1, [(0.07, (1, 0, 2), 0.2)]),
        ss=(0.05, [(0.04, (0, 1, 1), 1.0), (0.02, (2, -1, 0), 0.6)]),
        ax=(0.0, [(0.05, (1, 1, 1), 0.1)]), ay=(0.0, [(0.04, (-1, 2, 1), 0.7)]), az=(0.0, [(0.06, (1, -1, 2), 1.9)]))),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_rhs_of_plane_waves_matches_the_closed_form(oracle_mod, name):
    kw, spec = CASES[name]
    kw = dict(kw)
    cfg = pcfg.make_config(16, kw.pop("entropy"), kw.pop("magnetic"), **kw)
    W = {n: Waves(cfg, c0, [(a, k, p) for a, k, p in modes]) for n, (c0, modes) in spec.items()}
    f, slot = _fill(cfg, W)
    o = oracle_mod.Oracle(cfg)
    o.update_ghosts(f)
    df = np.zeros_like(f)
    o.rhs(f, df)
    ref = analytic_rhs(cfg, W)
    for n, r in ref.items():
        got = df[slot[n], 3:-3, 3:-3, 3:-3]
        scale = max(np.abs(r).max(), 1e-30)
        err = np.abs(got - r).max() / scale
        assert err <= 2e-12, (name, n, err)


def test_closed_form_derivatives_match_the_oracle_stencils(oracle_mod):
    """the three closed forms themselves against pc_oracle_deriv (der, der2, derij) for an oblique wave"""
    cfg = pcfg.make_config(16, False, False, ny=20, nz=24)
    w = Waves(cfg, 0.3, [(0.7, (2, -1, 3), 0.4), (0.2, (1, 3, -2), 1.5)])
    f = np.zeros((4, cfg["nz"] + 6, cfg["ny"] + 6, cfg["nx"] + 6))
    f[0, 3:-3, 3:-3, 3:-3] = w.val()
    o = oracle_mod.Oracle(cfg)
    o.update_ghosts(f)
    for j in range(3):
        assert np.abs(o.deriv(f, 0, j + 1, 1) - w.d1(j)).max() <= 1e-12
        assert np.abs(o.deriv(f, 0, j + 1, 2) - w.d2(j, j)).max() <= 1e-11
    for kind, (i, j) in ((12, (0, 1)), (23, (1, 2)), (31, (2, 0))):
        assert np.abs(o.deriv(f, 0, 1, kind) - w.d2(i, j)).max() <= 1e-11
    for j in range(3):  # der5, der6 and the upwind form of der6
        for kind, ref in ((5, w.d5(j)), (6, w.d6(j)), (7, w.d6(j, upwind=True))):
            assert np.abs(o.deriv(f, 0, j + 1, kind) - ref).max() <= 2e-12 * np.abs(ref).max(), (kind, j)
