// pc_oracle.cpp -- CPU ORACLE: restatement of Pencil Code's RHS + 2N-RK3 path.
//
// TEST INFRASTRUCTURE ONLY (see pc_oracle.h).  Never linked into the product.
// Every routine cites the reference file:line (relative to /root/reference/src)
// it follows; arithmetic is written in the reference's pencil formulation and
// evaluation order (one (m,n) pencil of length nx at a time).  Build with
//   g++ -O2 -ffp-contract=off -fno-fast-math -fopenmp
// so that no FMA contraction changes rounding relative to `gfortran -O`
// without -march (config/compilers/separate/GNU-gfortran.conf:4-9).
//
// Parity: "coarsely pinned" -- see header of pc_oracle.h.

#include "pc_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

constexpr int NG = 3; // nghost, deriv.f90:7

typedef std::vector<double> V;

struct Grid {
  int nx, ny, nz, mx, my, mz;
  size_t sy, sz, sv; // strides in doubles
  double dx_1, dy_1, dz_1;
  double dline_1[3];
  double dxyz_2, dxyz_4, dxyz_6; // grid.f90:3051-3053
  explicit Grid(const pc_oracle_pars &p) {
    nx = p.nx; ny = p.ny; nz = p.nz;
    mx = nx + 2 * NG; my = ny + 2 * NG; mz = nz + 2 * NG;
    sy = mx; sz = (size_t)mx * my; sv = sz * mz;
    dx_1 = 1.0 / p.dx; dy_1 = 1.0 / p.dy; dz_1 = 1.0 / p.dz;
    dline_1[0] = dx_1; dline_1[1] = dy_1; dline_1[2] = dz_1;
    // grid.f90:3051-3053: dxyz_2 = dline_1(:,1)**2+dline_1(:,2)**2+dline_1(:,3)**2 etc.
    dxyz_2 = dx_1 * dx_1 + dy_1 * dy_1 + dz_1 * dz_1;
    dxyz_4 = std::pow(dx_1, 4) + std::pow(dy_1, 4) + std::pow(dz_1, 4);
    dxyz_6 = std::pow(dx_1, 6) + std::pow(dy_1, 6) + std::pow(dz_1, 6);
  }
  inline size_t idx(int ix, int iy, int iz, int k) const { // 0-based incl. ghosts
    return (size_t)ix + sy * iy + sz * iz + sv * k;
  }
};

// ---------------------------------------------------------------------------
// deriv.f90 -- pencil derivatives at row (m,n) (0-based incl. ghosts)
// ---------------------------------------------------------------------------
struct Deriv {
  const Grid &g;
  const double *f;
  int m, n;
  Deriv(const Grid &g_, const double *f_) : g(g_), f(f_), m(0), n(0) {}

  inline const double *row(int k, int dm, int dn) const {
    return f + g.idx(NG, m + dm, n + dn, k);
  }
  inline size_t stride(int j) const { return j == 1 ? 1 : (j == 2 ? g.sy : g.sz); }
  inline double d1(int j) const { return g.dline_1[j - 1]; }

  // der_main, deriv.f90:89-171
  void der(int k, int j, double *df) const {
    const double a = 1.0 / 60.0;
    const double fac = a * d1(j);
    const double *c = row(k, 0, 0);
    const ptrdiff_t s = (ptrdiff_t)stride(j);
    for (int i = 0; i < g.nx; ++i) {
      const double *q = c + i;
      df[i] = fac * (+45.0 * (q[s] - q[-s]) - 9.0 * (q[2 * s] - q[-2 * s]) + (q[3 * s] - q[-3 * s]));
    }
  }
  // der2_main, deriv.f90:472-561 ; coefficients initialize_deriv :38-40
  void der2(int k, int j, double *df2) const {
    const double c0 = -490. / 180., c1 = 270. / 180., c2 = -27. / 180., c3 = 2. / 180.;
    const double fac = d1(j) * d1(j);
    const double *c = row(k, 0, 0);
    const ptrdiff_t s = (ptrdiff_t)stride(j);
    for (int i = 0; i < g.nx; ++i) {
      const double *q = c + i;
      df2[i] = fac * (c0 * q[0] + c1 * (q[s] + q[-s]) + c2 * (q[2 * s] + q[-2 * s]) + c3 * (q[3 * s] + q[-3 * s]));
    }
  }
  // der5, deriv.f90:855-935 : fac=(1/2)*1/dx**5
  void der5(int k, int j, double *df) const {
    const double h = 1.0 / d1(j);
    const double fac = (1.0 / 2) * 1 / std::pow(h, 5);
    const double *c = row(k, 0, 0);
    const ptrdiff_t s = (ptrdiff_t)stride(j);
    for (int i = 0; i < g.nx; ++i) {
      const double *q = c + i;
      df[i] = fac * (+5.0 * (q[s] - q[-s]) - 4.0 * (q[2 * s] - q[-2 * s]) + (q[3 * s] - q[-3 * s]));
    }
  }
  // der6_main, deriv.f90:936-1072 ; upwind: fac=(1/60)*dx_1 else dx_1**6
  void der6(int k, int j, double *df, bool upwind) const {
    const double fac = upwind ? (1.0 / 60) * d1(j) : std::pow(d1(j), 6);
    const double *c = row(k, 0, 0);
    const ptrdiff_t s = (ptrdiff_t)stride(j);
    for (int i = 0; i < g.nx; ++i) {
      const double *q = c + i;
      df[i] = fac * (-20.0 * q[0] + 15.0 * (q[s] + q[-s]) - 6.0 * (q[2 * s] + q[-2 * s]) + (q[3 * s] + q[-3 * s]));
    }
  }
  // derij_main bidiagonal, deriv.f90:1400-1452 (lbidiagonal_derij=T, cdata.f90:827)
  void derij(int k, int i_, int j_, double *df) const {
    if (i_ == j_) return; // derij leaves df untouched for i==j (sub.f90:2651-2653)
    ptrdiff_t sa, sb;
    double fac;
    if (i_ + j_ == 3) { // x,y
      sa = 1; sb = (ptrdiff_t)g.sy;
      fac = (1. / 720.) * g.dx_1 * g.dy_1;
    } else if (i_ + j_ == 5) { // y,z
      sa = (ptrdiff_t)g.sy; sb = (ptrdiff_t)g.sz;
      fac = (1. / 720.) * g.dy_1 * g.dz_1;
    } else { // x,z
      sa = 1; sb = (ptrdiff_t)g.sz;
      fac = (1. / 720.) * g.dz_1 * g.dx_1;
    }
    const double *c = row(k, 0, 0);
    for (int i = 0; i < g.nx; ++i) {
      const double *q = c + i;
      df[i] = fac * (270. * (q[sa + sb] - q[-sa + sb] + q[-sa - sb] - q[sa - sb]) -
                     27. * (q[2 * sa + 2 * sb] - q[-2 * sa + 2 * sb] + q[-2 * sa - 2 * sb] - q[2 * sa - 2 * sb]) +
                     2. * (q[3 * sa + 3 * sb] - q[-3 * sa + 3 * sb] + q[-3 * sa - 3 * sb] - q[3 * sa - 3 * sb]));
    }
  }
};

// ---------------------------------------------------------------------------
// pencil case (cparam_pencils.inc subset) -- arrays of length nx
// vec: [j*nx + i], mat: [(i_*3+j_)*nx + i] for p%uij(:,i_,j_)
// ---------------------------------------------------------------------------
struct Pencil {
  int nx;
  V uu, u2, uij, divu, sij, sij2, oo, ugu, del2u, graddivu, del6u, uij5;
  V lnrho, rho, rho1, glnrho, uglnrho, sglnrho, del2lnrho, del6lnrho, uij5glnrho;
  V ss, gss, del2ss, cs2, lnTT, TT1, glnTT, del2lnTT;
  V ugss, fpres, advec_cs2;
  V fvisc, visc_heat, diffus_nu, diffus_nu3, diffus_chi;
  V aa, aij, bb, b2, uxb, bij, del2a, jj, j2, jxb, jxbr, ab, jb, ou;
  V advec_uu, advec_va2;
  V tmp, tmp2;
  explicit Pencil(int n) : nx(n) {
    auto s = [&](V &v, int c) { v.assign((size_t)c * n, 0.0); };
    s(uu, 3); s(u2, 1); s(uij, 9); s(divu, 1); s(sij, 9); s(sij2, 1); s(oo, 3); s(ugu, 3);
    s(del2u, 3); s(graddivu, 3); s(del6u, 3); s(uij5, 9);
    s(lnrho, 1); s(rho, 1); s(rho1, 1); s(glnrho, 3); s(uglnrho, 1); s(sglnrho, 3);
    s(del2lnrho, 1); s(del6lnrho, 1); s(uij5glnrho, 3);
    s(ss, 1); s(gss, 3); s(del2ss, 1); s(cs2, 1); s(lnTT, 1); s(TT1, 1); s(glnTT, 3); s(del2lnTT, 1);
    s(ugss, 1); s(fpres, 3); s(advec_cs2, 1);
    s(fvisc, 3); s(visc_heat, 1); s(diffus_nu, 1); s(diffus_nu3, 1); s(diffus_chi, 1);
    s(aa, 3); s(aij, 9); s(bb, 3); s(b2, 1); s(uxb, 3); s(bij, 9); s(del2a, 3); s(jj, 3); s(j2, 1);
    s(jxb, 3); s(jxbr, 3); s(ab, 1); s(jb, 1); s(ou, 1);
    s(advec_uu, 1); s(advec_va2, 1); s(tmp, 1); s(tmp2, 1);
  }
};

// per-pencil CFL accumulators (equ.f90:1482-1494)
struct Cfl {
  V advec2, maxadvec, maxdiffus, maxdiffus2, maxdiffus3;
  explicit Cfl(int n) : advec2(n), maxadvec(n), maxdiffus(n), maxdiffus2(n), maxdiffus3(n) {}
  void zero() {
    std::fill(advec2.begin(), advec2.end(), 0.0);
    std::fill(maxadvec.begin(), maxadvec.end(), 0.0);
    std::fill(maxdiffus.begin(), maxdiffus.end(), 0.0);
    std::fill(maxdiffus2.begin(), maxdiffus2.end(), 0.0);
    std::fill(maxdiffus3.begin(), maxdiffus3.end(), 0.0);
  }
};

struct Eos { // eos_idealgas.f90:154-270 (units_eos with cp given)
  double cs20, gamma, gamma_m1, gamma1, cp, cp1, cv, cv1, lnrho0, lnTT0;
  explicit Eos(const pc_oracle_pars &p) {
    cs20 = p.cs20; gamma = p.gamma; lnrho0 = p.lnrho0;
    gamma_m1 = gamma - 1.0; gamma1 = 1 / gamma;
    cp = p.cp; cp1 = 1 / cp; cv = gamma1 * cp; cv1 = gamma * cp1;
    if (gamma_m1 != 0.0) lnTT0 = std::log(cs20 / (cp * gamma_m1)); // :266
    else lnTT0 = std::log(cs20 / cp);                               // :268
  }
};

// ----------------------------- sub.f90 helpers -----------------------------
// dot_mn sub.f90:599-640 : a1*b1+a2*b2+a3*b3
inline void dot_mn(int nx, const double *a, const double *b, double *c) {
  for (int i = 0; i < nx; ++i) c[i] = a[i] * b[i] + a[nx + i] * b[nx + i] + a[2 * nx + i] * b[2 * nx + i];
}
// dot2_mn sub.f90:872-900
inline void dot2_mn(int nx, const double *a, double *b) {
  for (int i = 0; i < nx; ++i) b[i] = a[i] * a[i] + a[nx + i] * a[nx + i] + a[2 * nx + i] * a[2 * nx + i];
}
// cross_mn sub.f90:1456-1480
inline void cross_mn(int nx, const double *a, const double *b, double *c) {
  for (int i = 0; i < nx; ++i) {
    c[i] = a[nx + i] * b[2 * nx + i] - a[2 * nx + i] * b[nx + i];
    c[nx + i] = a[2 * nx + i] * b[i] - a[i] * b[2 * nx + i];
    c[2 * nx + i] = a[i] * b[nx + i] - a[nx + i] * b[i];
  }
}
// multmv_mn sub.f90:1228-1260 : c_i = a_ij b_j, sum order j=1,2,3
inline void multmv_mn(int nx, const double *a, const double *b, double *c) {
  for (int i_ = 0; i_ < 3; ++i_)
    for (int i = 0; i < nx; ++i) {
      double t = a[(i_ * 3 + 0) * nx + i] * b[i];
      t = t + a[(i_ * 3 + 1) * nx + i] * b[nx + i];
      t = t + a[(i_ * 3 + 2) * nx + i] * b[2 * nx + i];
      c[i_ * nx + i] = t;
    }
}
// curl_mn sub.f90:2041-2097 (Cartesian)
inline void curl_mn(int nx, const double *aij, double *b) {
  for (int i = 0; i < nx; ++i) {
    b[i] = aij[(2 * 3 + 1) * nx + i] - aij[(1 * 3 + 2) * nx + i];
    b[nx + i] = aij[(0 * 3 + 2) * nx + i] - aij[(2 * 3 + 0) * nx + i];
    b[2 * nx + i] = aij[(1 * 3 + 0) * nx + i] - aij[(0 * 3 + 1) * nx + i];
  }
}
// gij sub.f90:1505-1544 : g(:,i,j) = der(f,k+i-1,j)
inline void gij(const Deriv &D, int k, double *g, int nx) {
  for (int i_ = 0; i_ < 3; ++i_)
    for (int j_ = 0; j_ < 3; ++j_) D.der(k + i_, j_ + 1, g + (i_ * 3 + j_) * nx);
}
// grad_main sub.f90:1610-1629
inline void grad(const Deriv &D, int k, double *g, int nx) {
  for (int j_ = 0; j_ < 3; ++j_) D.der(k, j_ + 1, g + j_ * nx);
}
// del2_main sub.f90:2304-2337 : d2fdx+d2fdy+d2fdz
inline void del2(const Deriv &D, int k, double *out, int nx, V &t1, V &t2) {
  D.der2(k, 1, out);
  D.der2(k, 2, t1.data());
  D.der2(k, 3, t2.data());
  for (int i = 0; i < nx; ++i) out[i] = out[i] + t1[i] + t2[i];
}
// del6 sub.f90:3222-3250
inline void del6(const Deriv &D, int k, double *out, int nx, V &t1, V &t2) {
  D.der6(k, 1, out, false);
  D.der6(k, 2, t1.data(), false);
  D.der6(k, 3, t2.data(), false);
  for (int i = 0; i < nx; ++i) out[i] = out[i] + t1[i] + t2[i];
}
// calc_del6_for_upwind + doupwind sub.f90:7491-7574
inline void doupwind(const Deriv &D, int k, const double *uu, double *ugradf, int mask, int nx) {
  V del6f((size_t)3 * nx, 0.0), tmp(nx);
  for (int j_ = 0; j_ < 3; ++j_) {
    if (j_ + 1 == mask) {
      for (int i = 0; i < nx; ++i) del6f[j_ * nx + i] = 0.;
    } else {
      D.der6(k, j_ + 1, tmp.data(), true);
      for (int i = 0; i < nx; ++i) del6f[j_ * nx + i] = std::fabs(uu[j_ * nx + i]) * tmp[i];
    }
  }
  for (int i = 0; i < nx; ++i) {
    double s = del6f[i] + del6f[nx + i] + del6f[2 * nx + i]; // sum(del6f,2)
    if (mask > 0) ugradf[i] = ugradf[i] + s;
    else ugradf[i] = ugradf[i] - s;
  }
}
// u_dot_grad_scl sub.f90:3650-3683
inline void u_dot_grad_scl(const Deriv &D, int k, const double *gradf, const double *uu, double *ugradf,
                           bool upwind, int nx) {
  dot_mn(nx, uu, gradf, ugradf);
  if (upwind) doupwind(D, k, uu, ugradf, 0, nx);
}

// step, der_step sub.f90:4957-5015
inline double step_fn(double x, double x0, double width) { return 0.5 * (1 + std::tanh((x - x0) / width)); }
inline double der_step_fn(double x, double x0, double width) {
  const double arg = std::fabs((x - x0) / width);
  if (arg >= 8.) return 2. / width * std::exp(-2. * arg);
  return 0.5 / (width * std::cosh(arg) * std::cosh(arg));
}

struct Ctx {
  const pc_oracle_pars &P;
  Grid g;
  Eos eos;
  V hcond_prof, dlnhcond_prof, profz_cool; // per interior z plane (entropy.f90:7741-7769, 1323+1367)
  explicit Ctx(const pc_oracle_pars &p) : P(p), g(p), eos(p) {
    if (p.lheatc_Kprof) {
      hcond_prof.resize(p.nz); dlnhcond_prof.resize(p.nz);
      for (int n = 0; n < p.nz; ++n) {
        const double z = p.z0 + n * p.dz;
        double h = 1. + (p.hcond1 - 1.) * step_fn(z, p.z1, -p.widthss) + (p.hcond2 - 1.) * step_fn(z, p.z2, p.widthss);
        dlnhcond_prof[n] = ((p.hcond1 - 1.) * der_step_fn(z, p.z1, -p.widthss) + (p.hcond2 - 1.) * der_step_fn(z, p.z2, p.widthss)) / h;
        hcond_prof[n] = p.hcond0 * h;
      }
    }
    if (p.cool != 0.0) {
      profz_cool.resize(p.nz);
      for (int n = 0; n < p.nz; ++n) {
        const double z = p.z0 + n * p.dz;
        const double q = (p.ztop - z) / p.wcool;
        profz_cool[n] = p.cool * std::exp(-0.5 * (q * q)); // cooling_profile='gaussian' (entropy.f90:1323,1367)
      }
    }
  }
  bool lhydro() const { return P.iux >= 0; }
  bool ldensity() const { return P.ilnrho >= 0; }
  bool lentropy() const { return P.iss >= 0; }
  bool lmagnetic() const { return P.iax >= 0; }
  bool lhyper_visc() const { return P.lvisc_hyper3_simplified || P.lvisc_hyper3_nu_const; }
};

// ---------------------------------------------------------------------------
// calc_all_pencils, equ.f90:887-993 : hydro -> density -> eos -> energy -> viscosity -> magnetic
// ---------------------------------------------------------------------------
void calc_pencils_hydro(const Ctx &c, const Deriv &D, Pencil &p, Cfl &cfl, bool lcourant) {
  const int nx = p.nx, iux = c.P.iux;
  // hydro.f90:3925-4063 / 3564-3809
  for (int j = 0; j < 3; ++j) {
    const double *r = D.row(iux + j, 0, 0);
    for (int i = 0; i < nx; ++i) p.uu[j * nx + i] = r[i];
  }
  dot2_mn(nx, p.uu.data(), p.u2.data());
  gij(D, iux, p.uij.data(), nx);
  // div_mn sub.f90:1981 : a(1,1)+a(2,2)+a(3,3)
  for (int i = 0; i < nx; ++i) p.divu[i] = p.uij[0 * nx + i] + p.uij[4 * nx + i] + p.uij[8 * nx + i];
  // traceless_strain sub.f90:8922-8990
  for (int j = 0; j < 3; ++j) {
    for (int i = 0; i < nx; ++i) p.sij[(j * 3 + j) * nx + i] = p.uij[(j * 3 + j) * nx + i] - (1. / 3.) * p.divu[i];
    for (int i_ = j + 1; i_ < 3; ++i_)
      for (int i = 0; i < nx; ++i) {
        double s = .5 * (p.uij[(i_ * 3 + j) * nx + i] + p.uij[(j * 3 + i_) * nx + i]);
        p.sij[(i_ * 3 + j) * nx + i] = s;
        p.sij[(j * 3 + i_) * nx + i] = s;
      }
  }
  // multm2_sym_mn sub.f90:1207-1226
  for (int i = 0; i < nx; ++i) {
    double b = p.sij[0 * nx + i] * p.sij[0 * nx + i];
    for (int i_ = 1; i_ < 3; ++i_) {
      b = b + p.sij[(i_ * 3 + i_) * nx + i] * p.sij[(i_ * 3 + i_) * nx + i];
      for (int j = 0; j < i_; ++j) b = b + 2 * (p.sij[(i_ * 3 + j) * nx + i] * p.sij[(i_ * 3 + j) * nx + i]);
    }
    p.sij2[i] = b;
  }
  curl_mn(nx, p.uij.data(), p.oo.data());
  dot_mn(nx, p.oo.data(), p.uu.data(), p.ou.data());
  // ugu: u_dot_grad_vec sub.f90:3482-3534, UPWIND=lupw_uu
  {
    V ff((size_t)3 * nx);
    for (int j = 0; j < 3; ++j) {
      for (int k = 0; k < 3; ++k)
        for (int i = 0; i < nx; ++i) ff[k * nx + i] = p.uij[(j * 3 + k) * nx + i];
      u_dot_grad_scl(D, iux + j, ff.data(), p.uu.data(), p.ugu.data() + j * nx, c.P.lupw_uu != 0, nx);
    }
  }
  // del2v_etc sub.f90:2610-2749
  {
    V fijj((size_t)9 * nx), fjji((size_t)9 * nx), tmp(nx);
    for (int i_ = 0; i_ < 3; ++i_)
      for (int j = 0; j < 3; ++j) {
        D.der2(iux + i_, j + 1, tmp.data());
        for (int i = 0; i < nx; ++i) fijj[(i_ * 3 + j) * nx + i] = tmp[i];
        D.derij(iux + j, j + 1, i_ + 1, tmp.data()); // untouched when i==j -> keeps der2 value
        for (int i = 0; i < nx; ++i) fjji[(i_ * 3 + j) * nx + i] = tmp[i];
      }
    for (int i_ = 0; i_ < 3; ++i_)
      for (int i = 0; i < nx; ++i) {
        p.del2u[i_ * nx + i] = fijj[(i_ * 3 + 0) * nx + i] + fijj[(i_ * 3 + 1) * nx + i] + fijj[(i_ * 3 + 2) * nx + i];
        p.graddivu[i_ * nx + i] = fjji[(i_ * 3 + 0) * nx + i] + fjji[(i_ * 3 + 1) * nx + i] + fjji[(i_ * 3 + 2) * nx + i];
      }
  }
  if (c.lhyper_visc()) {
    // del6v sub.f90:2834-2860
    for (int i_ = 0; i_ < 3; ++i_) del6(D, iux + i_, p.del6u.data() + i_ * nx, nx, p.tmp, p.tmp2);
  }
  if (c.P.lvisc_hyper3_nu_const) {
    // hydro.f90:3564-3809: uij5(:,i,j)=der5(f,iuu+i-1,j)
    for (int i_ = 0; i_ < 3; ++i_)
      for (int j = 0; j < 3; ++j) D.der5(iux + i_, j + 1, p.uij5.data() + (i_ * 3 + j) * nx);
  }
  // hydro.f90:3492-3531 advec_uu=sum(abs(uu)*dline_1,2); maxadvec=max(maxadvec,advec_uu)
  if (lcourant) {
    for (int i = 0; i < nx; ++i) {
      p.advec_uu[i] = std::fabs(p.uu[i]) * c.g.dline_1[0] + std::fabs(p.uu[nx + i]) * c.g.dline_1[1] +
                      std::fabs(p.uu[2 * nx + i]) * c.g.dline_1[2];
      cfl.maxadvec[i] = std::max(cfl.maxadvec[i], p.advec_uu[i]);
    }
  }
}

void calc_pencils_density(const Ctx &c, const Deriv &D, Pencil &p) {
  const int nx = p.nx, il = c.P.ilnrho;
  // density.f90:2727-2811
  const double *r = D.row(il, 0, 0);
  for (int i = 0; i < nx; ++i) p.lnrho[i] = r[i];
  for (int i = 0; i < nx; ++i) p.rho1[i] = std::exp(-r[i]); // :2750
  for (int i = 0; i < nx; ++i) p.rho[i] = std::exp(r[i]);   // rho=exp(lnrho)
  grad(D, il, p.glnrho.data(), nx);
  u_dot_grad_scl(D, il, p.glnrho.data(), p.uu.data(), p.uglnrho.data(), c.P.lupw_lnrho != 0, nx);
  if (c.lentropy() && (c.P.lheatc_chiconst || c.P.lheatc_Kprof)) del2(D, il, p.del2lnrho.data(), nx, p.tmp, p.tmp2);
  if (c.P.ldiff_hyper3lnrho) del6(D, il, p.del6lnrho.data(), nx, p.tmp, p.tmp2);
  multmv_mn(nx, p.sij.data(), p.glnrho.data(), p.sglnrho.data());
  if (c.P.lvisc_hyper3_nu_const) multmv_mn(nx, p.uij5.data(), p.glnrho.data(), p.uij5glnrho.data());
}

void calc_pencils_eos(const Ctx &c, const Deriv &D, Pencil &p) {
  const int nx = p.nx;
  const Eos &e = c.eos;
  if (c.lentropy()) {
    // eos_idealgas.f90:851-888 (ilnrho_ss)
    const double *r = D.row(c.P.iss, 0, 0);
    for (int i = 0; i < nx; ++i) p.ss[i] = r[i];
    grad(D, c.P.iss, p.gss.data(), nx);
    if (c.P.lheatc_chiconst || c.P.lheatc_Kprof) del2(D, c.P.iss, p.del2ss.data(), nx, p.tmp, p.tmp2);
    for (int i = 0; i < nx; ++i) p.cs2[i] = e.cs20 * std::exp(e.cv1 * p.ss[i] + e.gamma_m1 * (p.lnrho[i] - e.lnrho0));
    for (int i = 0; i < nx; ++i) p.lnTT[i] = e.lnTT0 + e.cv1 * p.ss[i] + e.gamma_m1 * (p.lnrho[i] - e.lnrho0);
    for (int i = 0; i < nx; ++i) p.TT1[i] = std::exp(-p.lnTT[i]);
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < nx; ++i) p.glnTT[j * nx + i] = e.gamma_m1 * p.glnrho[j * nx + i] + e.cv1 * p.gss[j * nx + i];
    if (c.P.lheatc_chiconst || c.P.lheatc_Kprof)
      for (int i = 0; i < nx; ++i) p.del2lnTT[i] = e.gamma_m1 * p.del2lnrho[i] + e.cv1 * p.del2ss[i];
  } else {
    // eos_idealgas.f90:969-1000 (ilnrho_cs2, isothermal): cs2=cs20
    for (int i = 0; i < nx; ++i) p.cs2[i] = e.cs20;
  }
}

void calc_pencils_energy(const Ctx &c, const Deriv &D, Pencil &p) {
  const int nx = p.nx;
  if (c.lentropy()) {
    // entropy.f90:3554-3653
    u_dot_grad_scl(D, c.P.iss, p.gss.data(), p.uu.data(), p.ugss.data(), c.P.lupw_ss != 0, nx);
    const double gamma1 = 1. / c.eos.gamma;
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < nx; ++i)
        p.fpres[j * nx + i] = -p.cs2[i] * (p.glnrho[j * nx + i] + p.glnTT[j * nx + i]) * gamma1; // :3596-3600
  } else {
    // noentropy.f90:252-325: fpres=-cs2*glnrho
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < nx; ++i) p.fpres[j * nx + i] = -p.cs2[i] * p.glnrho[j * nx + i];
  }
  for (int i = 0; i < nx; ++i) p.advec_cs2[i] = p.cs2[i] * c.g.dxyz_2;
}

void calc_pencils_viscosity(const Ctx &c, Pencil &p, Cfl &cfl, bool lcourant) {
  const int nx = p.nx;
  // viscosity.f90:1411-1476: zero fvisc, visc_heat, diffus_total
  std::fill(p.fvisc.begin(), p.fvisc.end(), 0.0);
  std::fill(p.visc_heat.begin(), p.visc_heat.end(), 0.0);
  std::fill(p.diffus_nu.begin(), p.diffus_nu.end(), 0.0);
  std::fill(p.diffus_nu3.begin(), p.diffus_nu3.end(), 0.0);
  if (c.P.lvisc_nu_const) {
    // :1633-1664
    const double nu = c.P.nu;
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < nx; ++i)
        p.fvisc[j * nx + i] = p.fvisc[j * nx + i] +
                              nu * (p.del2u[j * nx + i] + 2. * p.sglnrho[j * nx + i] + 1. / 3. * p.graddivu[j * nx + i]);
    for (int i = 0; i < nx; ++i) p.visc_heat[i] = p.visc_heat[i] + 2 * nu * p.sij2[i];
    for (int i = 0; i < nx; ++i) p.diffus_nu[i] = p.diffus_nu[i] + nu;
  }
  if (c.P.lvisc_hyper3_simplified) {
    // :1982-1989
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < nx; ++i) p.fvisc[j * nx + i] = p.fvisc[j * nx + i] + c.P.nu_hyper3 * p.del6u[j * nx + i];
    for (int i = 0; i < nx; ++i) p.diffus_nu3[i] = p.diffus_nu3[i] + c.P.nu_hyper3;
  }
  if (c.P.lvisc_hyper3_nu_const) {
    // :2250-2257
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < nx; ++i)
        p.fvisc[j * nx + i] = p.fvisc[j * nx + i] + c.P.nu_hyper3 * (p.del6u[j * nx + i] + p.uij5glnrho[j * nx + i]);
    for (int i = 0; i < nx; ++i) p.diffus_nu3[i] = p.diffus_nu3[i] + c.P.nu_hyper3;
  }
  if (lcourant) {
    // :2433-2459
    for (int i = 0; i < nx; ++i) {
      p.diffus_nu[i] = p.diffus_nu[i] * c.g.dxyz_2;
      p.diffus_nu3[i] = p.diffus_nu3[i] * c.g.dxyz_6;
      cfl.maxdiffus[i] = std::max(cfl.maxdiffus[i], p.diffus_nu[i]);
      cfl.maxdiffus3[i] = std::max(cfl.maxdiffus3[i], p.diffus_nu3[i]);
    }
  }
}

void calc_pencils_magnetic(const Ctx &c, const Deriv &D, Pencil &p, Cfl &cfl, bool lcourant) {
  const int nx = p.nx, iax = c.P.iax;
  const double mu01 = 1. / c.P.mu0;
  // magnetic.f90:4279-5225 (pencpar)
  for (int j = 0; j < 3; ++j) {
    const double *r = D.row(iax + j, 0, 0);
    for (int i = 0; i < nx; ++i) p.aa[j * nx + i] = r[i];
  }
  gij(D, iax, p.aij.data(), nx);
  curl_mn(nx, p.aij.data(), p.bb.data());
  dot_mn(nx, p.aa.data(), p.bb.data(), p.ab.data());
  dot2_mn(nx, p.bb.data(), p.b2.data());
  cross_mn(nx, p.uu.data(), p.bb.data(), p.uxb.data());
  // gij_etc sub.f90:2945-3113 -> bij, del2a
  {
    V d2A((size_t)27 * nx); // d2A(:,a,b,i) at [((a*3+b)*3+i)*nx]
    auto at = [&](int a, int b, int i_) { return d2A.data() + ((size_t)(a * 3 + b) * 3 + i_) * nx; };
    for (int i_ = 0; i_ < 3; ++i_) {
      for (int j = 0; j < 3; ++j) D.der2(iax + i_, j + 1, at(j, j, i_));
      D.derij(iax + i_, 2, 3, at(1, 2, i_)); std::memcpy(at(2, 1, i_), at(1, 2, i_), sizeof(double) * nx);
      D.derij(iax + i_, 3, 1, at(2, 0, i_)); std::memcpy(at(0, 2, i_), at(2, 0, i_), sizeof(double) * nx);
      D.derij(iax + i_, 1, 2, at(0, 1, i_)); std::memcpy(at(1, 0, i_), at(0, 1, i_), sizeof(double) * nx);
    }
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < nx; ++i) {
        p.bij[(0 * 3 + j) * nx + i] = at(1, j, 2)[i] - at(2, j, 1)[i];
        p.bij[(1 * 3 + j) * nx + i] = at(2, j, 0)[i] - at(0, j, 2)[i];
        p.bij[(2 * 3 + j) * nx + i] = at(0, j, 1)[i] - at(1, j, 0)[i];
      }
    for (int i_ = 0; i_ < 3; ++i_)
      for (int i = 0; i < nx; ++i) p.del2a[i_ * nx + i] = at(0, 0, i_)[i] + at(1, 1, i_)[i] + at(2, 2, i_)[i];
  }
  // jj=mu01*curl_mn(bij) :4735
  curl_mn(nx, p.bij.data(), p.jj.data());
  for (int i = 0; i < 3 * nx; ++i) p.jj[i] = mu01 * p.jj[i];
  dot2_mn(nx, p.jj.data(), p.j2.data());
  dot_mn(nx, p.jj.data(), p.bb.data(), p.jb.data());
  cross_mn(nx, p.jj.data(), p.bb.data(), p.jxb.data());
  // jxbr=rho1*jxb :4870-4888
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < nx; ++i) p.jxbr[j * nx + i] = p.rho1[i] * p.jxb[j * nx + i];
  // :5160-5217
  if (lcourant) {
    for (int i = 0; i < nx; ++i) {
      if (c.lhydro() && c.P.llorentzforce) {
        double s = (p.bb[i] * c.g.dline_1[0]) * (p.bb[i] * c.g.dline_1[0]) +
                   (p.bb[nx + i] * c.g.dline_1[1]) * (p.bb[nx + i] * c.g.dline_1[1]) +
                   (p.bb[2 * nx + i] * c.g.dline_1[2]) * (p.bb[2 * nx + i] * c.g.dline_1[2]);
        p.advec_va2[i] = s * mu01 * p.rho1[i];
      } else {
        p.advec_va2[i] = 0.;
      }
      cfl.advec2[i] = cfl.advec2[i] + p.advec_va2[i];
    }
  }
}

// ---------------------------------------------------------------------------
// calc_all_rhs, equ.f90:1223-1366 : duu_dt, dlnrho_dt, denergy_dt, daa_dt
// ---------------------------------------------------------------------------
void calc_all_rhs(const Ctx &c, const Deriv &D, Pencil &p, Cfl &cfl, double *df, bool lcourant) {
  const int nx = p.nx;
  const Grid &g = c.g;
  auto dfrow = [&](int k) { return df + g.idx(NG, D.m, D.n, k); };
  // duu_dt hydro.f90:4421-4684 -> advec_uu :4332-4419 : df_u -= ugu ; + fvisc
  for (int j = 0; j < 3; ++j) {
    double *d = dfrow(c.P.iux + j);
    for (int i = 0; i < nx; ++i) d[i] = d[i] - p.ugu[j * nx + i];
  }
  // coriolis_cartesian hydro.f90:6339-6366 (theta=0): c2=2*Omega
  if (c.P.Omega != 0.0) {
    const double c2 = 2 * c.P.Omega;
    double *dx_ = dfrow(c.P.iux), *dy_ = dfrow(c.P.iux + 1);
    for (int i = 0; i < nx; ++i) dx_[i] = dx_[i] + c2 * p.uu[nx + i];
    for (int i = 0; i < nx; ++i) dy_[i] = dy_[i] - c2 * p.uu[i];
  }
  // calc_viscous_force viscosity.f90:2791-2825: df_u += fvisc
  for (int j = 0; j < 3; ++j) {
    double *d = dfrow(c.P.iux + j);
    for (int i = 0; i < nx; ++i) d[i] = d[i] + p.fvisc[j * nx + i];
  }
  // dlnrho_dt density.f90:2887-3060: density_rhs = -divu - uglnrho ; df += density_rhs
  {
    double *d = dfrow(c.P.ilnrho);
    for (int i = 0; i < nx; ++i) {
      double density_rhs = -p.divu[i];
      density_rhs = density_rhs - p.uglnrho[i];
      d[i] = d[i] + density_rhs;
    }
    // hyper_diffusion density.f90:3197-3288
    if (c.P.ldiff_hyper3lnrho) {
      for (int i = 0; i < nx; ++i) d[i] = d[i] + c.P.diffrho_hyper3 * p.del6lnrho[i];
      if (lcourant) {
        // :3377-3390 diffus_diffrho3*dxyz_6
        for (int i = 0; i < nx; ++i) cfl.maxdiffus3[i] = std::max(cfl.maxdiffus3[i], c.P.diffrho_hyper3 * g.dxyz_6);
      }
    }
  }
  // denergy_dt entropy.f90:3671-3876 / noentropy.f90:364-402: df_u += fpres
  for (int j = 0; j < 3; ++j) {
    double *d = dfrow(c.P.iux + j);
    for (int i = 0; i < nx; ++i) d[i] = d[i] + p.fpres[j * nx + i];
  }
  if (c.lentropy()) {
    double *d = dfrow(c.P.iss);
    for (int i = 0; i < nx; ++i) d[i] = d[i] - p.ugss[i];
    // calc_viscous_heat viscosity.f90:2743-2789: df_ss += TT1*visc_heat
    if (c.P.lviscosity_heat)
      for (int i = 0; i < nx; ++i) d[i] = d[i] + p.TT1[i] * p.visc_heat[i];
    std::fill(p.diffus_chi.begin(), p.diffus_chi.end(), 0.0); // entropy.f90:3789
    // calc_heatcond (K-profile) entropy.f90:5913-6151: thdiff = rho1*hcond*(del2lnTT + glnTT.(glnTT+glhc))
    if (c.P.lheatc_Kprof) {
      const int nz_ = D.n - NG;
      const double hcond = c.hcond_prof[nz_], glhc3 = c.dlnhcond_prof[nz_];
      V glnThcond((size_t)3 * nx), g2(nx);
      for (int k = 0; k < 3 * nx; ++k) glnThcond[k] = p.glnTT[k] + (k >= 2 * nx ? glhc3 : 0.);
      dot_mn(nx, p.glnTT.data(), glnThcond.data(), g2.data());
      for (int i = 0; i < nx; ++i) {
        const double thdiff = p.rho1[i] * hcond * (p.del2lnTT[i] + g2[i]);
        const double chix = p.rho1[i] * hcond * c.eos.cv1;
        d[i] = d[i] + thdiff;
        if (lcourant) p.diffus_chi[i] = p.diffus_chi[i] + chix * g.dxyz_2;
      }
    }
    // calc_heatcond_constchi entropy.f90:4852-4947
    if (c.P.lheatc_chiconst) {
      V gsum((size_t)3 * nx), g2(nx);
      for (int k = 0; k < 3 * nx; ++k) gsum[k] = p.glnrho[k] + p.glnTT[k];
      dot_mn(nx, gsum.data(), p.glnTT.data(), g2.data());
      const double cp = 1 / c.eos.cp1; // p%cp=1/cp1 eos_idealgas.f90:811
      for (int i = 0; i < nx; ++i) d[i] = d[i] + c.P.chi * cp * (p.del2lnTT[i] + g2[i]);
      if (lcourant)
        for (int i = 0; i < nx; ++i) p.diffus_chi[i] = p.diffus_chi[i] + (c.eos.gamma * c.P.chi) * g.dxyz_2;
    }
    if (lcourant) // entropy.f90:3826-3829
      for (int i = 0; i < nx; ++i) cfl.maxdiffus[i] = std::max(cfl.maxdiffus[i], p.diffus_chi[i]);
    // calc_heat_cool entropy.f90:6387-6598 with get_heat_cool_gravz :6761-6828 (cooling_profile='gaussian',
    // lcooling_to_cs2cool=T, luminosity=0): heat = -profz_cool(n)*(cs2-cs2cool)/cs2cool ; df_ss += TT1*rho1*heat
    if (c.P.cool != 0.0) {
      const double pc = c.profz_cool[D.n - NG];
      for (int i = 0; i < nx; ++i) {
        double heat = 0.0;
        heat = heat - pc * (p.cs2[i] - c.P.cs2cool) / c.P.cs2cool;
        d[i] = d[i] + p.TT1[i] * p.rho1[i] * heat;
      }
    }
  }
  if (lcourant)
    for (int i = 0; i < nx; ++i) cfl.advec2[i] = cfl.advec2[i]; // advec_cs2 is added in calc_maxadvec
  // daa_dt magnetic.f90:5392-6640
  if (c.lmagnetic()) {
    // Lorentz force :5531
    if (c.P.llorentzforce)
      for (int j = 0; j < 3; ++j) {
        double *d = dfrow(c.P.iux + j);
        for (int i = 0; i < nx; ++i) d[i] = d[i] + p.jxbr[j * nx + i];
      }
    // fres = eta*del2a (non-Weyl) :5584-5590
    V fres((size_t)3 * nx, 0.0), dAdt((size_t)3 * nx, 0.0);
    if (c.P.lresi_eta_const)
      for (int k = 0; k < 3 * nx; ++k) fres[k] = fres[k] + c.P.eta * p.del2a[k];
    // ohmic heat :6138-6144
    if (c.lentropy() && c.P.lohmic_heat) {
      double *d = dfrow(c.P.iss);
      for (int i = 0; i < nx; ++i) d[i] = d[i] + c.P.eta * c.P.mu0 * p.j2[i] * p.rho1[i] * p.TT1[i];
    }
    if (!c.P.lupw_aa) {
      for (int k = 0; k < 3 * nx; ++k) dAdt[k] = dAdt[k] + p.uxb[k] + fres[k]; // :6247
    } else {
      // :6257-6313
      V uxb_upw((size_t)3 * nx, 0.0);
      for (int j = 0; j < 3; ++j) {
        for (int k = 0; k < 3; ++k)
          if (k != j)
            for (int i = 0; i < nx; ++i)
              uxb_upw[j * nx + i] = uxb_upw[j * nx + i] +
                                    p.uu[k * nx + i] * (p.aij[(k * 3 + j) * nx + i] - p.aij[(j * 3 + k) * nx + i]);
        doupwind(D, c.P.iax + j, p.uu.data(), uxb_upw.data() + j * nx, j + 1, nx);
      }
      for (int k = 0; k < 3 * nx; ++k) dAdt[k] = dAdt[k] + uxb_upw[k] + fres[k];
    }
    for (int j = 0; j < 3; ++j) { // :6524
      double *d = dfrow(c.P.iax + j);
      for (int i = 0; i < nx; ++i) d[i] = d[i] + dAdt[j * nx + i];
    }
    if (lcourant) // :6541-6573 diffus_eta=eta*dxyz_2
      for (int i = 0; i < nx; ++i) cfl.maxdiffus[i] = std::max(cfl.maxdiffus[i], c.P.eta * g.dxyz_2);
  }
  // addgravity gravity_simple.f90:1083-1162 (equ.f90:1316): df_uz += gravz_zpencil(n), profile 'const'
  if (c.P.gravz != 0.0) {
    double *d = dfrow(c.P.iux + 2);
    for (int i = 0; i < nx; ++i) d[i] = d[i] + c.P.gravz;
  }
}

struct Accum { // per-thread diagnostic accumulators
  double dt1 = 0, u2s = 0, u2m = 0, rhos = 0, b2s = 0, b2m = 0, j2s = 0, j2m = 0, abs_ = 0, jbs = 0, ous = 0,
         epsKs = 0, sss = 0, dtu = 0, dtc = 0, dtb = 0, dtnu = 0, dtchi = 0;
};

} // namespace

extern "C" {

int pc_oracle_sizeof_pars(void) { return (int)sizeof(pc_oracle_pars); }

// bc_per_x/y/z boundcond.f90:2387-2433 (nproc=1): f(1:l1-1)=f(l2i:l2); f(l2+1:mx)=f(l1:l1i)
void pc_oracle_update_ghosts(const pc_oracle_pars *P, double *f) {
  Grid g(*P);
  for (int k = 0; k < P->nvar; ++k) {
    // x
    for (int iz = NG; iz < NG + g.nz; ++iz)
      for (int iy = NG; iy < NG + g.ny; ++iy)
        for (int q = 0; q < NG; ++q) {
          f[g.idx(q, iy, iz, k)] = f[g.idx(g.nx + q, iy, iz, k)];
          f[g.idx(NG + g.nx + q, iy, iz, k)] = f[g.idx(NG + q, iy, iz, k)];
        }
    // y (full x range, as boundconds_y operates on f(:,...))
    for (int iz = NG; iz < NG + g.nz; ++iz)
      for (int q = 0; q < NG; ++q)
        for (int ix = 0; ix < g.mx; ++ix) {
          f[g.idx(ix, q, iz, k)] = f[g.idx(ix, g.ny + q, iz, k)];
          f[g.idx(ix, NG + g.ny + q, iz, k)] = f[g.idx(ix, NG + q, iz, k)];
        }
    // z
    if (P->bcz[0][k] == PC_BC_P)
      for (int q = 0; q < NG; ++q)
        for (int iy = 0; iy < g.my; ++iy)
          for (int ix = 0; ix < g.mx; ++ix) {
            f[g.idx(ix, iy, q, k)] = f[g.idx(ix, iy, g.nz + q, k)];
            f[g.idx(ix, iy, NG + g.nz + q, k)] = f[g.idx(ix, iy, NG + q, k)];
          }
  }
  // non-periodic z: boundconds_z boundcond.f90:1421-1787 -- bottom for all variables in slot order, then top; every
  // condition acts on whole (mx,my) planes, x/y ghosts included
  Eos e(*P);
  const int n1 = NG, n2 = NG + g.nz - 1;
  for (int tb = 0; tb < 2; ++tb)
    for (int k = 0; k < P->nvar; ++k) {
      const int bc = P->bcz[tb][k];
      if (bc == PC_BC_P) continue;
      const int nb = tb == 0 ? n1 : n2, dir = tb == 0 ? +1 : -1; // ghost plane nb - dir*i mirrors nb + dir*i
      for (int iy = 0; iy < g.my; ++iy)
        for (int ix = 0; ix < g.mx; ++ix) {
          auto F = [&](int iz, int kk) -> double & { return f[g.idx(ix, iy, iz, kk)]; };
          switch (bc) {
          case PC_BC_S: // bc_sym_z sgn=+1 :3842-3900
            for (int i = 1; i <= NG; ++i) F(nb - dir * i, k) = F(nb + dir * i, k);
            break;
          case PC_BC_A: // sgn=-1: antisymmetry, boundary value 0
            for (int i = 1; i <= NG; ++i) F(nb - dir * i, k) = -F(nb + dir * i, k);
            F(nb, k) = 0.;
            break;
          case PC_BC_A2: // sgn=-1, rel=T: bottom 2*f(n1)+sgn*f(n1+i) ; top f(n2)+(f(n2)+sgn*f(n2-i))
            for (int i = 1; i <= NG; ++i) {
              if (tb == 0) F(nb - i, k) = 2 * F(nb, k) - F(nb + i, k);
              else F(nb + i, k) = F(nb, k) + (F(nb, k) - F(nb - i, k));
            }
            break;
          case PC_BC_C1: { // bc_ss_flux entropy_bcs.f90:93-235 (not lheatc_chiconst, not Kramers)
            if (k != P->iss || P->lheatc_chiconst) { std::fprintf(stderr, "pc_oracle: bcz 'c1' only for ss with K-profile conduction\n"); std::exit(1); }
            const int il = P->ilnrho;
            double cs2_xy = F(nb, k);
            cs2_xy = e.cs20 * std::exp(e.gamma_m1 * (F(nb, il) - e.lnrho0) + e.cv1 * cs2_xy);
            const double tmp_xy = e.cp * (tb == 0 ? P->FbotKbot : P->FtopKtop) / cs2_xy;
            for (int i = 1; i <= NG; ++i) {
              const double dz2_bound = 2. * (i * P->dz); // grid.f90:3149-3151
              if (tb == 0) {
                const double dlnrho = F(nb + i, il) - F(nb - i, il); // getdlnrho_z density_methods.f90:409-426
                F(nb - i, k) = F(nb + i, k) + (e.cp - e.cv) * (dlnrho + dz2_bound * tmp_xy);
              } else {
                const double dlnrho = F(nb + i, il) - F(nb - i, il);
                F(nb + i, k) = F(nb - i, k) + (e.cp - e.cv) * (-dlnrho - dz2_bound * tmp_xy);
              }
            }
          } break;
          case PC_BC_CT: { // bc_ss_temp_z eos_idealgas.f90:3112-3260 (entropy, logarithmic density)
            if (k != P->iss) { std::fprintf(stderr, "pc_oracle: bcz 'cT' only for ss\n"); std::exit(1); }
            const int il = P->ilnrho;
            const double tmp = 2 * e.cv * std::log((tb == 0 ? P->cs2bot : P->cs2top) / e.cs20);
            F(nb, k) = 0.5 * tmp - (e.cp - e.cv) * (F(nb, il) - e.lnrho0);
            for (int i = 1; i <= NG; ++i)
              F(nb - dir * i, k) = -F(nb + dir * i, k) + tmp - (e.cp - e.cv) * (F(nb + dir * i, il) + F(nb - dir * i, il) - 2 * e.lnrho0);
          } break;
          default: std::fprintf(stderr, "pc_oracle: unknown bcz %d\n", bc); std::exit(1);
          }
        }
    }
}

void pc_oracle_deriv(const pc_oracle_pars *P, const double *f, int k, int j, int kind, double *out) {
  Grid g(*P);
  Deriv D(g, f);
  for (int n = 0; n < g.nz; ++n)
    for (int m = 0; m < g.ny; ++m) {
      D.m = m + NG; D.n = n + NG;
      double *o = out + ((size_t)n * g.ny + m) * g.nx;
      switch (kind) {
      case 1: D.der(k, j, o); break;
      case 2: D.der2(k, j, o); break;
      case 5: D.der5(k, j, o); break;
      case 6: D.der6(k, j, o, false); break;
      case 7: D.der6(k, j, o, true); break;
      case 12: D.derij(k, 1, 2, o); break;
      case 23: D.derij(k, 2, 3, o); break;
      case 31: D.derij(k, 3, 1, o); break;
      default: std::fprintf(stderr, "pc_oracle_deriv: bad kind %d\n", kind); std::exit(1);
      }
    }
}

// rhs_cpu equ.f90:1386-1567 (mn loop), set_dt1_max :1872-1972
double pc_oracle_rhs(const pc_oracle_pars *P, const double *f, double *df, int lcourant_, pc_oracle_diag *diag) {
  Ctx c(*P);
  const Grid &g = c.g;
  const bool lcourant = lcourant_ != 0;
  const int nyz = g.ny * g.nz;
  int nth = P->nthreads;
#ifdef _OPENMP
  if (nth <= 0) nth = omp_get_max_threads();
#else
  nth = 1;
#endif
  std::vector<Accum> acc(nth);
  const double dt1_init = (P->dtmax != 0.0) ? 1.0 / P->dtmax : 0.0; // equ.f90:270-280
#pragma omp parallel num_threads(nth)
  {
#ifdef _OPENMP
    const int tid = omp_get_thread_num();
#else
    const int tid = 0;
#endif
    Pencil p(g.nx);
    Cfl cfl(g.nx);
    Deriv D(g, f);
    Accum a;
    a.dt1 = dt1_init;
#pragma omp for schedule(static)
    for (int imn = 0; imn < nyz; ++imn) {
      D.n = imn / g.ny + NG;
      D.m = imn % g.ny + NG;
      cfl.zero();
      // calc_all_pencils equ.f90:887-993
      calc_pencils_hydro(c, D, p, cfl, lcourant);
      calc_pencils_density(c, D, p);
      calc_pencils_eos(c, D, p);
      calc_pencils_energy(c, D, p);
      calc_pencils_viscosity(c, p, cfl, lcourant);
      if (c.lmagnetic()) calc_pencils_magnetic(c, D, p, cfl, lcourant);
      // calc_all_rhs equ.f90:1223-1366
      calc_all_rhs(c, D, p, cfl, df, lcourant);
      if (lcourant) {
        // calc_maxadvec + set_dt1_max equ.f90:1858-1972
        for (int i = 0; i < g.nx; ++i) {
          cfl.advec2[i] = cfl.advec2[i] + p.advec_cs2[i];
          cfl.maxadvec[i] = cfl.maxadvec[i] + std::sqrt(cfl.advec2[i]);
          double dt1_advec = cfl.maxadvec[i] / P->cdt;
          double dt1_diffus = cfl.maxdiffus[i] / P->cdtv + cfl.maxdiffus2[i] / P->cdtv2 + cfl.maxdiffus3[i] / P->cdtv3;
          double dt1_src = 0.0;
          double loc = std::sqrt(dt1_advec * dt1_advec + dt1_diffus * dt1_diffus + dt1_src * dt1_src);
          a.dt1 = std::max(a.dt1, loc);
        }
      }
      if (diag) {
        for (int i = 0; i < g.nx; ++i) {
          a.u2s += p.u2[i]; a.u2m = std::max(a.u2m, p.u2[i]);
          a.rhos += p.rho[i];
          a.ous += p.ou[i];
          a.epsKs += p.visc_heat[i] * p.rho[i]; // viscosity.f90: epsK=<visc_heat*rho>
          if (c.lentropy()) a.sss += p.ss[i];
          if (c.lmagnetic()) {
            a.b2s += p.b2[i]; a.b2m = std::max(a.b2m, p.b2[i]);
            a.j2s += p.j2[i]; a.j2m = std::max(a.j2m, p.j2[i]);
            a.abs_ += p.ab[i]; a.jbs += p.jb[i];
          }
          if (lcourant) {
            a.dtu = std::max(a.dtu, p.advec_uu[i] / P->cdt);
            a.dtc = std::max(a.dtc, std::sqrt(p.advec_cs2[i]) / P->cdt);
            if (c.lmagnetic()) a.dtb = std::max(a.dtb, std::sqrt(p.advec_va2[i]) / P->cdt);
            a.dtnu = std::max(a.dtnu, p.diffus_nu[i] / P->cdtv);
            if (c.lentropy()) a.dtchi = std::max(a.dtchi, p.diffus_chi[i] / P->cdtv);
          }
        }
      }
    }
    acc[tid] = a;
  }
  Accum t;
  t.dt1 = dt1_init;
  for (const Accum &a : acc) {
    t.dt1 = std::max(t.dt1, a.dt1);
    t.u2s += a.u2s; t.u2m = std::max(t.u2m, a.u2m); t.rhos += a.rhos; t.b2s += a.b2s;
    t.b2m = std::max(t.b2m, a.b2m); t.j2s += a.j2s; t.j2m = std::max(t.j2m, a.j2m);
    t.abs_ += a.abs_; t.jbs += a.jbs; t.ous += a.ous; t.epsKs += a.epsKs; t.sss += a.sss;
    t.dtu = std::max(t.dtu, a.dtu); t.dtc = std::max(t.dtc, a.dtc); t.dtb = std::max(t.dtb, a.dtb);
    t.dtnu = std::max(t.dtnu, a.dtnu); t.dtchi = std::max(t.dtchi, a.dtchi);
  }
  if (diag) {
    const double nw = (double)g.nx * g.ny * g.nz; // nwgrid, diagnostics.f90:840-1051
    diag->urms = std::sqrt(t.u2s / nw); diag->umax = std::sqrt(t.u2m);
    diag->rhom = t.rhos / nw;
    diag->brms = std::sqrt(t.b2s / nw); diag->bmax = std::sqrt(t.b2m);
    diag->jrms = std::sqrt(t.j2s / nw); diag->jmax = std::sqrt(t.j2m);
    diag->abm = t.abs_ / nw; diag->jbm = t.jbs / nw; diag->oum = t.ous / nw;
    diag->epsK = t.epsKs / nw; diag->ssm = t.sss / nw;
    diag->dtu = t.dtu; diag->dtc = t.dtc; diag->dtb = t.dtb; diag->dtnu = t.dtnu;
    diag->dt1_max = t.dt1; diag->dtchi = t.dtchi;
  }
  return lcourant ? t.dt1 : 0.0;
}

// advance_substeps timestep.f90:77-204 ; coefficients initialize_timestep :39-44
void pc_oracle_substep(const pc_oracle_pars *P, double *f, double *df, int itsub, int ldt, double *dt, double *t,
                       pc_oracle_diag *diag) {
  static const double alpha_ts[3] = {0.0, -5.0 / 9.0, -153.0 / 128.0};
  static const double beta_ts[3] = {1.0 / 3.0, 15.0 / 16.0, 8.0 / 15.0};
  Grid g(*P);
  const size_t ntot = g.sv * P->nvar;
  // ds follows the same recurrence as df (:109-116,146): ds=alpha*ds+1 from ds=0
  double ds = 0.0;
  for (int s = 1; s < itsub; ++s) ds = alpha_ts[s - 1] * ds + 1.0;
  if (itsub == 1) {
    std::memset(df, 0, ntot * sizeof(double)); // :109-111 df=0
    ds = 0.0;
  } else {
    const double a = alpha_ts[itsub - 1];
#pragma omp parallel for schedule(static)
    for (size_t q = 0; q < ntot; ++q) df[q] = a * df[q]; // :113
    ds = a * ds;
  }
  pc_oracle_update_ghosts(P, f); // pde: boundconds_x/y/z equ.f90:208-227
  const bool lcourant = (itsub == 1) && ldt; // lupdate_courant_dt equ.f90:106
  double dt1 = pc_oracle_rhs(P, f, df, lcourant ? 1 : 0, itsub == 1 ? diag : nullptr);
  if (lcourant) *dt = 1.0 / dt1; // set_dt sub.f90:4223-4254
  ds = ds + 1.0;                  // :146
  const double dt_beta = (*dt) * beta_ts[itsub - 1]; // :150
  const double dtsub = ds * dt_beta;                 // :152
  // :160-161 f(l1:l2,m1:m2,n1:n2,1:mvar) += dt_beta_ts(itsub)*df(...)
  for (int k = 0; k < P->nvar; ++k)
#pragma omp parallel for collapse(2) schedule(static)
    for (int iz = NG; iz < NG + g.nz; ++iz)
      for (int iy = NG; iy < NG + g.ny; ++iy) {
        size_t o = g.idx(NG, iy, iz, k);
        for (int i = 0; i < g.nx; ++i) f[o + i] = f[o + i] + dt_beta * df[o + i];
      }
  *t = *t + dtsub; // :201
}

void pc_oracle_time_step(const pc_oracle_pars *P, double *f, double *df, int ldt, double *dt, double *t,
                         const pc_oracle_forcing *frc, pc_oracle_diag *diag) {
  for (int itsub = 1; itsub <= 3; ++itsub) pc_oracle_substep(P, f, df, itsub, ldt, dt, t, diag);
  if (frc) pc_oracle_forcing_hel(P, f, frc); // run.f90:575 addforce
}

// forcing_hel forcing.f90:2066-2345 with fxyz_coefs_hel :2003-2064
void pc_oracle_forcing_hel(const pc_oracle_pars *P, double *f, const pc_oracle_forcing *F) {
  Grid g(*P);
  struct Cx { double re, im; };
  auto cmul = [](Cx a, Cx b) { return Cx{a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; };
  std::vector<Cx> fx(g.nx), fy(g.ny), fz(g.nz);
  for (int i = 0; i < g.nx; ++i) {
    double x = P->x0 + i * P->dx;
    double arg = F->kk[0] * F->k1_ff * x + F->phase;
    fx[i] = Cx{std::cos(arg) * F->fact, std::sin(arg) * F->fact};
  }
  for (int i = 0; i < g.ny; ++i) {
    double y = P->y0 + i * P->dy;
    double arg = F->kk[1] * F->k1_ff * y;
    fy[i] = Cx{std::cos(arg), std::sin(arg)};
  }
  for (int i = 0; i < g.nz; ++i) {
    double z = P->z0 + i * P->dz;
    double arg = F->kk[2] * F->k1_ff * z;
    fz[i] = Cx{std::cos(arg), std::sin(arg)};
  }
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < g.nz; ++n)
    for (int m = 0; m < g.ny; ++m)
      for (int j = 0; j < 3; ++j) {
        double *r = f + g.idx(NG, m + NG, n + NG, P->iux + j);
        for (int i = 0; i < g.nx; ++i) {
          Cx fxyz = cmul(cmul(fx[i], fy[m]), fz[n]);
          // real(cmplx(coef1(j),coef2(j))*fxyz)
          double re = F->coef1[j] * fxyz.re - F->coef2[j] * fxyz.im;
          double rhs = 1.0 * F->fda[j] * 1.0 * re; // force_ampl*fda(j)*cos(omega_ff*t), omega_ff=0
          r[i] = r[i] + rhs * 1.0;                 // *force1_scl
        }
      }
}

// fconst_coefs_hel forcing.f90:1773-2001 (default branch: no shear, old_forcing_evector=F)
void pc_oracle_forcing_coefs(double force, double relhel, double kav, double cs0eff, double dt, const double kvec[3],
                             double phase, double phi, pc_oracle_forcing *o) {
  const double kx = kvec[0], ky = kvec[1], kz = kvec[2];
  const double k2 = kx * kx + ky * ky + kz * kz;
  const double k = std::sqrt(k2);
  double ex, ey, ez;
  if ((ky == 0) && (kz == 0)) { ex = 0; ey = 1; ez = 0; } else { ex = 1; ey = 0; ez = 0; }
  double kk[3] = {kx, ky, kz}, ee[3] = {ex, ey, ez}, e1[3], e2[3];
  auto cross = [](const double *a, const double *b, double *c) {
    c[0] = a[1] * b[2] - a[2] * b[1]; c[1] = a[2] * b[0] - a[0] * b[2]; c[2] = a[0] * b[1] - a[1] * b[0];
  };
  cross(kk, ee, e1);
  double norm = e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2];
  for (double &v : e1) v = v / std::sqrt(norm);
  cross(kk, e1, e2);
  norm = e2[0] * e2[0] + e2[1] * e2[1] + e2[2] * e2[2];
  for (double &v : e2) v = v / std::sqrt(norm);
  for (int q = 0; q < 3; ++q) ee[q] = std::cos(phi) * e1[q] + std::sin(phi) * e2[q];
  ex = ee[0]; ey = ee[1]; ez = ee[2];
  const double kde = kk[0] * ee[0] + kk[1] * ee[1] + kk[2] * ee[2];
  const double kex = ky * ez - kz * ey, key = kz * ex - kx * ez, kez = kx * ey - ky * ex;
  const double kkex = ky * kez - kz * key, kkey = kz * kex - kx * kez, kkez = kx * key - ky * kex;
  const double slope_ff = 0.0;
  const double ffnorm = std::sqrt(1. + relhel * relhel) * k * std::sqrt(k2 - kde * kde) /
                        std::sqrt(kav * cs0eff * cs0eff * cs0eff) * std::pow(k / kav, slope_ff);
  o->fact = force / ffnorm * std::sqrt(dt);
  o->coef1[0] = k * kex; o->coef1[1] = k * key; o->coef1[2] = k * kez;
  o->coef2[0] = relhel * kkex; o->coef2[1] = relhel * kkey; o->coef2[2] = relhel * kkez;
  o->fda[0] = o->fda[1] = o->fda[2] = 1.;
  o->kk[0] = kx; o->kk[1] = ky; o->kk[2] = kz;
  o->phase = phase;
  o->k1_ff = 1.0;
}

} // extern "C"
