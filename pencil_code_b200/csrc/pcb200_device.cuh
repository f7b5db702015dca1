// pcb200_device.cuh -- device-side parameters, 6th-order stencils and the per-cell MHD
// right-hand side for the fused rhs+RK kernels (sm_100a).
//
// What is computed follows Pencil Code's Cartesian/equidistant/periodic branches
// (citations relative to /root/reference/src):
//   stencils  deriv.f90:89-171 (der), :472-561 (der2), :1400-1452 (derij, bidiagonal),
//             :936-1072 (der6), :855-935 (der5)
//   pencils   hydro.f90:3925-4063, density.f90:2727-2811, eos_idealgas.f90:851-888/:969-1000,
//             entropy.f90:3554-3653, noentropy.f90:252-325, viscosity.f90:1633-1664,
//             :1982-1989,:2250-2257, magnetic.f90:4279-5225
//   rhs       hydro.f90:4332-4684, density.f90:2887-3060,3197-3288, entropy.f90:3671-3876,
//             :4852-4947, magnetic.f90:5392-6640
//   cfl       equ.f90:1858-1972
// HOW it is computed is per-cell (not per-pencil): only the quantities the RHS needs are
// formed (6 of 9 aij, 6 of 9 mixed d2A), stencil differences are shared, FMA contraction on.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pcb200 {

constexpr int NG = 3;
constexpr int MAXVAR = 8;

struct Pars {
  int nx, ny, nz, mx, my, mz, nvar;
  long long sy, sz, sv; // strides (doubles) of y, z, variable
  double d1fac[3];      // (1/60)*d_j^-1            der
  double d2fac[3];      // d_j^-2                   der2
  double dmfac[3];      // (1/720)/(d_a d_b): xy, yz, zx   derij
  double d6fac[3];      // d_j^-6                   der6
  double d6upfac[3];    // (1/60)*d_j^-1            der6 upwind
  double d5fac[3];      // (1/2)/d_j^5              der5
  // metric factors folded into the stencil weights (variants 2 and 3: one multiply less per derivative)
  double d1c[3][3];     // d1fac * (45, -9, 1)
  double d2c[3][4];     // d2fac * (-490, 270, -27, 2)/180
  double dmc[3][3];     // dmfac * (270, -27, 2)
  double mcs20g1, chicp, nu2, etamu0, nu13; // -cs20/gamma, chi*cp, 2 nu, eta*mu0, nu/3 (variant 9: no loop-invariant products in registers)
  double dmcu[3];       // (nu/3) * dmc[0][:] (variant 9: z-mixed part of grad div u scattered straight into du/dt)
  double d2c0x2, d2c0x3; // 2x / 3x the centre weight of der2 (isotropic grids: second derivatives summed over directions)
  double dline_1[3];
  double dxyz_2, dxyz_4, dxyz_6;
  double cs20, gamma, gamma1, gamma_m1, cp, cp1, cv1, lnrho0, lnTT0, TT0_1; // TT0_1 = exp(-lnTT0)
  double expc[14];      // 1/n!, n = 0..13 (exp_fast)
  double expk[3];       // 32/ln2, -(ln2/32) high part (32 bits), -(ln2/32) low part
  double exptab[64];    // 2^(j/32), j = 0..31, then 2^(-j/32) (variant 9: table-based exp, copied to shared memory)
  double mu0, mu01;
  double nu, nu_hyper3, diffrho_hyper3, chi, eta;
  double cdt, cdtv, cdtv2, cdtv3, dt1_init;
  int lvisc_nu_const, lvisc_hyper3_simplified, lvisc_hyper3_nu_const, ldiff_hyper3lnrho;
  int lheatc_chiconst, lviscosity_heat, lresi_eta_const, lohmic_heat, llorentzforce;
  int lupw_uu, lupw_lnrho, lupw_ss, lupw_aa;
  // vertically stratified set-ups (template flag STRAT of variant 1): constant gravity, Coriolis force (theta=0),
  // K-profile conduction, cooling layer; closed z boundaries
  int lstrat, lheatc_Kprof, lcool;
  int bcz[2][MAXVAR];
  double gravz, c2omega, cs2cool1;      // 2*Omega, 1/cs2cool (cs2cool itself is cs2cool_)
  double cs2cool_;
  const double *zprof;                   // device: hcond(n), dlnhcond/dz(n), profz_cool(n), n = 0..nz-1 (interior planes)
  double cp_cv, cpFbotKbot, cpFtopKtop;  // cp-cv ; cp*Fbot/Kbot, cp*Ftop/Ktop ('c1')
  double cT_tmp[2];                      // 2*cv*log(cs2bot/cs20), 2*cv*log(cs2top/cs20) ('cT')
  double dz;
};

__constant__ Pars cP; // single translation unit

// helical forcing of the step (forcing.f90:2170-2345); tables are complex (re,im)
struct Forcing {
  int on;
  double coef1[3], coef2[3], fda[3];
  const double2 *fx, *fy, *fz; // indexed by local index incl. ghosts
  // variant 9: per-plane coefficients {P_0..2, Q_0..2}(z) with force_i = Re(fx fy) P_i + Im(fx fy) Q_i,
  // P_i = fda_i (coef1_i Re fz - coef2_i Im fz), Q_i = -fda_i (coef1_i Im fz + coef2_i Re fz)
  const double *pq;
};

// ---------------------------------------------------------------------------
// Tap policy 1: straight from global memory (L1/L2 cached).  c points at the
// centre cell of variable 0.
// ---------------------------------------------------------------------------
struct GlobalTaps {
  const double *__restrict__ c;
  __device__ __forceinline__ long long stride(int dir) const {
    return dir == 0 ? 1LL : (dir == 1 ? cP.sy : cP.sz);
  }
  __device__ __forceinline__ double ld(const double *p) const { return __ldg(p); }
  __device__ __forceinline__ double val(int k) const { return ld(c + k * cP.sv); }
  // der_main deriv.f90:89-171
  __device__ __forceinline__ double d1(int k, int dir) const {
    const double *q = c + k * cP.sv;
    const long long s = stride(dir);
    return cP.d1fac[dir] *
           (45.0 * (ld(q + s) - ld(q - s)) - 9.0 * (ld(q + 2 * s) - ld(q - 2 * s)) + (ld(q + 3 * s) - ld(q - 3 * s)));
  }
  // der2_main deriv.f90:472-561
  __device__ __forceinline__ double d2(int k, int dir) const {
    const double *q = c + k * cP.sv;
    const long long s = stride(dir);
    return cP.d2fac[dir] * ((-490. / 180.) * ld(q) + (270. / 180.) * (ld(q + s) + ld(q - s)) +
                            (-27. / 180.) * (ld(q + 2 * s) + ld(q - 2 * s)) + (2. / 180.) * (ld(q + 3 * s) + ld(q - 3 * s)));
  }
  // derij_main (bidiagonal) deriv.f90:1400-1452 ; pair: 0=xy 1=yz 2=zx
  __device__ __forceinline__ double dij(int k, int pair) const {
    const double *q = c + k * cP.sv;
    const long long sa = pair == 0 ? 1LL : (pair == 1 ? cP.sy : 1LL);
    const long long sb = pair == 0 ? cP.sy : cP.sz;
    return cP.dmfac[pair] *
           (270. * (ld(q + sa + sb) - ld(q - sa + sb) + ld(q - sa - sb) - ld(q + sa - sb)) -
            27. * (ld(q + 2 * sa + 2 * sb) - ld(q - 2 * sa + 2 * sb) + ld(q - 2 * sa - 2 * sb) - ld(q + 2 * sa - 2 * sb)) +
            2. * (ld(q + 3 * sa + 3 * sb) - ld(q - 3 * sa + 3 * sb) + ld(q - 3 * sa - 3 * sb) - ld(q + 3 * sa - 3 * sb)));
  }
  // der6_main deriv.f90:936-1072 (without the metric factor)
  __device__ __forceinline__ double d6raw(int k, int dir) const {
    const double *q = c + k * cP.sv;
    const long long s = stride(dir);
    return -20.0 * ld(q) + 15.0 * (ld(q + s) + ld(q - s)) - 6.0 * (ld(q + 2 * s) + ld(q - 2 * s)) +
           (ld(q + 3 * s) + ld(q - 3 * s));
  }
  __device__ __forceinline__ double d6(int k, int dir) const { return cP.d6fac[dir] * d6raw(k, dir); }
  __device__ __forceinline__ double d6up(int k, int dir) const { return cP.d6upfac[dir] * d6raw(k, dir); }
  // der5 deriv.f90:855-935
  __device__ __forceinline__ double d5(int k, int dir) const {
    const double *q = c + k * cP.sv;
    const long long s = stride(dir);
    return cP.d5fac[dir] *
           (5.0 * (ld(q + s) - ld(q - s)) - 4.0 * (ld(q + 2 * s) - ld(q - 2 * s)) + (ld(q + 3 * s) - ld(q - 3 * s)));
  }
};

// variable slots in Pencil Code's registration order (register.f90:135-147)
// SW (kernel variant 10 with entropy): on the DEVICE the slots of lnrho and s are exchanged (u, s, lnrho, A), so that the
// variables the fused kernel finishes three planes late (u, and s with a magnetic field) and the ones it finishes at once
// (lnrho, A; s without a magnetic field) are neighbours and move as ONE TMA box per array and stage each
// (pcb200.cu: State::swap; host <-> device copies permute the two variables, copy_rows)
template <bool ENT, bool SW = false> struct Slots {
  static constexpr int IUX = 0, ILN = (ENT && SW) ? 4 : 3, ISS = SW ? 3 : 4, IAX = ENT ? 5 : 4;
};

__device__ __forceinline__ int pair_of(int i, int j) { // 0=xy 1=yz 2=zx
  return (i + j == 1) ? 0 : ((i + j == 3) ? 1 : 2);
}

// upwind correction sum_j |u_j| * der6(f,j,upwind), j != mask  (sub.f90:7491-7574)
template <class T> __device__ __forceinline__ double upwind_sum(const T &tp, int k, const double u[3], int mask) {
  double s = 0.0;
#pragma unroll
  for (int j = 0; j < 3; ++j)
    if (j != mask) s += fabs(u[j]) * tp.d6up(k, j);
  return s;
}

// ---------------------------------------------------------------------------
// Full right-hand side of all variables at one cell.
// ---------------------------------------------------------------------------
// STRAT (with n = interior z index of the cell): gravity, Coriolis force, K-profile conduction, cooling layer
template <bool ENT, bool MAG, bool HYP, bool UPW, class T, bool STRAT = false>
__device__ __forceinline__ void rhs_cell(const T &tp, double *__restrict__ rhs, int n = 0) {
  typedef Slots<ENT> S;
  const Pars &P = cP;

  // ---- hydro pencils (hydro.f90:3925-4063)
  double u[3], uij[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    u[i] = tp.val(S::IUX + i);
#pragma unroll
    for (int j = 0; j < 3; ++j) uij[i][j] = tp.d1(S::IUX + i, j);
  }
  const double divu = uij[0][0] + uij[1][1] + uij[2][2];
  // traceless_strain sub.f90:8922-8990
  double sij[3][3];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    sij[j][j] = uij[j][j] - (1. / 3.) * divu;
#pragma unroll
    for (int i = j + 1; i < 3; ++i) {
      sij[i][j] = .5 * (uij[i][j] + uij[j][i]);
      sij[j][i] = sij[i][j];
    }
  }
  // u.grad u (sub.f90:3482-3534)
  double ugu[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    ugu[i] = u[0] * uij[i][0] + u[1] * uij[i][1] + u[2] * uij[i][2];
    if (UPW && P.lupw_uu) ugu[i] -= upwind_sum(tp, S::IUX + i, u, -1);
  }
  // del2v_etc sub.f90:2610-2749
  double del2u[3], graddivu[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double dii = tp.d2(S::IUX + i, i);
    double d2s = dii, gd = dii;
#pragma unroll
    for (int j = 0; j < 3; ++j)
      if (j != i) {
        d2s += tp.d2(S::IUX + i, j);
        gd += tp.dij(S::IUX + j, pair_of(i, j));
      }
    del2u[i] = d2s;
    graddivu[i] = gd;
  }

  // ---- density pencils (density.f90:2727-2811)
  const double lnrho = tp.val(S::ILN);
  double gl[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) gl[j] = tp.d1(S::ILN, j);
  double uglnrho = u[0] * gl[0] + u[1] * gl[1] + u[2] * gl[2];
  if (UPW && P.lupw_lnrho) uglnrho -= upwind_sum(tp, S::ILN, u, -1);
  double sgl[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) sgl[i] = sij[i][0] * gl[0] + sij[i][1] * gl[1] + sij[i][2] * gl[2];

  // ---- eos + energy pencils
  double cs2, TT1 = 0.0, rho1 = 0.0, ss = 0.0;
  double gss[3] = {0., 0., 0.}, glnTT[3] = {0., 0., 0.};
  double fpres[3];
  if (ENT) {
    ss = tp.val(S::ISS);
#pragma unroll
    for (int j = 0; j < 3; ++j) gss[j] = tp.d1(S::ISS, j);
    // eos_idealgas.f90:851-888
    const double e = P.cv1 * ss + P.gamma_m1 * (lnrho - P.lnrho0);
    cs2 = P.cs20 * exp(e);
    TT1 = exp(-(P.lnTT0 + e));
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      glnTT[j] = P.gamma_m1 * gl[j] + P.cv1 * gss[j];
      fpres[j] = -cs2 * (gl[j] + glnTT[j]) * P.gamma1; // entropy.f90:3596-3600
    }
  } else {
    cs2 = P.cs20; // eos_idealgas.f90:969-1000
#pragma unroll
    for (int j = 0; j < 3; ++j) fpres[j] = -cs2 * gl[j]; // noentropy.f90:252-325
  }
  if (MAG) rho1 = exp(-lnrho); // density.f90:2750

  // ---- viscosity (viscosity.f90:1633-1664, :1982-1989, :2250-2257)
  double fvisc[3] = {0., 0., 0.};
  double sij2 = 0.0;
  if (P.lvisc_nu_const) {
#pragma unroll
    for (int i = 0; i < 3; ++i) fvisc[i] += P.nu * (del2u[i] + 2. * sgl[i] + (1. / 3.) * graddivu[i]);
    if (ENT) { // multm2_sym_mn sub.f90:1207-1226
      sij2 = sij[0][0] * sij[0][0];
      sij2 += sij[1][1] * sij[1][1];
      sij2 += 2 * (sij[1][0] * sij[1][0]);
      sij2 += sij[2][2] * sij[2][2];
      sij2 += 2 * (sij[2][0] * sij[2][0]);
      sij2 += 2 * (sij[2][1] * sij[2][1]);
    }
  }
  if (HYP && (P.lvisc_hyper3_simplified || P.lvisc_hyper3_nu_const)) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      double h = tp.d6(S::IUX + i, 0) + tp.d6(S::IUX + i, 1) + tp.d6(S::IUX + i, 2); // del6v sub.f90:2834
      if (P.lvisc_hyper3_nu_const) // uij5glnrho
        h += tp.d5(S::IUX + i, 0) * gl[0] + tp.d5(S::IUX + i, 1) * gl[1] + tp.d5(S::IUX + i, 2) * gl[2];
      fvisc[i] += P.nu_hyper3 * h;
    }
  }

  // ---- momentum: df_u = -ugu + fvisc + fpres (+ jxbr)
  double ru[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) ru[i] = -ugu[i];
  if (STRAT) { // coriolis_cartesian hydro.f90:6339-6366 (theta=0)
    ru[0] += P.c2omega * u[1];
    ru[1] -= P.c2omega * u[0];
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) ru[i] = (ru[i] + fvisc[i]) + fpres[i];

  // ---- continuity (density.f90:2887-3060, 3197-3288)
  double rl = (-divu) - uglnrho;
  if (HYP && P.ldiff_hyper3lnrho) rl += P.diffrho_hyper3 * (tp.d6(S::ILN, 0) + tp.d6(S::ILN, 1) + tp.d6(S::ILN, 2));
  rhs[S::ILN] = rl;

  // ---- entropy (entropy.f90:3671-3876)
  double rs = 0.0;
  if (ENT) {
    double ugss = u[0] * gss[0] + u[1] * gss[1] + u[2] * gss[2];
    if (UPW && P.lupw_ss) ugss -= upwind_sum(tp, S::ISS, u, -1);
    rs = -ugss;
    if (P.lviscosity_heat) rs += TT1 * (2 * P.nu * sij2); // viscosity.f90:2743-2789
    const bool kprof = STRAT && P.lheatc_Kprof;
    double del2lnTT = 0.0;
    if (P.lheatc_chiconst || kprof) {
      const double del2lnrho = tp.d2(S::ILN, 0) + tp.d2(S::ILN, 1) + tp.d2(S::ILN, 2);
      const double del2ss = tp.d2(S::ISS, 0) + tp.d2(S::ISS, 1) + tp.d2(S::ISS, 2);
      del2lnTT = P.gamma_m1 * del2lnrho + P.cv1 * del2ss;
    }
    if (kprof) { // calc_heatcond_arrays entropy.f90:5913-6119: rho1*hcond*(del2lnTT + glnTT.(glnTT+glhc))
      const double hcond = __ldg(P.zprof + n), glhc = __ldg(P.zprof + P.nz + n);
      const double g2 = glnTT[0] * glnTT[0] + glnTT[1] * glnTT[1] + glnTT[2] * (glnTT[2] + glhc);
      rs += exp(-lnrho) * hcond * (del2lnTT + g2);
    }
    if (P.lheatc_chiconst) {                             // entropy.f90:4852-4947
      const double g2 = (gl[0] + glnTT[0]) * glnTT[0] + (gl[1] + glnTT[1]) * glnTT[1] + (gl[2] + glnTT[2]) * glnTT[2];
      rs += P.chi * P.cp * (del2lnTT + g2);
    }
    if (STRAT && P.lcool) { // get_heat_cool_gravz entropy.f90:6761-6828 + calc_heat_cool :6553
      const double heat = -__ldg(P.zprof + 2 * P.nz + n) * (cs2 - P.cs2cool_) * P.cs2cool1;
      rs += TT1 * exp(-lnrho) * heat;
    }
  }

  // ---- magnetic (magnetic.f90:4279-5225, 5392-6640)
  if (MAG) {
    // only the 6 off-diagonal aij are needed for bb (all 9 when upwinding)
    double aij[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) aij[i][j] = (i != j || (UPW && P.lupw_aa)) ? tp.d1(S::IAX + i, j) : 0.0;
    double bb[3];
    bb[0] = aij[2][1] - aij[1][2];
    bb[1] = aij[0][2] - aij[2][0];
    bb[2] = aij[1][0] - aij[0][1];
    // second derivatives of A: d2[i][j]=d_jj A_i ; mixed m(i,pair)
    double del2a[3], jj[3];
    {
      // jj = mu01*curl(bij), with bij from gij_etc (sub.f90:2945-3113):
      // curlb_1 = b32-b23 = (d2A(1,2,2)-d2A(2,2,1)) - (d2A(3,3,1)-d2A(1,3,3))
      double dd[3][3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) dd[i][j] = tp.d2(S::IAX + i, j);
#pragma unroll
      for (int i = 0; i < 3; ++i) del2a[i] = dd[i][0] + dd[i][1] + dd[i][2];
      const double a1_xy = tp.dij(S::IAX + 0, 0), a1_zx = tp.dij(S::IAX + 0, 2);
      const double a2_xy = tp.dij(S::IAX + 1, 0), a2_yz = tp.dij(S::IAX + 1, 1);
      const double a3_yz = tp.dij(S::IAX + 2, 1), a3_zx = tp.dij(S::IAX + 2, 2);
      // bij(i,j)=eps_ikl d2A(k,j,l):  b32=d2A(1,2,2)-d2A(2,2,1), b23=d2A(3,3,1)-d2A(1,3,3) ...
      const double b32 = a2_xy - dd[0][1], b23 = dd[0][2] - a3_zx;
      const double b13 = a3_yz - dd[1][2], b31 = dd[1][0] - a1_xy;
      const double b21 = a1_zx - dd[2][0], b12 = dd[2][1] - a2_yz;
      jj[0] = P.mu01 * (b32 - b23);
      jj[1] = P.mu01 * (b13 - b31);
      jj[2] = P.mu01 * (b21 - b12);
    }
    const double j2 = jj[0] * jj[0] + jj[1] * jj[1] + jj[2] * jj[2];
    if (P.llorentzforce) { // jxbr=rho1*(jj x bb) :4870-4888, df_u += jxbr :5531
      ru[0] += rho1 * (jj[1] * bb[2] - jj[2] * bb[1]);
      ru[1] += rho1 * (jj[2] * bb[0] - jj[0] * bb[2]);
      ru[2] += rho1 * (jj[0] * bb[1] - jj[1] * bb[0]);
    }
    if (ENT && P.lohmic_heat) rs += P.eta * P.mu0 * j2 * rho1 * TT1; // :6138-6144
    double fres[3] = {0., 0., 0.};
    if (P.lresi_eta_const) {
#pragma unroll
      for (int i = 0; i < 3; ++i) fres[i] = P.eta * del2a[i]; // :5584-5590
    }
    if (!(UPW && P.lupw_aa)) {
      // dAdt = uxb + fres :6247
      rhs[S::IAX + 0] = (u[1] * bb[2] - u[2] * bb[1]) + fres[0];
      rhs[S::IAX + 1] = (u[2] * bb[0] - u[0] * bb[2]) + fres[1];
      rhs[S::IAX + 2] = (u[0] * bb[1] - u[1] * bb[0]) + fres[2];
    } else {
      // upwinded induction :6257-6313
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        double x = 0.0;
#pragma unroll
        for (int k = 0; k < 3; ++k)
          if (k != j) x += u[k] * (aij[k][j] - aij[j][k]);
        x += upwind_sum(tp, S::IAX + j, u, j);
        rhs[S::IAX + j] = x + fres[j];
      }
    }
  }
  if (ENT) rhs[S::ISS] = rs;
  if (STRAT) ru[2] += P.gravz; // addgravity gravity_simple.f90:1083-1162
#pragma unroll
  for (int i = 0; i < 3; ++i) rhs[S::IUX + i] = ru[i];
}

// ---------------------------------------------------------------------------
// Courant time step per cell (equ.f90:1858-1972, hydro.f90:3492-3531,
// magnetic.f90:5160-5217, viscosity.f90:2433-2459, entropy.f90:4852-4947)
// ---------------------------------------------------------------------------
struct CflParts {
  double dt1, advec_uu, advec_cs2, advec_va2, diffus_chi;
};
// n: interior z index of the cell (K-profile conduction only)
template <bool ENT, bool MAG, class T, bool SW = false> __device__ __forceinline__ CflParts cfl_cell(const T &tp, int n) {
  typedef Slots<ENT, SW> S;
  const Pars &P = cP;
  CflParts r;
  const double u0 = tp.val(S::IUX), u1 = tp.val(S::IUX + 1), u2 = tp.val(S::IUX + 2);
  r.advec_uu = fabs(u0) * P.dline_1[0] + fabs(u1) * P.dline_1[1] + fabs(u2) * P.dline_1[2];
  const double lnrho = tp.val(S::ILN);
  double cs2 = P.cs20;
  if (ENT) cs2 = P.cs20 * exp(P.cv1 * tp.val(S::ISS) + P.gamma_m1 * (lnrho - P.lnrho0));
  r.advec_cs2 = cs2 * P.dxyz_2;
  r.advec_va2 = 0.0;
  if (MAG && P.llorentzforce) {
    const double b0 = tp.d1(S::IAX + 2, 1) - tp.d1(S::IAX + 1, 2);
    const double b1 = tp.d1(S::IAX + 0, 2) - tp.d1(S::IAX + 2, 0);
    const double b2 = tp.d1(S::IAX + 1, 0) - tp.d1(S::IAX + 0, 1);
    const double t0 = b0 * P.dline_1[0], t1 = b1 * P.dline_1[1], t2 = b2 * P.dline_1[2];
    r.advec_va2 = (t0 * t0 + t1 * t1 + t2 * t2) * P.mu01 * exp(-lnrho);
  }
  const double maxadvec = r.advec_uu + sqrt(r.advec_va2 + r.advec_cs2);
  double maxdiffus = 0.0, maxdiffus3 = 0.0;
  if (P.lvisc_nu_const) maxdiffus = fmax(maxdiffus, P.nu * P.dxyz_2);
  if (P.lvisc_hyper3_simplified || P.lvisc_hyper3_nu_const) {
    double d3 = 0.0;
    if (P.lvisc_hyper3_simplified) d3 += P.nu_hyper3;
    if (P.lvisc_hyper3_nu_const) d3 += P.nu_hyper3;
    maxdiffus3 = fmax(maxdiffus3, d3 * P.dxyz_6);
  }
  if (P.ldiff_hyper3lnrho) maxdiffus3 = fmax(maxdiffus3, P.diffrho_hyper3 * P.dxyz_6);
  r.diffus_chi = 0.0; // entropy.f90:3789,3826-3829: the conduction terms add up before the max
  if (ENT && P.lstrat && P.lheatc_Kprof) r.diffus_chi += (exp(-lnrho) * __ldg(P.zprof + n) * P.cv1) * P.dxyz_2;
  if (ENT && P.lheatc_chiconst) r.diffus_chi += (P.gamma * P.chi) * P.dxyz_2;
  maxdiffus = fmax(maxdiffus, r.diffus_chi);
  if (MAG) maxdiffus = fmax(maxdiffus, P.eta * P.dxyz_2);
  const double a = maxadvec / P.cdt;
  const double d = maxdiffus / P.cdtv + maxdiffus3 / P.cdtv3;
  r.dt1 = sqrt(a * a + d * d);
  return r;
}

// ---------------------------------------------------------------------------
// time-series diagnostics per cell
// ---------------------------------------------------------------------------
enum { DS_U2 = 0, DS_RHO, DS_OU, DS_EPSK, DS_SS, DS_B2, DS_J2, DS_AB, DS_JB, DS_NSUM };
enum { DM_U2 = 0, DM_B2, DM_J2, DM_DTU, DM_DTC, DM_DTB, DM_DTNU, DM_DTCHI, DM_NMAX };

template <bool ENT, bool MAG, class T, bool SW = false>
__device__ __forceinline__ void diag_cell(const T &tp, double *__restrict__ sums, double *__restrict__ maxs, int n) {
  typedef Slots<ENT, SW> S;
  const Pars &P = cP;
  double u[3], uij[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    u[i] = tp.val(S::IUX + i);
#pragma unroll
    for (int j = 0; j < 3; ++j) uij[i][j] = tp.d1(S::IUX + i, j);
  }
  const double u2 = u[0] * u[0] + u[1] * u[1] + u[2] * u[2];
  const double divu = uij[0][0] + uij[1][1] + uij[2][2];
  double sij2;
  {
    const double s00 = uij[0][0] - (1. / 3.) * divu, s11 = uij[1][1] - (1. / 3.) * divu, s22 = uij[2][2] - (1. / 3.) * divu;
    const double s10 = .5 * (uij[1][0] + uij[0][1]), s20 = .5 * (uij[2][0] + uij[0][2]), s21 = .5 * (uij[2][1] + uij[1][2]);
    sij2 = s00 * s00 + s11 * s11 + 2 * (s10 * s10) + s22 * s22 + 2 * (s20 * s20) + 2 * (s21 * s21);
  }
  const double oo0 = uij[2][1] - uij[1][2], oo1 = uij[0][2] - uij[2][0], oo2 = uij[1][0] - uij[0][1];
  const double lnrho = tp.val(S::ILN);
  const double rho = exp(lnrho);
  sums[DS_U2] = u2;
  sums[DS_RHO] = rho;
  sums[DS_OU] = oo0 * u[0] + oo1 * u[1] + oo2 * u[2];
  sums[DS_EPSK] = P.lvisc_nu_const ? (2 * P.nu * sij2) * rho : 0.0;
  sums[DS_SS] = ENT ? tp.val(S::ISS) : 0.0;
  sums[DS_B2] = sums[DS_J2] = sums[DS_AB] = sums[DS_JB] = 0.0;
  maxs[DM_U2] = u2;
  maxs[DM_B2] = maxs[DM_J2] = 0.0;
  if (MAG) {
    double a[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) a[i] = tp.val(S::IAX + i);
    const double b0 = tp.d1(S::IAX + 2, 1) - tp.d1(S::IAX + 1, 2);
    const double b1 = tp.d1(S::IAX + 0, 2) - tp.d1(S::IAX + 2, 0);
    const double b2c = tp.d1(S::IAX + 1, 0) - tp.d1(S::IAX + 0, 1);
    double dd[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) dd[i][j] = tp.d2(S::IAX + i, j);
    const double a1_xy = tp.dij(S::IAX + 0, 0), a1_zx = tp.dij(S::IAX + 0, 2);
    const double a2_xy = tp.dij(S::IAX + 1, 0), a2_yz = tp.dij(S::IAX + 1, 1);
    const double a3_yz = tp.dij(S::IAX + 2, 1), a3_zx = tp.dij(S::IAX + 2, 2);
    const double j0 = P.mu01 * ((a2_xy - dd[0][1]) - (dd[0][2] - a3_zx));
    const double j1 = P.mu01 * ((a3_yz - dd[1][2]) - (dd[1][0] - a1_xy));
    const double j2c = P.mu01 * ((a1_zx - dd[2][0]) - (dd[2][1] - a2_yz));
    const double b2 = b0 * b0 + b1 * b1 + b2c * b2c;
    const double jsq = j0 * j0 + j1 * j1 + j2c * j2c;
    sums[DS_B2] = b2;
    sums[DS_J2] = jsq;
    sums[DS_AB] = a[0] * b0 + a[1] * b1 + a[2] * b2c;
    sums[DS_JB] = j0 * b0 + j1 * b1 + j2c * b2c;
    maxs[DM_B2] = b2;
    maxs[DM_J2] = jsq;
  }
  const CflParts c = cfl_cell<ENT, MAG, T, SW>(tp, n);
  maxs[DM_DTCHI] = c.diffus_chi / P.cdtv;
  maxs[DM_DTU] = c.advec_uu / P.cdt;
  maxs[DM_DTC] = sqrt(c.advec_cs2) / P.cdt;
  maxs[DM_DTB] = sqrt(c.advec_va2) / P.cdt;
  maxs[DM_DTNU] = P.lvisc_nu_const ? (P.nu * P.dxyz_2) / P.cdtv : 0.0;
}

// warp/block reductions (warp shuffles; fixed tree order -> bitwise reproducible)
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

} // namespace pcb200
