// pcb200_fast.cuh -- kernel variant 2: (32x8)-cell tiles marched along z, the 7 z-planes of the
// vector fields staged in a shared-memory ring by TMA (cp.async.bulk.tensor), scalar fields' z taps
// in a per-thread register pipeline, one fused RHS + 2N-RK3 update per cell.
//
// Same arithmetic as rhs_cell (pcb200_device.cuh; reference citations there) but organised around
// "lines": the 6 taps of one variable along one axis are loaded once and feed der, der2 (and
// der5/der6/upwind when enabled), deriv.f90:89-171,472-561,855-935,936-1072.
#pragma once
#include "pcb200_device.cuh"
#include <cuda.h>

namespace pcb200 {

// ---- derivatives from a loaded line (c = centre, m[i]/p[i] = taps at -(i+1)/+(i+1)) ------------
__device__ __forceinline__ double d1_line(const double m[3], const double p[3], int dir) {
  return cP.d1c[dir][0] * (p[0] - m[0]) + (cP.d1c[dir][1] * (p[1] - m[1]) + cP.d1c[dir][2] * (p[2] - m[2]));
}
__device__ __forceinline__ double d2_line(double c, const double m[3], const double p[3], int dir) {
  return (cP.d2c[dir][0] * c + cP.d2c[dir][1] * (p[0] + m[0])) + (cP.d2c[dir][2] * (p[1] + m[1]) + cP.d2c[dir][3] * (p[2] + m[2]));
}
__device__ __forceinline__ double d6raw_line(double c, const double m[3], const double p[3]) {
  return -20.0 * c + 15.0 * (p[0] + m[0]) - 6.0 * (p[1] + m[1]) + (p[2] + m[2]);
}
__device__ __forceinline__ double d5_line(const double m[3], const double p[3], int dir) {
  return cP.d5fac[dir] * (5.0 * (p[0] - m[0]) - 4.0 * (p[1] - m[1]) + (p[2] - m[2]));
}

// exp(x) for the equation of state: x = k ln2 + r, |r| <= ln2/2 (Cody-Waite, two-part ln2), degree-13
// Taylor polynomial of exp(r) (truncation 4e-18 relative), coefficients read from constant memory as
// FMA operands; 2^k by an integer add on the exponent field.  ~1 ulp for |x| < 700 (the EOS arguments
// are O(1)); no special-case branches.  Half the instructions of the library exp, whose coefficients
// are materialised by two moves each.
__device__ __forceinline__ double exp_fast(double x) {
  const double shifter = 6755399441055744.0; // 1.5 * 2^52: the low word of x*log2e + shifter is rint(x*log2e)
  const double t = fma(x, 1.4426950408889634074, shifter);
  const int k = __double2loint(t);
  const double kd = t - shifter;
  double r = fma(kd, -6.93147180369123816490e-01, x);
  r = fma(kd, -1.90821492927058770002e-10, r);
  double p = cP.expc[13];
#pragma unroll
  for (int n = 12; n >= 0; --n) p = fma(p, r, cP.expc[n]);
  return __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));
}

// Isotropic grids (one set of der2 weights): the symmetric tap sums p_k + m_k of several directions are added up
// first and weighted once -- sum_j der2_j(f) = n c0 f + c1 S0 + c2 S1 + c3 S2 with S_k = sum_j (p_jk + m_jk) --
// which saves 4 FP64 operations per Laplacian and 2-3 per (diagonal + rest) pair.
struct SymSum {
  double s0, s1, s2;
  bool first = true;
  __device__ __forceinline__ void add(const double m[3], const double p[3]) {
    if (first) { s0 = p[0] + m[0]; s1 = p[1] + m[1]; s2 = p[2] + m[2]; first = false; }
    else { s0 += p[0] + m[0]; s1 += p[1] + m[1]; s2 += p[2] + m[2]; }
  }
  // c0n = n * d2c[0][0] for the n directions that were added
  __device__ __forceinline__ double weighted(double c, double c0n) const {
    return (c0n * c + cP.d2c[0][1] * s0) + (cP.d2c[0][2] * s1 + cP.d2c[0][3] * s2);
  }
};

// Policy adapter: global memory lines (lets variant-2 physics run on the variant-1 access path)
struct GlobalLines {
  static constexpr bool TRAIL = false, ISO = false;
  static __device__ __forceinline__ int dr(int j) { return j; } // index of the per-direction stencil weights
  static __device__ __forceinline__ int pr(int p) { return p; } // index of the per-pair mixed-derivative weights
  const double *__restrict__ c0;
  __device__ __forceinline__ double c(int k) const { return __ldg(c0 + k * cP.sv); }
  __device__ __forceinline__ void line(int k, int dir, double m[3], double p[3]) const {
    const double *q = c0 + k * cP.sv;
    const long long s = dir == 0 ? 1LL : (dir == 1 ? cP.sy : cP.sz);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      m[i] = __ldg(q - (i + 1) * s);
      p[i] = __ldg(q + (i + 1) * s);
    }
  }
  __device__ __forceinline__ double dij(int k, int pair) const {
    GlobalTaps t{c0};
    return t.dij(k, pair);
  }
};

// ---------------------------------------------------------------------------
// physics on lines, in two halves so that the magnetic half can run in its own warps (variant 3):
//   hydro_part : u, lnrho, ss  -> rhs of u (without the Lorentz force), lnrho, ss (without Ohmic heat)
//   mag_part   : A (+ u centre) -> rhs of A, j x b and j^2
//   combine    : ru += rho1*(j x b), rs += eta*mu0*j2*rho1*TT1
// The accumulation order is that of rhs_cell (pcb200_device.cuh).
// ---------------------------------------------------------------------------
struct HydroOut {
  double u[3], lnrho, TT1;
};

// UMIX=false: the mixed derivatives of grad div u are computed elsewhere (variant 3: magnetic role)
template <bool ENT, bool HYP, bool UPW, class T, bool UMIX = true>
__device__ __forceinline__ void hydro_part(T &tp, double *__restrict__ rhs, double *__restrict__ fc, HydroOut &ho) {
  typedef Slots<ENT> S;
  const Pars &P = cP;
  double tg[3] = {0., 0., 0.}; // trailing halves of the z-mixed derivatives of grad div u (TRAIL policies)
  if (T::TRAIL && UMIX) tp.trail_pop(0, tg);
  const bool hyp_visc = HYP && (P.lvisc_hyper3_simplified || P.lvisc_hyper3_nu_const);

  double u[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) { u[i] = tp.c(S::IUX + i); fc[S::IUX + i] = u[i]; ho.u[i] = u[i]; }
  const double au[3] = {fabs(u[0]), fabs(u[1]), fabs(u[2])};

  // ---- lnrho first (its gradient enters sglnrho, hyper3 viscosity and the pressure force)
  const double lnrho = tp.c(S::ILN);
  fc[S::ILN] = lnrho;
  ho.lnrho = lnrho;
  double gl[3], del2lnrho = 0.0, del6lnrho = 0.0, upw_l = 0.0;
  SymSum sl;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    double m[3], p[3];
    tp.line(S::ILN, j, m, p);
    gl[j] = d1_line(m, p, T::dr(j));
    if (ENT && T::ISO) sl.add(m, p);
    else if (ENT) del2lnrho += d2_line(lnrho, m, p, T::dr(j));
    if (HYP || UPW) {
      const double r6 = d6raw_line(lnrho, m, p);
      if (HYP) del6lnrho += P.d6fac[T::dr(j)] * r6;
      if (UPW) upw_l += au[j] * (P.d6upfac[T::dr(j)] * r6);
    }
  }

  if (ENT && T::ISO) del2lnrho = sl.weighted(lnrho, P.d2c0x3);

  // ---- velocity: gradient tensor, laplacian, diagonal second derivatives
  double uij[3][3], del2u[3], graddivu[3], hyp[3] = {0., 0., 0.}, upw_u[3] = {0., 0., 0.};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    double d2s = 0.0;
    SymSum su;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double m[3], p[3];
      tp.line(S::IUX + i, j, m, p);
      uij[i][j] = d1_line(m, p, T::dr(j));
      if (T::TRAIL && UMIX) { // in-plane differences of this plane feed the outputs of planes z+1..z+3
        if (i == 2 && j == 0) tp.trail_add(0, 0, 2, m, p);
        if (i == 2 && j == 1) tp.trail_add(0, 1, 1, m, p);
        if (i == 0 && j == 0) tp.trail_add(0, 2, 2, m, p);
        if (i == 1 && j == 1) tp.trail_add(0, 2, 1, m, p);
      }
      if (T::ISO && j != i) {
        su.add(m, p); // the two off-diagonal second derivatives only enter through their sum
      } else {
        const double d2 = d2_line(u[i], m, p, T::dr(j));
        d2s += d2;
        if (j == i) graddivu[i] = d2; // fjji(i,i)=der2 (sub.f90:2651-2653)
      }
      if (HYP || UPW) {
        const double r6 = d6raw_line(u[i], m, p);
        if (HYP) {
          hyp[i] += P.d6fac[T::dr(j)] * r6;                                           // del6v
          if (P.lvisc_hyper3_nu_const) hyp[i] += d5_line(m, p, T::dr(j)) * gl[j];     // uij5glnrho
        }
        if (UPW) upw_u[i] += au[j] * (P.d6upfac[T::dr(j)] * r6);
      }
    }
    if (T::ISO) d2s += su.weighted(u[i], P.d2c0x2);
    del2u[i] = d2s;
  }
  // mixed derivatives for grad div u: fjji(i,j)=d_i d_j u_j
  if (UMIX && T::TRAIL) {
    tp.lead(0, tg); // + leading halves, read from the planes z+1..z+3 of the ring
    graddivu[0] += tp.dij(S::IUX + 1, 0) + tg[0];
    graddivu[1] += tp.dij(S::IUX + 0, 0) + tg[1];
    graddivu[2] += tg[2];
  } else if (UMIX) {
    const double u1_xy = tp.dij(S::IUX + 0, 0), u1_zx = tp.dij(S::IUX + 0, 2);
    const double u2_xy = tp.dij(S::IUX + 1, 0), u2_yz = tp.dij(S::IUX + 1, 1);
    const double u3_yz = tp.dij(S::IUX + 2, 1), u3_zx = tp.dij(S::IUX + 2, 2);
    graddivu[0] += u2_xy + u3_zx;
    graddivu[1] += u1_xy + u3_yz;
    graddivu[2] += u1_zx + u2_yz;
  }
  const double divu = uij[0][0] + uij[1][1] + uij[2][2];
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
  double ru[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    double ugu = u[0] * uij[i][0] + u[1] * uij[i][1] + u[2] * uij[i][2];
    if (UPW && P.lupw_uu) ugu -= upw_u[i];
    ru[i] = -ugu;
  }
  double uglnrho = u[0] * gl[0] + u[1] * gl[1] + u[2] * gl[2];
  if (UPW && P.lupw_lnrho) uglnrho -= upw_l;

  // ---- viscosity
  double sij2 = 0.0;
  if (P.lvisc_nu_const) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double sgl = sij[i][0] * gl[0] + sij[i][1] * gl[1] + sij[i][2] * gl[2];
      ru[i] += P.nu * (del2u[i] + 2. * sgl + (1. / 3.) * graddivu[i]);
    }
    if (ENT) {
      sij2 = sij[0][0] * sij[0][0] + sij[1][1] * sij[1][1] + sij[2][2] * sij[2][2] +
             2 * (sij[1][0] * sij[1][0] + sij[2][0] * sij[2][0] + sij[2][1] * sij[2][1]);
    }
  }
  if (hyp_visc) {
#pragma unroll
    for (int i = 0; i < 3; ++i) ru[i] += P.nu_hyper3 * hyp[i];
  }

  // ---- continuity
  double rl = (-divu) - uglnrho;
  if (HYP && P.ldiff_hyper3lnrho) rl += P.diffrho_hyper3 * del6lnrho;
  rhs[S::ILN] = rl;

  // ---- eos / energy
  ho.TT1 = 0.0;
  if (ENT) {
    const double ss = tp.c(S::ISS);
    fc[S::ISS] = ss;
    double gss[3], del2ss = 0.0, upw_s = 0.0;
    SymSum sss;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double m[3], p[3];
      tp.line(S::ISS, j, m, p);
      gss[j] = d1_line(m, p, T::dr(j));
      if (T::ISO) sss.add(m, p);
      else del2ss += d2_line(ss, m, p, T::dr(j));
      if (UPW) upw_s += au[j] * (P.d6upfac[T::dr(j)] * d6raw_line(ss, m, p));
    }
    if (T::ISO) del2ss = sss.weighted(ss, P.d2c0x3);
    const double e = P.cv1 * ss + P.gamma_m1 * (lnrho - P.lnrho0);
    const double ex = exp_fast(e);
    const double cs2 = P.cs20 * ex;
    const double TT1 = P.TT0_1 * __drcp_rn(ex); // = exp(-(lnTT0+e)) to 1 ulp, one exp less (eos_idealgas.f90:851-888)
    ho.TT1 = TT1;
    double glnTT[3], g2 = 0.0;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      glnTT[j] = P.gamma_m1 * gl[j] + P.cv1 * gss[j];
      ru[j] += -cs2 * (gl[j] + glnTT[j]) * P.gamma1;
      g2 += (gl[j] + glnTT[j]) * glnTT[j];
    }
    double ugss = u[0] * gss[0] + u[1] * gss[1] + u[2] * gss[2];
    if (UPW && P.lupw_ss) ugss -= upw_s;
    double rs = -ugss;
    if (P.lviscosity_heat) rs += TT1 * (2 * P.nu * sij2);
    if (P.lheatc_chiconst) rs += P.chi * P.cp * ((P.gamma_m1 * del2lnrho + P.cv1 * del2ss) + g2);
    rhs[S::ISS] = rs;
  } else {
#pragma unroll
    for (int j = 0; j < 3; ++j) ru[j] += -P.cs20 * gl[j];
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) rhs[S::IUX + i] = ru[i];
}

struct MagOut {
  double jxb[3], j2;
};

// u[] = velocity at the cell centre; writes rhs[IAX..IAX+2], fc[IAX..IAX+2]
template <bool ENT, bool UPW, class T>
__device__ __forceinline__ void mag_part(T &tp, const double u[3], double *__restrict__ rhs, double *__restrict__ fc,
                                         MagOut &mo) {
  typedef Slots<ENT> S;
  const Pars &P = cP;
  double tg[3] = {0., 0., 0.};
  if (T::TRAIL) tp.trail_pop(1, tg);
  const bool upa = UPW && P.lupw_aa;
  const double au[3] = {fabs(u[0]), fabs(u[1]), fabs(u[2])};
  constexpr bool SUM = T::ISO && T::TRAIL; // diagonal second derivative + summed rest (see SymSum)
  double aij[3][3], dd[3][3], dia[3], rest[3], upw_a[3] = {0., 0., 0.};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double a = tp.c(S::IAX + i);
    fc[S::IAX + i] = a;
    SymSum sa;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double m[3], p[3];
      tp.line(S::IAX + i, j, m, p);
      if (T::TRAIL) {
        if (i == 2 && j == 0) tp.trail_add(1, 0, 2, m, p);
        if (i == 2 && j == 1) tp.trail_add(1, 1, 1, m, p);
        if (i == 0 && j == 0) tp.trail_add(1, 2, 2, m, p);
        if (i == 1 && j == 1) tp.trail_add(1, 2, 1, m, p);
      }
      aij[i][j] = (i != j || upa) ? d1_line(m, p, T::dr(j)) : 0.0;
      if (SUM) {
        if (j == i) dia[i] = d2_line(a, m, p, T::dr(j));
        else sa.add(m, p);
      } else {
        dd[i][j] = d2_line(a, m, p, T::dr(j));
      }
      if (UPW && j != i) upw_a[i] += au[j] * (P.d6upfac[T::dr(j)] * d6raw_line(a, m, p)); // mask=i (magnetic.f90:6301)
    }
    if (SUM) rest[i] = sa.weighted(a, P.d2c0x2);
  }
  const double bb[3] = {aij[2][1] - aij[1][2], aij[0][2] - aij[2][0], aij[1][0] - aij[0][1]};
  double jj[3];
  if (T::TRAIL) { // jj = mu01*(grad div A - del2 A); z-mixed parts = leading (ring) + trailing (accumulators) halves
    tp.lead(1, tg);
    if (SUM) {
      jj[0] = P.mu01 * ((tp.dij(S::IAX + 1, 0) + tg[0]) - rest[0]);
      jj[1] = P.mu01 * ((tp.dij(S::IAX + 0, 0) + tg[1]) - rest[1]);
      jj[2] = P.mu01 * (tg[2] - rest[2]);
    } else {
      jj[0] = P.mu01 * ((tp.dij(S::IAX + 1, 0) + tg[0]) - (dd[0][1] + dd[0][2]));
      jj[1] = P.mu01 * ((tp.dij(S::IAX + 0, 0) + tg[1]) - (dd[1][2] + dd[1][0]));
      jj[2] = P.mu01 * (tg[2] - (dd[2][0] + dd[2][1]));
    }
  } else {
    const double a1_xy = tp.dij(S::IAX + 0, 0), a1_zx = tp.dij(S::IAX + 0, 2);
    const double a2_xy = tp.dij(S::IAX + 1, 0), a2_yz = tp.dij(S::IAX + 1, 1);
    const double a3_yz = tp.dij(S::IAX + 2, 1), a3_zx = tp.dij(S::IAX + 2, 2);
    jj[0] = P.mu01 * ((a2_xy - dd[0][1]) - (dd[0][2] - a3_zx));
    jj[1] = P.mu01 * ((a3_yz - dd[1][2]) - (dd[1][0] - a1_xy));
    jj[2] = P.mu01 * ((a1_zx - dd[2][0]) - (dd[2][1] - a2_yz));
  }
  mo.jxb[0] = jj[1] * bb[2] - jj[2] * bb[1];
  mo.jxb[1] = jj[2] * bb[0] - jj[0] * bb[2];
  mo.jxb[2] = jj[0] * bb[1] - jj[1] * bb[0];
  mo.j2 = jj[0] * jj[0] + jj[1] * jj[1] + jj[2] * jj[2];
  double fres[3] = {0., 0., 0.};
  if (P.lresi_eta_const) {
#pragma unroll
    for (int i = 0; i < 3; ++i) fres[i] = SUM ? P.eta * (dia[i] + rest[i]) : P.eta * (dd[i][0] + dd[i][1] + dd[i][2]);
  }
  if (!upa) {
    rhs[S::IAX + 0] = (u[1] * bb[2] - u[2] * bb[1]) + fres[0];
    rhs[S::IAX + 1] = (u[2] * bb[0] - u[0] * bb[2]) + fres[1];
    rhs[S::IAX + 2] = (u[0] * bb[1] - u[1] * bb[0]) + fres[2];
  } else {
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double x = 0.0;
#pragma unroll
      for (int k = 0; k < 3; ++k)
        if (k != j) x += u[k] * (aij[k][j] - aij[j][k]);
      rhs[S::IAX + j] = (x + upw_a[j]) + fres[j];
    }
  }
}

// Lorentz force and Ohmic heating (magnetic.f90:4870-4888,5531,6138-6144)
template <bool ENT>
__device__ __forceinline__ void combine_mag(const HydroOut &ho, const double jxb[3], double j2, double *__restrict__ rhs) {
  typedef Slots<ENT> S;
  const Pars &P = cP;
  const double rho1 = exp_fast(-ho.lnrho);
  if (P.llorentzforce) {
#pragma unroll
    for (int i = 0; i < 3; ++i) rhs[S::IUX + i] += rho1 * jxb[i];
  }
  if (ENT && P.lohmic_heat) rhs[S::ISS] += P.eta * P.mu0 * j2 * rho1 * ho.TT1;
}

template <bool ENT, bool MAG, bool HYP, bool UPW, class T>
__device__ __forceinline__ void rhs_cell_lines(T &tp, double *__restrict__ rhs, double *__restrict__ fc) {
  HydroOut ho;
  hydro_part<ENT, HYP, UPW>(tp, rhs, fc, ho);
  if (MAG) {
    MagOut mo;
    mag_part<ENT, UPW>(tp, ho.u, rhs, fc, mo);
    combine_mag<ENT>(ho, mo.jxb, mo.j2, rhs);
  }
}

// ---------------------------------------------------------------------------
// variant 2 tiling constants
// ---------------------------------------------------------------------------
constexpr int V2_TX = 32, V2_TY = 8;
constexpr int V2_HX = V2_TX + 2 * NG, V2_HY = V2_TY + 2 * NG; // 38 x 14 halo'd plane tile
constexpr int V2_PLANE = V2_HX * V2_HY;                       // 532 doubles (TMA box)
constexpr int V2_PLANE_PAD = 544;                             // 4352 B = 34 x 128 B (TMA smem alignment)
constexpr int V2_NSLOT = 8;                                   // planes z-3..z+3 in use + 1 in flight
constexpr int V2_THREADS = V2_TX * V2_TY;
constexpr unsigned V2_BOX_BYTES = V2_PLANE * sizeof(double);

template <bool ENT, bool MAG> struct V2Layout {
  static constexpr int NRING = 3 + (MAG ? 3 : 0); // u (and A): need z-mixed taps -> ring of planes
  static constexpr int NSCAL = 1 + (ENT ? 1 : 0); // lnrho (and ss): only in-plane taps from smem
  static constexpr int RING_DOUBLES = V2_NSLOT * NRING * V2_PLANE_PAD;
  static constexpr int SCAL_DOUBLES = 2 * NSCAL * V2_PLANE_PAD;
  static constexpr size_t SMEM_BYTES = (size_t)(RING_DOUBLES + SCAL_DOUBLES) * sizeof(double) + (V2_NSLOT + 2) * 8 + 128;
};

// ---- PTX wrappers (mbarrier + TMA) -----------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void *dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// TMA prefetch of a tile into L2 (no shared-memory destination, no LSU instructions)
__device__ __forceinline__ void tma_prefetch_4d(const CUtensorMap *map, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.prefetch.tensor.4d.L2.global.tile [%0, {%1, %2, %3, %4}];" ::"l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

// named barriers: producers arrive without blocking, only the consumer waits
__device__ __forceinline__ void named_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void named_bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// ---- tap policy on the shared-memory ring ------------------------------------
// All pointers below are derived from the `extern __shared__` array by plain pointer arithmetic, so
// that the compiler keeps them in the shared address space (LDS with immediate offsets, 32-bit
// address registers) -- one base pointer per ring plane, every tap a compile-time offset from it.
// TRAIL_ = true (kernel variant 4): "lead/trail" evaluation of the z-mixed bidiagonal derivatives
//   d_zx f(z) = sum_k w_k [Dx_k f(z+k) - Dx_k f(z-k)],  Dx_k f(z') = f(x+k,y,z') - f(x-k,y,z')   (deriv.f90:1400-1452)
// The leading half is read from the ring planes z+1..z+3; the trailing half is accumulated, one plane at a
// time, from the in-plane differences that the first derivatives of the centre plane form anyway
// (9 accumulators per vector field) -- 48 of 282 LDS.64 per cell less.
template <bool ENT, bool MAG, bool TRAIL_ = false, bool ISO_ = false> struct SmemLines {
  typedef V2Layout<ENT, MAG> L;
  typedef Slots<ENT> S;
  static constexpr bool TRAIL = TRAIL_;
  // ISO_: dx == dy == dz bit for bit -> one set of stencil weights serves all directions (30 -> 10 constants)
  static constexpr bool ISO = ISO_;
  static __device__ __forceinline__ int dr(int j) { return ISO ? 0 : j; }
  static __device__ __forceinline__ int pr(int p) { return ISO ? 0 : p; }
  const double *pl[7];  // planes z-3..z+3 of ring field 0 at this thread's cell
  const double *scal;   // current scalar planes (field 0) at this thread's cell
  double lz[7], sz[7];  // register pipelines of lnrho / ss along z
  double T[MAG ? 2 : 1][3][3]; // [u|A][component of grad div][output plane z+j]: trailing halves (TRAIL only)

  static __device__ __forceinline__ int gfield(int g) { return g == 0 ? 0 : 3 * V2_PLANE_PAD; } // ring offset of v_x
  __device__ __forceinline__ void trail_zero() {
#pragma unroll
    for (int g = 0; g < (MAG ? 2 : 1); ++g)
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int j = 0; j < 3; ++j) T[g][c][j] = 0.0;
  }
  // accumulators of the output plane -> out[], then shift by one plane
  __device__ __forceinline__ void trail_pop(int g, double out[3]) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      out[c] = T[g][c][0];
      T[g][c][0] = T[g][c][1];
      T[g][c][1] = T[g][c][2];
      T[g][c][2] = 0.0;
    }
  }
  // differences p[k]-m[k] of the centre plane contribute -w_k D_k to the outputs of planes z+k
  __device__ __forceinline__ void trail_add(int g, int c, int pair, const double m[3], const double p[3]) {
#pragma unroll
    for (int j = 0; j < 3; ++j) T[g][c][j] -= cP.dmc[pr(pair)][j] * (p[j] - m[j]);
  }
  // leading halves: planes z+1..z+3 at in-plane offsets +-k
  __device__ __forceinline__ void lead(int g, double out[3]) const {
    const int fx = gfield(g), fy = fx + V2_PLANE_PAD, fz = fy + V2_PLANE_PAD;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const double *q = pl[4 + j];
      const int a = j + 1, b = (j + 1) * V2_HX;
      out[0] += cP.dmc[pr(2)][j] * (q[fz + a] - q[fz - a]);
      out[1] += cP.dmc[pr(1)][j] * (q[fz + b] - q[fz - b]);
      out[2] += cP.dmc[pr(2)][j] * (q[fx + a] - q[fx - a]) + cP.dmc[pr(1)][j] * (q[fy + b] - q[fy - b]);
    }
  }
  // first plane of a z chunk: contributions of the planes z-3..z-1 (ring planes 0..2) to the outputs z..z+2
  __device__ __forceinline__ void trail_prologue(int g) {
    const int fx = gfield(g), fy = fx + V2_PLANE_PAD, fz = fy + V2_PLANE_PAD;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const double *q = pl[d];
#pragma unroll
      for (int k = 3 - d; k <= 3; ++k) {
        const int j = d - 3 + k, a = k, b = k * V2_HX;
        T[g][0][j] -= cP.dmc[pr(2)][k - 1] * (q[fz + a] - q[fz - a]);
        T[g][1][j] -= cP.dmc[pr(1)][k - 1] * (q[fz + b] - q[fz - b]);
        T[g][2][j] -= cP.dmc[pr(2)][k - 1] * (q[fx + a] - q[fx - a]) + cP.dmc[pr(1)][k - 1] * (q[fy + b] - q[fy - b]);
      }
    }
  }

  static __device__ __forceinline__ bool is_ring(int k) { return k < 3 || (MAG && k >= S::IAX); }
  static __device__ __forceinline__ int ridx(int k) { return k < 3 ? k : 3 + (k - S::IAX); }
  static __device__ __forceinline__ int sidx(int k) { return k == S::ILN ? 0 : 1; }

  __device__ __forceinline__ double c(int k) const {
    if (is_ring(k)) return pl[3][ridx(k) * V2_PLANE_PAD];
    return k == S::ILN ? lz[3] : sz[3];
  }
  __device__ __forceinline__ void line(int k, int dir, double m[3], double p[3]) const {
    if (is_ring(k)) {
      const int f = ridx(k) * V2_PLANE_PAD;
      if (dir == 2) {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = pl[2 - i][f];
          p[i] = pl[4 + i][f];
        }
      } else {
        const double *q = pl[3] + f;
        const int s = dir == 0 ? 1 : V2_HX;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = q[-(i + 1) * s];
          p[i] = q[(i + 1) * s];
        }
      }
    } else {
      if (dir == 2) {
        const double *r = (k == S::ILN) ? lz : sz;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = r[2 - i];
          p[i] = r[4 + i];
        }
      } else {
        const double *q = scal + sidx(k) * V2_PLANE_PAD;
        const int s = dir == 0 ? 1 : V2_HX;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = q[-(i + 1) * s];
          p[i] = q[(i + 1) * s];
        }
      }
    }
  }
  // bidiagonal mixed derivative (ring variables only)
  __device__ __forceinline__ double dij(int k, int pair) const {
    const int f = ridx(k) * V2_PLANE_PAD;
    double g[3];
    if (pair == 0) {
      const double *q = pl[3] + f;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const int a = (i + 1), b = (i + 1) * V2_HX;
        g[i] = (q[a + b] - q[-a + b]) + (q[-a - b] - q[a - b]);
      }
    } else {
      const int a1 = pair == 1 ? V2_HX : 1; // in-plane axis: y for yz, x for zx
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const double *qp = pl[4 + i] + f, *qm = pl[2 - i] + f;
        const int a = (i + 1) * a1;
        g[i] = (qp[a] - qp[-a]) + (qm[-a] - qm[a]);
      }
    }
    return cP.dmc[pr(pair)][0] * g[0] + (cP.dmc[pr(pair)][1] * g[1] + cP.dmc[pr(pair)][2] * g[2]);
  }
};

struct V2Args {
  const double *fin;
  double *fout;
  double *df;
  const double *dt_dev;
  double alpha, beta;
  int read_df, write_df;
  int z_lo, z_hi, lz; // z range (incl. ghost offset) and chunk length per blockIdx.z
  int df_prefetch;    // 1: TMA-prefetch df tiles into L2 two planes ahead
  Forcing frc;
};

extern __shared__ __align__(1024) double v2_smem[];

template <bool ENT, bool MAG, bool HYP, bool UPW, bool TRAIL = false>
__global__ void __launch_bounds__(V2_THREADS, 1)
    rhs_update_v2(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_df, const V2Args a) {
  typedef V2Layout<ENT, MAG> L;
  typedef Slots<ENT> S;
  constexpr int NV = 4 + (ENT ? 1 : 0) + (MAG ? 3 : 0);
  // carve-up: ring | scalar planes | mbarriers.  The dynamic shared window starts 1024-byte aligned
  // (no static shared memory in this kernel), which satisfies TMA's 128-byte destination alignment;
  // checked once below.
  double *ring = v2_smem;
  double *scal = ring + L::RING_DOUBLES;
  uint64_t *bar_ring = reinterpret_cast<uint64_t *>(scal + L::SCAL_DOUBLES);
  uint64_t *bar_scal = bar_ring + V2_NSLOT;

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * V2_TX + tx;
  const int bx0 = blockIdx.x * V2_TX, by0 = blockIdx.y * V2_TY; // origin of the halo'd tile (incl. ghosts)
  const int zc0 = a.z_lo + blockIdx.z * a.lz;
  const int zc1 = min(zc0 + a.lz, a.z_hi);
  if (zc0 >= zc1) return;
  const int ix = bx0 + tx + NG, iy = by0 + ty + NG;
  const bool active = ix < cP.nx + NG && iy < cP.ny + NG;
  const int ixc = min(ix, cP.nx + NG - 1), iyc = min(iy, cP.ny + NG - 1); // clamped (register pipelines)
  const int last_plane = zc1 - 1 + NG;

  if (tid == 0) {
    if (smem_u32(ring) & 127u) __trap();
    for (int s = 0; s < V2_NSLOT; ++s) mbar_init(&bar_ring[s], 1);
    mbar_init(&bar_scal[0], 1);
    mbar_init(&bar_scal[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  auto issue_ring = [&](int j) { // ring plane j <-> z = zc0-3+j
    const int p = zc0 - NG + j;
    if (p > last_plane) return;
    const int slot = j % V2_NSLOT;
    mbar_expect_tx(&bar_ring[slot], L::NRING * V2_BOX_BYTES);
#pragma unroll
    for (int f = 0; f < L::NRING; ++f) {
      const int var = f < 3 ? f : S::IAX + (f - 3);
      tma_load_4d(ring + (slot * L::NRING + f) * V2_PLANE_PAD, &tmap, &bar_ring[slot], bx0, by0, p, var);
    }
  };
  auto issue_scal = [&](int j) { // scalar plane j <-> z = zc0+j
    const int p = zc0 + j;
    if (p >= zc1) return;
    const int slot = j & 1;
    mbar_expect_tx(&bar_scal[slot], L::NSCAL * V2_BOX_BYTES);
    tma_load_4d(scal + (slot * L::NSCAL + 0) * V2_PLANE_PAD, &tmap, &bar_scal[slot], bx0, by0, p, S::ILN);
    if (ENT) tma_load_4d(scal + (slot * L::NSCAL + 1) * V2_PLANE_PAD, &tmap, &bar_scal[slot], bx0, by0, p, S::ISS);
  };

  if (tid == 0) {
#pragma unroll 1
    for (int j = 0; j < 7; ++j) issue_ring(j);
    issue_scal(0);
  }

  SmemLines<ENT, MAG, TRAIL> tp;
  if (TRAIL) tp.trail_zero();
  const int toff = (ty + NG) * V2_HX + (tx + NG);
  // register pipelines for the scalars (own column, straight from global; coalesced)
  const long long col = ixc + cP.sy * iyc;
  {
    const double *gl = a.fin + col + S::ILN * cP.sv;
#pragma unroll
    for (int d = 0; d < 7; ++d) tp.lz[d] = __ldg(gl + cP.sz * (zc0 - NG + d));
    if (ENT) {
      const double *gs = a.fin + col + S::ISS * cP.sv;
#pragma unroll
      for (int d = 0; d < 7; ++d) tp.sz[d] = __ldg(gs + cP.sz * (zc0 - NG + d));
    }
  }
  const double dtb = (*a.dt_dev) * a.beta;
  const bool rdf = a.read_df != 0 && active;
  // helical forcing (substep 3 only): fx(x)*fy(y) is fixed for this thread, fz(z) is fetched at the top of each plane
  double fxy_r = 0.0, fxy_i = 0.0;
  if (a.frc.on) {
    const double2 fx = a.frc.fx[ixc], fy = a.frc.fy[iyc];
    fxy_r = fx.x * fy.x - fx.y * fy.y;
    fxy_i = fx.x * fy.y + fx.y * fy.x;
  }

#pragma unroll 1
  for (int z = zc0; z < zc1; ++z) {
    const int j0 = z - zc0;
    // slot release + TMA issue: only the first warp waits until every warp has finished plane z-1
    // (named barriers 1/2 alternate, because a warp may run one plane ahead of the slowest one)
    if (tid < 32) {
      if (j0 > 0) named_bar_sync(1 + ((j0 - 1) & 1), V2_THREADS);
      if (tid == 0) {
        issue_ring(j0 + 7);
        issue_scal(j0 + 1);
        // (UTMAPF.L2 = cp.async.bulk.prefetch.tensor raised "illegal instruction" on sm_100a/driver 580 with this
        //  FP64 4-D map; df is prefetched with per-thread prefetch.global.L2 below instead)
        if (a.read_df && a.df_prefetch == 2 && z + 2 < zc1) tma_prefetch_4d(&tmap_df, bx0 + NG, by0 + NG, z + 2, 0); // debug only
      }
      __syncwarp();
    }
    const long long o = ix + cP.sy * iy + cP.sz * z;
    // early global loads: this cell's df (consumed at the very end of the iteration) and the scalars' next
    // pipeline entries (plane z+4).  Measured: loading df late instead (even after an L2 prefetch two planes
    // ahead) costs 20-40 %, the 16 registers are well spent.
    double dfv[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) dfv[k] = rdf ? __ldcs(a.df + o + k * cP.sv) : 0.0;
    if (rdf && a.df_prefetch == 1 && z + 2 < zc1) {
#pragma unroll
      for (int k = 0; k < NV; ++k) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.df + o + 2 * cP.sz + k * cP.sv));
    }
    double2 fzv = make_double2(0.0, 0.0);
    if (a.frc.on) fzv = __ldg(a.frc.fz + z);
    double nl = 0.0, ns = 0.0;
    if (z + 4 < cP.mz) {
      nl = __ldg(a.fin + col + S::ILN * cP.sv + cP.sz * (z + 4));
      if (ENT) ns = __ldg(a.fin + col + S::ISS * cP.sv + cP.sz * (z + 4));
    }
    if (j0 == 0) {
#pragma unroll 1
      for (int d = 0; d < 6; ++d) mbar_wait(&bar_ring[d], 0);
    }
    mbar_wait(&bar_ring[(j0 + 6) % V2_NSLOT], ((j0 + 6) / V2_NSLOT) & 1);
    mbar_wait(&bar_scal[j0 & 1], (j0 >> 1) & 1);
#pragma unroll
    for (int d = 0; d < 7; ++d) tp.pl[d] = ring + ((j0 + d) % V2_NSLOT) * (L::NRING * V2_PLANE_PAD) + toff;
    tp.scal = scal + (j0 & 1) * (L::NSCAL * V2_PLANE_PAD) + toff;
    if (TRAIL && j0 == 0) {
      tp.trail_prologue(0);
      if (MAG) tp.trail_prologue(1);
    }

    if (active) {
      double rhs[NV], fc[NV];
      rhs_cell_lines<ENT, MAG, HYP, UPW>(tp, rhs, fc);
      double force[3] = {0., 0., 0.};
      if (a.frc.on) { // forcing.f90:2170-2345
        const double re = fxy_r * fzv.x - fxy_i * fzv.y, im = fxy_r * fzv.y + fxy_i * fzv.x;
#pragma unroll
        for (int j = 0; j < 3; ++j) force[j] = a.frc.fda[j] * (a.frc.coef1[j] * re - a.frc.coef2[j] * im);
      }
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const long long q = o + k * cP.sv;
        const double d = a.alpha * dfv[k] + rhs[k]; // timestep.f90:113 (alpha=0 in substep 1)
        if (a.write_df) __stcs(a.df + q, d);
        double fn = fc[k] + dtb * d; // timestep.f90:160-161
        if (k < 3) fn += force[k];
        a.fout[q] = fn;
      }
    }
    // advance the scalar pipelines
#pragma unroll
    for (int d = 0; d < 6; ++d) {
      tp.lz[d] = tp.lz[d + 1];
      if (ENT) tp.sz[d] = tp.sz[d + 1];
    }
    tp.lz[6] = nl;
    if (ENT) tp.sz[6] = ns;
    // this warp is done with plane z-3's ring slot and with this scalar slot
    if (z + 1 < zc1 && tid >= 32) named_bar_arrive(1 + (j0 & 1), V2_THREADS);
  }
}

} // namespace pcb200
